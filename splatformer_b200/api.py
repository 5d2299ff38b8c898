"""The gsplat v0.1.11 Python API that SplatFormer calls, on top of libsplat_b200.so.

Same names, argument order, return tensors, error behaviour and autograd contract as the three
functions used at /root/reference/utils/gs_utils.py:78 (spherical_harmonics), :82-95
(project_gaussians, 12 positional arguments, 7-tuple return) and :96-109 (rasterize_gaussians with
background= and return_alpha=True).  Each autograd Function is a thin host wrapper: it allocates
outputs with torch, passes raw device pointers and the current CUDA stream through the C ABI
(include/splat_b200.h) and never touches the oracle or any CPU path.
"""
from __future__ import annotations

from typing import Optional

import torch
from torch import Tensor
from torch.autograd import Function

from . import _abi, utils
from .sh import deg_from_sh, num_sh_bases

F32, I32 = torch.float32, torch.int32


def _f(t: Tensor) -> Tensor:
    return _abi.aligned16(t.contiguous())


# --------------------------------------------------------------------------------------------
# spherical_harmonics
# --------------------------------------------------------------------------------------------
class _SphericalHarmonics(Function):
    @staticmethod
    def forward(ctx, degrees_to_use: int, viewdirs: Tensor, coeffs: Tensor):
        lib = _abi.load()
        n = coeffs.shape[0]
        degree = deg_from_sh(coeffs.shape[-2])
        viewdirs, coeffs = _f(viewdirs), _f(coeffs)
        colors = torch.empty(n, 3, dtype=F32, device=coeffs.device)
        _abi.check(lib.sb_sh_fwd(1, n, degree, degrees_to_use, _abi.ptr(viewdirs, F32, "viewdirs"),
                                 _abi.ptr(coeffs, F32, "coeffs"), _abi.ptr(colors, F32, "colors"),
                                 _abi.stream()), "sb_sh_fwd")
        ctx.save_for_backward(viewdirs)
        ctx.meta = (n, degree, degrees_to_use)
        return colors

    @staticmethod
    def backward(ctx, v_colors: Tensor):
        lib = _abi.load()
        (viewdirs,) = ctx.saved_tensors
        n, degree, degrees_to_use = ctx.meta
        v_colors = _f(v_colors)
        v_coeffs = torch.empty(n, num_sh_bases(degree), 3, dtype=F32, device=v_colors.device)
        _abi.check(lib.sb_sh_bwd(1, n, degree, degrees_to_use, _abi.ptr(viewdirs, F32, "viewdirs"),
                                 _abi.ptr(v_colors, F32, "v_colors"), _abi.ptr(v_coeffs, F32, "v_coeffs"),
                                 _abi.stream()), "sb_sh_bwd")
        return None, None, v_coeffs


def spherical_harmonics(degrees_to_use: int, viewdirs: Tensor, coeffs: Tensor,
                        method: str = "poly") -> Tensor:
    """Evaluate SH colours: (N,3) dirs x (N,K,3) coeffs -> (N,3).  No +0.5, no clamp.
    `method` ("poly" | "fast") is accepted for signature parity; both give the same values."""
    assert method in ("poly", "fast"), "Invalid method."
    assert coeffs.shape[-2] >= num_sh_bases(degrees_to_use)
    assert viewdirs.shape[-1] == 3 and coeffs.shape[-1] == 3
    return _SphericalHarmonics.apply(degrees_to_use, viewdirs, coeffs)


# --------------------------------------------------------------------------------------------
# project_gaussians
# --------------------------------------------------------------------------------------------
class _ProjectGaussians(Function):
    @staticmethod
    def forward(ctx, means3d, scales, glob_scale, quats, viewmat, fx, fy, cx, cy, img_height,
                img_width, block_width, clip_thresh):
        lib = _abi.load()
        n = means3d.shape[-2]
        dev = means3d.device
        means3d, scales, quats, viewmat = _f(means3d), _f(scales), _f(quats), _f(viewmat)
        if viewmat.numel() < 12:
            raise ValueError("viewmat must be (3,4) or (4,4)")
        xys = torch.empty(n, 2, dtype=F32, device=dev)
        depths = torch.empty(n, dtype=F32, device=dev)
        radii = torch.empty(n, dtype=I32, device=dev)
        conics = torch.empty(n, 3, dtype=F32, device=dev)
        compensation = torch.empty(n, dtype=F32, device=dev)
        num_tiles_hit = torch.empty(n, dtype=I32, device=dev)
        cov3d = torch.empty(n, 6, dtype=F32, device=dev)
        _abi.check(lib.sb_project_fwd(
            1, n, _abi.ptr(means3d, F32, "means3d"), _abi.ptr(scales, F32, "scales"), float(glob_scale),
            _abi.ptr(quats, F32, "quats"), _abi.ptr(viewmat, F32, "viewmat"), 12, None,
            float(fx), float(fy), float(cx), float(cy), int(img_height), int(img_width), int(block_width),
            float(clip_thresh), _abi.ptr(xys, F32, "xys"), _abi.ptr(depths, F32, "depths"),
            _abi.ptr(radii, I32, "radii"), _abi.ptr(conics, F32, "conics"),
            _abi.ptr(compensation, F32, "compensation"), _abi.ptr(num_tiles_hit, I32, "num_tiles_hit"),
            _abi.ptr(cov3d, F32, "cov3d"), _abi.stream()), "sb_project_fwd")
        ctx.scalars = (n, float(glob_scale), float(fx), float(fy), float(cx), float(cy), int(img_height),
                       int(img_width), int(block_width), float(clip_thresh))
        ctx.save_for_backward(means3d, scales, quats, viewmat, radii)
        ctx.mark_non_differentiable(radii, num_tiles_hit)
        ctx.set_materialize_grads(False)
        return xys, depths, radii, conics, compensation, num_tiles_hit, cov3d

    @staticmethod
    def backward(ctx, v_xys, v_depths, v_radii, v_conics, v_compensation, v_num_tiles_hit, v_cov3d):
        lib = _abi.load()
        means3d, scales, quats, viewmat, radii = ctx.saved_tensors
        n, glob_scale, fx, fy, cx, cy, img_height, img_width, block_width, clip_thresh = ctx.scalars
        dev = means3d.device
        v_xys = torch.zeros(n, 2, dtype=F32, device=dev) if v_xys is None else _f(v_xys)
        v_conics = torch.zeros(n, 3, dtype=F32, device=dev) if v_conics is None else _f(v_conics)
        v_depths = None if v_depths is None else _f(v_depths)
        v_compensation = None if v_compensation is None else _f(v_compensation)
        v_cov3d = None if v_cov3d is None else _f(v_cov3d)
        v_means = torch.empty(n, 3, dtype=F32, device=dev)
        v_scales = torch.empty(n, 3, dtype=F32, device=dev)
        v_quats = torch.empty(n, 4, dtype=F32, device=dev)
        v_viewmat12 = torch.zeros(12, dtype=F32, device=dev) if ctx.needs_input_grad[4] else None
        _abi.check(lib.sb_project_bwd(
            1, n, _abi.ptr(means3d, F32, "means3d"), _abi.ptr(scales, F32, "scales"), glob_scale,
            _abi.ptr(quats, F32, "quats"), _abi.ptr(viewmat, F32, "viewmat"), 12, None, fx, fy, cx, cy,
            img_height, img_width, block_width, clip_thresh, _abi.ptr(radii, I32, "radii"),
            _abi.ptr(v_xys, F32, "v_xys"), _abi.ptr(v_depths, F32, "v_depths"),
            _abi.ptr(v_conics, F32, "v_conics"), _abi.ptr(v_compensation, F32, "v_compensation"),
            _abi.ptr(v_cov3d, F32, "v_cov3d"), _abi.ptr(v_means, F32, "v_means"),
            _abi.ptr(v_scales, F32, "v_scales"), _abi.ptr(v_quats, F32, "v_quats"),
            _abi.ptr(v_viewmat12, F32, "v_viewmat"), _abi.stream()), "sb_project_bwd")
        v_viewmat = None
        if v_viewmat12 is not None:
            v_viewmat = torch.zeros_like(viewmat)
            v_viewmat.view(-1)[:12] = v_viewmat12
        return (v_means, v_scales, None, v_quats, v_viewmat, None, None, None, None, None, None, None, None)


def project_gaussians(means3d: Tensor, scales: Tensor, glob_scale: float, quats: Tensor,
                      viewmat: Tensor, fx: float, fy: float, cx: float, cy: float, img_height: int,
                      img_width: int, block_width: int, clip_thresh: float = 0.01):
    """Project 3D Gaussians to 2D.  Returns (xys (N,2), depths (N,), radii (N,) int32, conics
    (N,3), compensation (N,), num_tiles_hit (N,) int32, cov3d (N,6)); culled Gaussians have
    radii = num_tiles_hit = 0 and zeros elsewhere."""
    assert block_width > 1 and block_width <= 16, "block_width must be between 2 and 16"
    assert (quats.norm(dim=-1) - 1 < 1e-6).all(), "quats must be normalized"
    return _ProjectGaussians.apply(means3d, scales, glob_scale, quats, viewmat, fx, fy, cx, cy,
                                   img_height, img_width, block_width, clip_thresh)


# --------------------------------------------------------------------------------------------
# rasterize_gaussians
# --------------------------------------------------------------------------------------------
def _pack(xys: Tensor, conics: Tensor, colors: Tensor, opacity: Tensor) -> Tensor:
    lib = _abi.load()
    n = xys.shape[0]
    packed = torch.empty(n, 12, dtype=F32, device=xys.device)
    _abi.check(lib.sb_pack_splats(n, _abi.ptr(xys, F32, "xys"), _abi.ptr(conics, F32, "conics"),
                                  _abi.ptr(colors, F32, "colors"), _abi.ptr(opacity, F32, "opacity"),
                                  _abi.ptr(packed, F32, "packed"), _abi.stream()), "sb_pack_splats")
    return packed


# Segment capacity / rect-list pool of the per-view rect path (block width 16).  A frame whose tiles are deeper,
# or whose lists need more room, is detected through the totals the tile sort reports and takes the global path.
RECT_TILE_CAPACITY = 2048
USE_RECT_PATH = True


class _RasterizeGaussians(Function):
    @staticmethod
    def forward(ctx, xys, depths, radii, conics, num_tiles_hit, colors, opacity, img_height,
                img_width, block_width, background, return_alpha):
        lib = _abi.load()
        n = xys.shape[0]
        dev = xys.device
        channels = colors.shape[-1]
        tiles_x = (img_width + block_width - 1) // block_width
        tiles_y = (img_height + block_width - 1) // block_width
        groups = _channel_groups(channels)
        p = _abi.ptr

        def empty_frame():
            out_img = torch.ones(img_height, img_width, channels, dtype=F32, device=dev) * background
            final_ts = torch.zeros(img_height, img_width, dtype=F32, device=dev)
            ctx.empty = True
            ctx.num_intersects = 0
            ctx.set_materialize_grads(False)
            out_alpha = 1 - final_ts
            return (out_img, out_alpha) if return_alpha else out_img

        xys_c, conics_c, colors_c = _f(xys), _f(conics), _f(colors)
        opacity_c = _f(opacity).view(-1)
        background_c = _f(background.to(F32))
        rect_ids = rect_meta = gaussian_ids_sorted = tile_bins = None
        packs = None
        num_intersects = -1
        num_tiles = tiles_x * tiles_y
        cap = RECT_TILE_CAPACITY
        if USE_RECT_PATH and block_width == 16 and n > 0 and num_tiles * cap < 2 ** 31:
            # Per-tile segments + per-tile depth sort + rect lists (csrc/tile_sort.cu), as the fused path: no prefix
            # sum, no emission, no 64-bit global sort, no bin-edge pass; the one host read (intersection count and
            # overflow flags) takes the place of upstream's cum_tiles_hit[-1].item().
            st = _abi.stream()
            packs = [_pack(xys_c, conics_c, _group3(colors_c, lo, width), opacity_c) for lo, width in groups]
            state = torch.zeros(num_tiles + 1 + 4, dtype=I32, device=dev)
            tile_counts, view_has, totals, work = state[:num_tiles], state[num_tiles:-4], state[-4:-1], state[-1:]
            tile_keys = torch.empty(num_tiles * cap, dtype=torch.int64, device=dev)
            _abi.check(lib.sb_bin_to_tile_segments(
                1, n, p(xys_c, F32, "xys"), 2, p(_f(depths), F32, "depths"), p(_f(radii), I32, "radii"), tiles_x,
                tiles_y, block_width, p(tile_counts, I32, "tile_counts"), p(tile_keys, torch.int64, "tile_keys"), cap,
                st), "sb_bin_to_tile_segments")
            pool_cap = int(min(num_tiles * 8 * cap, max(12 * n, 1 << 18)))
            rect_ids = torch.empty(pool_cap + 32, dtype=I32, device=dev)
            rect_meta = torch.empty(num_tiles * 8, 2, dtype=I32, device=dev)
            tile_bins = torch.empty(num_tiles, 2, dtype=I32, device=dev)
            _abi.check(lib.sb_sort_tiles(
                1, num_tiles, cap, p(tile_counts, I32, "tile_counts"), p(tile_keys, torch.int64, "tile_keys"), None,
                p(tile_bins, I32, "tile_bins"), p(view_has, I32, "view_has_isects"), p(totals, I32, "totals"), n,
                tiles_x, p(packs[0], F32, "packed"), p(rect_ids, I32, "rect_ids"), p(rect_meta, I32, "rect_meta"),
                p(totals[2:], I32, "pool_head"), pool_cap, st), "sb_sort_tiles")
            num_intersects, overflow, _ = (int(x) for x in totals.tolist())          # the one host read
            if overflow:
                rect_ids = rect_meta = None                 # deep tiles / many tiles per Gaussian: global path below
            elif num_intersects < 1:
                return empty_frame()
        final_ts = torch.empty(img_height, img_width, dtype=F32, device=dev)
        final_idx = torch.empty(img_height, img_width, dtype=I32, device=dev)
        out_alpha = torch.empty(img_height, img_width, dtype=F32, device=dev) if return_alpha else None
        if rect_ids is None:
            num_intersects, cum_tiles_hit = utils.compute_cumulative_intersects(num_tiles_hit)
            if num_intersects < 1:
                return empty_frame()
            (_, _, _, gaussian_ids_sorted, tile_bins) = utils.bin_and_sort_gaussians(
                n, num_intersects, xys_c, _f(depths), _f(radii), cum_tiles_hit, (tiles_x, tiles_y, 1),
                block_width)
        # The kernels composite three channels per pass; other channel counts (upstream's nd_rasterize,
        # SURVEY.md 8f-4) run ceil(C/3) passes over the same sorted lists, zero-padding the last group.
        parts = []
        if packs is None:
            packs = [_pack(xys_c, conics_c, _group3(colors_c, lo, width), opacity_c) for lo, width in groups]
        for gi, (lo, width) in enumerate(groups):
            bg_g = _group3(background_c, lo, width)
            img_g = torch.empty(img_height, img_width, 3, dtype=F32, device=dev)
            if rect_ids is not None:
                if gi > 0:
                    work.zero_()              # the persistent kernel's queue head, consumed by the previous pass
                _abi.check(lib.sb_rasterize_rect_fwd(
                    1, n, p(packs[gi], F32, "packed"), p(rect_ids, I32, "rect_ids"), p(rect_meta, I32, "rect_meta"),
                    p(bg_g, F32, "background"), img_height, img_width, None, p(img_g, F32, "out_img"),
                    p(final_ts, F32, "final_Ts"), p(final_idx, I32, "final_idx"), p(out_alpha, F32, "out_alpha"), 0,
                    None, 0, None, p(work, I32, "work_counter"), _abi.stream()), "sb_rasterize_rect_fwd")
            else:
                _abi.check(lib.sb_rasterize_fwd(
                    1, n, p(packs[gi], F32, "packed"), p(gaussian_ids_sorted, I32, "gaussian_ids_sorted"),
                    p(tile_bins, I32, "tile_bins"), p(bg_g, F32, "background"), img_height,
                    img_width, block_width, None, p(img_g, F32, "out_img"), p(final_ts, F32, "final_Ts"),
                    p(final_idx, I32, "final_idx"), p(out_alpha, F32, "out_alpha"), 0, None, 0, None,
                    _abi.stream()), "sb_rasterize_fwd")
            parts.append(img_g if width == 3 else img_g[..., :width])
        out_img = parts[0] if len(parts) == 1 else torch.cat(parts, dim=-1)
        ctx.empty = False
        ctx.num_intersects = num_intersects
        ctx.rect = rect_ids is not None
        ctx.meta = (n, img_height, img_width, block_width, channels)
        lists = (rect_ids, rect_meta) if ctx.rect else (gaussian_ids_sorted, tile_bins)
        ctx.save_for_backward(lists[0], lists[1], background_c, final_ts, final_idx, conics_c, opacity_c, *packs)
        ctx.set_materialize_grads(False)
        return (out_img, out_alpha) if return_alpha else out_img

    @staticmethod
    def backward(ctx, v_out_img, v_out_alpha=None):
        none12 = (None,) * 12
        if ctx.empty or (v_out_img is None and v_out_alpha is None):
            return none12
        lib = _abi.load()
        p = _abi.ptr
        (list_a, list_b, background, final_ts, final_idx, conics, opacity, *packs) = ctx.saved_tensors
        n, img_height, img_width, block_width, channels = ctx.meta
        dev = conics.device
        v_out_img = torch.zeros(img_height, img_width, channels, dtype=F32, device=dev) if v_out_img is None \
            else _f(v_out_img)
        v_out_alpha = None if v_out_alpha is None else _f(v_out_alpha)
        v_xys = v_conics = v_opacity = None
        v_color_parts = []
        for gi, (lo, width) in enumerate(_channel_groups(channels)):
            v_img_g, bg_g = _group3(v_out_img, lo, width), _group3(background, lo, width)
            v_packed = torch.zeros(n, 12, dtype=F32, device=dev)
            # d/d alpha enters once: with the first group
            if ctx.rect:
                _abi.check(lib.sb_rasterize_rect_bwd(
                    1, n, p(packs[gi], F32, "packed"), p(list_a, I32, "rect_ids"), p(list_b, I32, "rect_meta"),
                    p(bg_g, F32, "background"), img_height, img_width, p(final_ts, F32, "final_Ts"),
                    p(final_idx, I32, "final_idx"), p(v_img_g, F32, "v_out_img"),
                    p(v_out_alpha if gi == 0 else None, F32, "v_out_alpha"), None, 0, None, 0, None,
                    p(v_packed, F32, "v_packed"), _abi.stream()), "sb_rasterize_rect_bwd")
            else:
                _abi.check(lib.sb_rasterize_bwd(
                    1, n, p(packs[gi], F32, "packed"), p(list_a, I32, "gaussian_ids_sorted"),
                    p(list_b, I32, "tile_bins"), p(bg_g, F32, "background"), img_height,
                    img_width, block_width, p(final_ts, F32, "final_Ts"), p(final_idx, I32, "final_idx"),
                    p(v_img_g, F32, "v_out_img"), p(v_out_alpha if gi == 0 else None, F32, "v_out_alpha"),
                    None, 0, None, 0, None, p(v_packed, F32, "v_packed"), _abi.stream()), "sb_rasterize_bwd")
            g_xys = torch.empty(n, 2, dtype=F32, device=dev)
            g_conics = torch.empty(n, 3, dtype=F32, device=dev)
            g_colors = torch.empty(n, 3, dtype=F32, device=dev)
            g_opacity = torch.empty(n, 1, dtype=F32, device=dev)
            _abi.check(lib.sb_unpack_splat_grads(
                n, p(v_packed, F32, "v_packed"), p(conics, F32, "conics"),
                p(opacity, F32, "opacity"), p(g_xys, F32, "v_xys"),
                p(g_conics, F32, "v_conics"), p(g_colors, F32, "v_colors"),
                p(g_opacity, F32, "v_opacity"), _abi.stream()), "sb_unpack_splat_grads")
            v_xys = g_xys if v_xys is None else v_xys + g_xys
            v_conics = g_conics if v_conics is None else v_conics + g_conics
            v_opacity = g_opacity if v_opacity is None else v_opacity + g_opacity
            v_color_parts.append(g_colors if width == 3 else g_colors[:, :width])
        v_colors = v_color_parts[0] if len(v_color_parts) == 1 else torch.cat(v_color_parts, dim=-1)
        return (v_xys, None, None, v_conics, None, v_colors, v_opacity, None, None, None, None, None)


def _channel_groups(channels: int):
    """[(first channel, width <= 3)] covering `channels`."""
    return [(lo, min(3, channels - lo)) for lo in range(0, channels, 3)]


def _group3(t: Tensor, lo: int, width: int) -> Tensor:
    """Channels [lo, lo+width) of the last dimension as a contiguous 3-channel tensor (zero-padded)."""
    if lo == 0 and width == 3 and t.shape[-1] == 3:
        return t
    part = t[..., lo:lo + width]
    if width < 3:
        part = torch.nn.functional.pad(part, (0, 3 - width))
    return _f(part)


def rasterize_gaussians(xys: Tensor, depths: Tensor, radii: Tensor, conics: Tensor,
                        num_tiles_hit: Tensor, colors: Tensor, opacity: Tensor, img_height: int,
                        img_width: int, block_width: int, background: Optional[Tensor] = None,
                        return_alpha: Optional[bool] = False):
    """Rasterise 2D Gaussians: sort by (tile, depth), composite front to back per 16x16 tile.
    Returns out_img (H,W,C) [and out_alpha (H,W) = 1 - final transmittance if return_alpha].
    C = 3 is what SplatFormer uses (/root/reference/utils/gs_utils.py:65,78); any other channel count
    (upstream's nd_rasterize path) is composited three channels per pass over the same sorted lists."""
    assert block_width > 1 and block_width <= 16, "block_width must be between 2 and 16"
    if colors.dtype == torch.uint8:
        colors = colors.float() / 255
    if background is not None:
        assert background.shape[0] == colors.shape[-1], \
            f"incorrect shape of background color tensor, expected shape {colors.shape[-1]}"
    else:
        background = torch.ones(colors.shape[-1], dtype=torch.float32, device=colors.device)
    if xys.ndimension() != 2 or xys.size(1) != 2:
        raise ValueError("xys must have dimensions (N, 2)")
    if colors.ndimension() != 2:
        raise ValueError("colors must have dimensions (N, D)")
    if colors.shape[-1] < 1:
        raise ValueError("colors must have at least one channel")
    res = _RasterizeGaussians.apply(xys, depths, radii, conics, num_tiles_hit, colors,
                                    opacity.reshape(-1, 1), img_height, img_width, block_width,
                                    background, bool(return_alpha))
    return res
