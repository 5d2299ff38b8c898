"""Fused multi-view render: all V views of a scene in one pass per stage.

This is the B200-native form of the loop SplatFormer runs in
`rasterize_gaussians_to_multiimgs` (/root/reference/utils/gs_utils.py:20-27): instead of V times
(SH, project, cumsum + .item(), emit, sort, bins, rasterise) -- about ten launches and nine host
synchronisations per view -- every stage is launched once for the whole (V, N) / (V, H, W) batch
through the C ABI, view-independent data (SH coefficients, Gaussian parameters) is read from HBM
once per scene, and the per-Gaussian gradients are summed over the views in registers.
Results are the ones the per-view gsplat-compatible API (`splatformer_b200.api`) produces: the
same device functions run underneath.
"""
from __future__ import annotations

import contextlib
import os
from typing import Dict, List, Optional, Tuple

import torch
from torch import Tensor
from torch.autograd import Function

from . import _abi, utils

F32, I32 = torch.float32, torch.int32


class StageTimer:
    """CUDA-event timing of the individual ABI calls (used by bench.py for the roofline of the
    dominant kernel).  Events are recorded on the current stream around each stage."""

    def __init__(self):
        self.events: List[Tuple[str, torch.cuda.Event, torch.cuda.Event]] = []

    @contextlib.contextmanager
    def stage(self, name: str):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        yield
        b.record()
        self.events.append((name, a, b))

    def totals_ms(self) -> Dict[str, float]:
        torch.cuda.synchronize()
        out: Dict[str, float] = {}
        for name, a, b in self.events:
            out[name] = out.get(name, 0.0) + a.elapsed_time(b)
        return out

    def counts(self) -> Dict[str, int]:
        out: Dict[str, int] = {}
        for name, _, _ in self.events:
            out[name] = out.get(name, 0) + 1
        return out


_timer: Optional[StageTimer] = None
_last_num_intersects = 0


def last_num_intersects() -> int:
    """Intersections (over all views) of the most recent render_views call (bench.py reports it)."""
    return _last_num_intersects


@contextlib.contextmanager
def profile_stages():
    """with profile_stages() as t: ...render...; t.totals_ms()"""
    global _timer
    prev, _timer = _timer, StageTimer()
    try:
        yield _timer
    finally:
        _timer = prev


_NVTX = bool(os.environ.get("SB_NVTX"))      # SB_NVTX=1: an NVTX range per stage (nsys / ncu --nvtx timelines)


def _stage(name: str):
    if _timer is not None:
        return _timer.stage(name)
    if _NVTX:
        return torch.cuda.nvtx.range("sb." + name)
    return contextlib.nullcontext()


def _c(t: Tensor) -> Tensor:
    return _abi.aligned16(t.contiguous())


_side_streams: Dict[int, torch.cuda.Stream] = {}
_pinned: Dict[int, Tensor] = {}
_overflowed: Dict[Tuple[int, int, int], int] = {}
USE_RECT_LISTS = True               # fused path, block width 16: rect-list rasteriser (csrc/raster_rect.cu)
DEFAULT_TILE_CAPACITY = 2048        # doubled per overflow of a (views, tiles, Gaussians) shape, up to the library's limit
# Scenes whose Gaussians cover many tiles each (large screen radii: a room seen from inside, the 4K stress
# shape) are binned faster by the depth-ordered global sort than by per-tile slot claims -- hundreds of
# millions of contended atomics and scattered 8-byte stores.  A shape that averaged more tiles per
# (view, Gaussian) than this goes to the global sort from its next call on.
GLOBAL_SORT_TILES_PER_GAUSSIAN = 16.0
_prefer_global: Dict[Tuple[int, int, int], bool] = {}


def _pinned_pair(device) -> Tensor:
    idx = device.index if device.index is not None else torch.cuda.current_device()
    if idx not in _pinned:
        _pinned[idx] = torch.zeros(2, dtype=I32).pin_memory()
    return _pinned[idx]


def _tile_capacity(lib, v: int, num_tiles: int, n: int) -> int:
    """Segment capacity for the tile sort, or 0 to use the global sort.  A (views, tiles, Gaussians)
    shape whose tiles overflowed the capacity gets twice the capacity next time; past the library's
    limit it goes straight to the global sort."""
    if _prefer_global.get((v, num_tiles, n)):
        return 0
    cap = DEFAULT_TILE_CAPACITY << _overflowed.get((v, num_tiles, n), 0)
    if cap > lib.sb_tile_sort_capacity() or v * num_tiles * cap >= 2 ** 31:
        return 0
    return cap


def _note_overflow(v: int, num_tiles: int, n: int) -> None:
    _overflowed[(v, num_tiles, n)] = _overflowed.get((v, num_tiles, n), 0) + 1


def _side_stream(device) -> torch.cuda.Stream:
    idx = device.index if device.index is not None else torch.cuda.current_device()
    if idx not in _side_streams:
        _side_streams[idx] = torch.cuda.Stream(device=device)
    return _side_streams[idx]


class _RenderViews(Function):
    """One autograd node for the whole scene.

    `flags` (see _abi.FLAG_*) tell the kernels which of the wrapper's activations they apply
    themselves; `color_rest` is features_rest when the SH coefficients arrive split as SplatFormer
    stores them.  With `gt` given the node returns the per-view L1 and squared-error sums instead of
    leaving the loss to torch (train.py:275,278) and the backward derives the image gradient inside
    the raster kernel."""

    @staticmethod
    def forward(ctx, means, scales, quats, opacities, color_src, color_rest, viewmats, campos, intrins, gt,
                sh_degree, flags, img_height, img_width, block_width, background, clip_thresh, clamp_rgb):
        lib = _abi.load()
        st = _abi.stream()                      # one current-stream lookup per pass (it costs ~20 us)
        dev = means.device
        n = means.shape[0]
        v = viewmats.shape[0]
        means, scales, quats = _c(means), _c(scales), _c(quats)
        opacities = _c(opacities).view(-1)
        color_src = _c(color_src)
        color_rest = None if color_rest is None else _c(color_rest)
        viewmats = _c(viewmats.to(F32))
        vm_stride = viewmats.shape[1] * viewmats.shape[2]
        campos = None if campos is None else _c(campos.to(F32))
        intrins = _c(intrins.to(F32))
        background = _c(background.to(F32))
        gt_ready = None
        if gt is not None:
            if gt.dtype not in (torch.uint8, F32):
                gt = gt.to(F32)
            if not gt.is_cuda:
                # targets in (pinned) host memory: the upload runs on a side stream underneath the
                # projection / sort stages and is awaited right before the raster kernel needs it
                side = _side_stream(dev)
                with torch.cuda.stream(side):
                    gt = gt.contiguous().to(dev, non_blocking=True)
                    gt_ready = torch.cuda.Event()
                    gt_ready.record(side)
                gt.record_stream(torch.cuda.current_stream())
            gt = _c(gt)
        gt_dtype = None if gt is None else gt.dtype
        tiles_x = (img_width + block_width - 1) // block_width
        tiles_y = (img_height + block_width - 1) // block_width

        packed = torch.empty(v, n, 12, dtype=F32, device=dev)
        depths = torch.empty(v, n, dtype=F32, device=dev)
        radii = torch.empty(v, n, dtype=I32, device=dev)
        tiles_hit = torch.empty(v, n, dtype=I32, device=dev)
        num_tiles = tiles_x * tiles_y
        out_img = final_ts = final_idx = out_alpha = None

        def alloc_images():
            # after the projection kernel is on its way: the host reaches the first launch sooner
            nonlocal out_img, final_ts, final_idx, out_alpha
            out_img = torch.empty(v, img_height, img_width, 3, dtype=F32, device=dev)
            final_ts = torch.empty(v, img_height, img_width, dtype=F32, device=dev)
            final_idx = torch.empty(v, img_height, img_width, dtype=I32, device=dev)
            out_alpha = None if gt is not None else torch.empty(v, img_height, img_width, dtype=F32, device=dev)
        # per-tile loss sums (fallback raster kernel) or per-rectangle sums (rect-list kernel)
        loss_buf = [None]

        def loss_partials(per_tile):
            loss_buf[0] = None if gt is None else torch.empty(v, num_tiles * per_tile, 2, dtype=F32, device=dev)
            return loss_buf[0]

        def project(tile_counts, tile_keys, cap):
            with _stage("project_sh_fwd"):
                _abi.check(lib.sb_views_project_fwd(
                    v, n, _abi.ptr(means, F32, "means"), _abi.ptr(scales, F32, "scales"), 1.0,
                    _abi.ptr(quats, F32, "quats"), _abi.ptr(opacities, F32, "opacities"),
                    _abi.ptr(color_src, F32, "color_src"), _abi.ptr(color_rest, F32, "color_rest"), sh_degree,
                    flags, _abi.ptr(viewmats, F32, "viewmats"), vm_stride, _abi.ptr(intrins, F32, "intrins"), 0.0,
                    0.0, 0.0, 0.0, _abi.ptr(campos, F32, "campos"), img_height, img_width, block_width,
                    clip_thresh, _abi.ptr(packed, F32, "packed"), _abi.ptr(depths, F32, "depths"),
                    _abi.ptr(radii, I32, "radii"), _abi.ptr(tiles_hit, I32, "num_tiles_hit"),
                    _abi.ptr(tile_counts, I32, "tile_counts"), _abi.ptr(tile_keys, torch.int64, "tile_keys"), cap,
                    st), "sb_views_project_fwd")

        def raster_rect(rect_ids, rect_counts, cap, view_has, work):
            if gt_ready is not None:
                torch.cuda.current_stream().wait_event(gt_ready)
            loss_partial = loss_partials(8)
            with _stage("raster_fwd"):
                _abi.check(lib.sb_rasterize_rect_fwd(
                    v, n, _abi.ptr(packed, F32, "packed"), _abi.ptr(rect_ids, I32, "rect_ids"),
                    _abi.ptr(rect_counts, I32, "rect_counts"), cap, _abi.ptr(background, F32, "background"),
                    img_height, img_width, _abi.ptr(view_has, I32, "view_has_isects"),
                    _abi.ptr(out_img, F32, "out_img"), _abi.ptr(final_ts, F32, "final_Ts"),
                    _abi.ptr(final_idx, I32, "final_idx"), _abi.ptr(out_alpha, F32, "out_alpha"),
                    1 if clamp_rgb else 0, _abi.ptr(gt, gt_dtype, "gt"), 1 if gt_dtype == torch.uint8 else 0,
                    _abi.ptr(loss_partial, F32, "loss_partial"), _abi.ptr(work, I32, "work_counter"), st),
                    "sb_rasterize_rect_fwd")

        def raster(gids_sorted, tile_bins, view_has):
            if gt_ready is not None:
                torch.cuda.current_stream().wait_event(gt_ready)
            loss_partial = loss_partials(1)
            with _stage("raster_fwd"):
                _abi.check(lib.sb_rasterize_fwd(
                    v, n, _abi.ptr(packed, F32, "packed"), _abi.ptr(gids_sorted, I32, "gaussian_ids_sorted"),
                    _abi.ptr(tile_bins, I32, "tile_bins"), _abi.ptr(background, F32, "background"), img_height,
                    img_width, block_width, _abi.ptr(view_has, I32, "view_has_isects"),
                    _abi.ptr(out_img, F32, "out_img"), _abi.ptr(final_ts, F32, "final_Ts"),
                    _abi.ptr(final_idx, I32, "final_idx"), _abi.ptr(out_alpha, F32, "out_alpha"),
                    1 if clamp_rgb else 0, _abi.ptr(gt, gt_dtype, "gt"), 1 if gt_dtype == torch.uint8 else 0,
                    _abi.ptr(loss_partial, F32, "loss_partial"), st), "sb_rasterize_fwd")

        global _last_num_intersects
        cap = _tile_capacity(lib, v, num_tiles, n)
        done = False
        rect_ids = rect_counts = None
        if cap:
            # fixed-capacity tile segments + per-tile sort (csrc/tile_sort.cu).  The raster kernel is
            # enqueued optimistically; the host then reads the overflow flag, which the GPU produced
            # BEFORE the raster kernel started -- the device never idles on this round trip.
            with _stage("zero_counts"):
                state = torch.zeros(v * num_tiles + v + 4, dtype=I32, device=dev)
            tile_counts, view_has, totals = state[:v * num_tiles], state[v * num_tiles:-4], state[-4:-2]
            work = state[-2:]               # rectangle queue heads of the persistent raster kernels
            tile_keys = torch.empty(v * num_tiles * cap, dtype=torch.int64, device=dev)
            project(tile_counts, tile_keys, cap)
            alloc_images()
            gids_sorted = torch.empty(v * num_tiles * cap, dtype=I32, device=dev)
            tile_bins = torch.empty(v * num_tiles, 2, dtype=I32, device=dev)
            use_rect = block_width == 16 and USE_RECT_LISTS
            if use_rect:
                # per-rectangle lists for the rect-list rasteriser (one warp per 8x4 rectangle)
                rect_ids = torch.empty(v * num_tiles * 8 * cap, dtype=I32, device=dev)
                rect_counts = torch.empty(v * num_tiles * 8, dtype=I32, device=dev)
            with _stage("sort_tiles"):
                _abi.check(lib.sb_sort_tiles(
                    v, num_tiles, cap, _abi.ptr(tile_counts, I32, "tile_counts"),
                    _abi.ptr(tile_keys, torch.int64, "tile_keys"), _abi.ptr(gids_sorted, I32, "gaussian_ids_sorted"),
                    _abi.ptr(tile_bins, I32, "tile_bins"), _abi.ptr(view_has, I32, "view_has_isects"),
                    _abi.ptr(totals, I32, "totals"), n, tiles_x, _abi.ptr(packed, F32, "packed"),
                    _abi.ptr(rect_ids, I32, "rect_ids"), _abi.ptr(rect_counts, I32, "rect_counts"),
                    st), "sb_sort_tiles")
            host_totals = _pinned_pair(dev)
            host_totals.copy_(totals, non_blocking=True)
            flag_ready = torch.cuda.Event()
            flag_ready.record()
            if use_rect:
                raster_rect(rect_ids, rect_counts, cap, view_has, work[:1])
            else:
                raster(gids_sorted, tile_bins, view_has)
            flag_ready.synchronize()
            num_intersects, overflow = int(host_totals[0]), int(host_totals[1])
            done = overflow == 0
            if num_intersects > GLOBAL_SORT_TILES_PER_GAUSSIAN * v * n:
                _prefer_global[(v, num_tiles, n)] = True
            if not done:
                _note_overflow(v, num_tiles, n)
                rect_ids = rect_counts = None
        if not done:
            # very deep tiles / many tiles per Gaussian: global sort over all (tile, Gaussian) pairs
            if not cap:
                project(None, None, 0)
                alloc_images()
            # depth pre-order of the Gaussians, emission in that order, radix sort over the tile bits only
            with _stage("global_sort"):
                num_intersects, _, gids_sorted, tile_bins, view_ends = utils.bin_depth_ordered(
                    v, n, packed, 12, depths.view(-1), radii.view(-1), tiles_hit, (tiles_x, tiles_y, 1), block_width,
                    compact=True)
            view_has = (torch.diff(view_ends, prepend=view_ends.new_zeros(1)) > 0).to(I32)
            raster(gids_sorted, tile_bins, view_has)
        _last_num_intersects = num_intersects
        ctx.meta = (v, n, sh_degree, flags, img_height, img_width, block_width, clip_thresh, vm_stride,
                    tuple(color_src.shape), None if color_rest is None else tuple(color_rest.shape),
                    num_intersects, clamp_rgb, cap)
        keep_img = clamp_rgb or gt is not None
        ctx.save_for_backward(means, scales, quats, viewmats, campos, intrins, background, packed,
                              gids_sorted, tile_bins, final_ts, final_idx, out_img if keep_img else None, gt,
                              rect_ids, rect_counts)
        ctx.set_materialize_grads(False)
        if gt is not None:
            sums = loss_buf[0].sum(dim=1)                       # (V, 2): sum |rgb - gt|, sum (rgb - gt)^2
            l1_sum, sq_sum = sums[:, 0], sums[:, 1]
            ctx.mark_non_differentiable(sq_sum, out_img)
            return l1_sum, sq_sum, out_img
        return out_img, out_alpha

    @staticmethod
    def backward(ctx, *grads):
        lib = _abi.load()
        st = _abi.stream()
        (means, scales, quats, viewmats, campos, intrins, background, packed, gids_sorted, tile_bins,
         final_ts, final_idx, out_img, gt, rect_ids, rect_counts) = ctx.saved_tensors
        (v, n, sh_degree, flags, img_height, img_width, block_width, clip_thresh, vm_stride, color_shape,
         rest_shape, num_intersects, clamp_rgb, cap) = ctx.meta
        dev = means.device
        nin = 18
        if gt is not None:
            v_l1 = grads[0]
            if v_l1 is None:
                return (None,) * nin
            v_loss = _c(v_l1.to(F32))
            v_out_img = v_out_alpha = None
        else:
            v_out_img, v_out_alpha = grads
            if v_out_img is None and v_out_alpha is None:
                return (None,) * nin
            v_out_img = None if v_out_img is None else _c(v_out_img)
            v_out_alpha = None if v_out_alpha is None else _c(v_out_alpha)
            v_loss = None
        with _stage("zero_grads"):
            v_packed = torch.zeros(v, n, 12, dtype=F32, device=dev)
        gt_dt = None if gt is None else gt.dtype
        if rect_ids is not None:
            with _stage("raster_bwd"):
                _abi.check(lib.sb_rasterize_rect_bwd(
                    v, n, _abi.ptr(packed, F32, "packed"), _abi.ptr(rect_ids, I32, "rect_ids"),
                    _abi.ptr(rect_counts, I32, "rect_counts"), cap, _abi.ptr(background, F32, "background"),
                    img_height, img_width, _abi.ptr(final_ts, F32, "final_Ts"),
                    _abi.ptr(final_idx, I32, "final_idx"), _abi.ptr(v_out_img, F32, "v_out_img"),
                    _abi.ptr(v_out_alpha, F32, "v_out_alpha"), _abi.ptr(out_img, F32, "out_img"),
                    1 if clamp_rgb else 0, _abi.ptr(gt, gt_dt, "gt"), 1 if gt_dt == torch.uint8 else 0,
                    _abi.ptr(v_loss, F32, "v_loss"), _abi.ptr(v_packed, F32, "v_packed"), st),
                    "sb_rasterize_rect_bwd")
        else:
          with _stage("raster_bwd"):
            _abi.check(lib.sb_rasterize_bwd(
                v, n, _abi.ptr(packed, F32, "packed"), _abi.ptr(gids_sorted, I32, "gaussian_ids_sorted"),
                _abi.ptr(tile_bins, I32, "tile_bins"), _abi.ptr(background, F32, "background"), img_height,
                img_width, block_width, _abi.ptr(final_ts, F32, "final_Ts"),
                _abi.ptr(final_idx, I32, "final_idx"), _abi.ptr(v_out_img, F32, "v_out_img"),
                _abi.ptr(v_out_alpha, F32, "v_out_alpha"), _abi.ptr(out_img, F32, "out_img"),
                1 if clamp_rgb else 0, _abi.ptr(gt, None if gt is None else gt.dtype, "gt"),
                1 if (gt is not None and gt.dtype == torch.uint8) else 0, _abi.ptr(v_loss, F32, "v_loss"),
                _abi.ptr(v_packed, F32, "v_packed"), st), "sb_rasterize_bwd")
        v_means = torch.empty(n, 3, dtype=F32, device=dev)
        v_scales = torch.empty(n, 3, dtype=F32, device=dev)
        v_quats = torch.empty(n, 4, dtype=F32, device=dev)
        v_opac = torch.empty(n, 1, dtype=F32, device=dev)
        v_color = torch.empty(color_shape, dtype=F32, device=dev)
        v_rest = None if rest_shape is None else torch.empty(rest_shape, dtype=F32, device=dev)
        with _stage("project_sh_bwd"):
            _abi.check(lib.sb_views_project_bwd(
                v, n, _abi.ptr(means, F32, "means"), _abi.ptr(scales, F32, "scales"), 1.0,
                _abi.ptr(quats, F32, "quats"), sh_degree, flags, _abi.ptr(viewmats, F32, "viewmats"), vm_stride,
                _abi.ptr(intrins, F32, "intrins"), 0.0, 0.0, 0.0, 0.0, _abi.ptr(campos, F32, "campos"),
                img_height, img_width, block_width, clip_thresh, _abi.ptr(packed, F32, "packed"),
                _abi.ptr(v_packed, F32, "v_packed"), _abi.ptr(v_means, F32, "v_means"),
                _abi.ptr(v_scales, F32, "v_scales"), _abi.ptr(v_quats, F32, "v_quats"),
                _abi.ptr(v_opac, F32, "v_opacities"), _abi.ptr(v_color, F32, "v_color_src"),
                _abi.ptr(v_rest, F32, "v_color_rest"), st), "sb_views_project_bwd")
        return (v_means, v_scales, v_quats, v_opac, v_color, v_rest) + (None,) * (nin - 6)


def render_views(means: Tensor, scales: Tensor, quats: Tensor, opacities: Tensor, color_src: Tensor,
                 sh_degree: int, viewmats: Tensor, campos: Optional[Tensor], intrins: Tensor,
                 img_height: int, img_width: int, block_width: int = 16,
                 background: Optional[Tensor] = None, clip_thresh: float = 0.01, clamp_rgb: bool = False,
                 color_rest: Optional[Tensor] = None, flags: int = 0, gt: Optional[Tensor] = None):
    """Render all V views of one scene.

    means (N,3), scales (N,3), quats (N,4), opacities (N,1|N); activated values unless `flags`
    (_abi.FLAG_RAW_SCALES | FLAG_RAW_OPACITIES | FLAG_COLOR_SIGMOID) say the kernels apply
    exp / sigmoid themselves.  color_src: (N,K,3) SH coefficients with sh_degree in 0..4 (colour =
    max(SH + 0.5, 0) as gs_utils.py:78-79; or features_dc (N,3) with color_rest = features_rest
    (N,K-1,3)), or (N,3) rgb with sh_degree = -1.  viewmats (V,3|4,4) world->camera with campos (V,3)
    camera centres, or OpenGL camera-to-world matrices with FLAG_C2W_OPENGL; intrins (V,4) = fx, fy, cx, cy,
    or (1,4) shared by all views with FLAG_INTRINS_SHARED.

    Returns rgb (V,H,W,3) and alpha (V,H,W); rgb is not clamped unless clamp_rgb=True, which folds the
    wrapper's torch.clamp(rgb, max=1) (gs_utils.py:111) into the kernels.  With gt (V,H,W,3) given it
    returns instead (l1_sum (V,), sq_sum (V,), rgb): the per-view sums of |rgb - gt| (differentiable)
    and (rgb - gt)^2 (for PSNR, not differentiable) computed inside the raster kernel.
    Differentiable in means, scales, quats, opacities, color_src and color_rest."""
    assert 2 <= block_width <= 16, "block_width must be between 2 and 16"
    if not means.is_cuda:
        raise RuntimeError("render_views needs CUDA tensors (there is no CPU path)")
    if background is None:
        background = torch.ones(3, dtype=F32, device=means.device)
    want = (1, 4) if flags & _abi.FLAG_INTRINS_SHARED else (viewmats.shape[0], 4)
    if tuple(intrins.shape) != want:
        raise ValueError(f"intrins must be {want}, got {tuple(intrins.shape)}")
    if campos is None and not (flags & _abi.FLAG_C2W_OPENGL):
        campos = torch.zeros(viewmats.shape[0], 3, dtype=F32, device=means.device)
    return _RenderViews.apply(means, scales, quats, opacities, color_src, color_rest, viewmats, campos, intrins,
                              gt, int(sh_degree), int(flags), int(img_height), int(img_width), int(block_width),
                              background, float(clip_thresh), bool(clamp_rgb))
