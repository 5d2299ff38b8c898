"""Drop-in for the two render wrappers of SplatFormer's `utils/gs_utils.py` (rows a1/a2 of the
hot-path table): same names, arguments, return values and conventions as
`rasterize_gaussians_to_multiimgs` (/root/reference/utils/gs_utils.py:20-27) and
`rasterize_gaussians_to_singleimg` (:29-114), implemented on the fused multi-view path.

What is kept exactly: OpenGL camera-to-world -> OpenCV world-to-camera (:32-42), exp / normalise /
sigmoid activations (:45-56), the `opacities` | `opacities_sigmoid` keys, features_dc + features_rest
concatenation (:59-62), degree-0 colours = sigmoid(dc) without SH (:64-65), clamp(SH + 0.5, min 0)
(:78-79), clamp(rgb, max 1) and alpha[..., None] (:111-112), lists of per-view tensors as return value.
What changes: the view-independent work is hoisted out of the view loop, all views go through one
launch per stage, and the nine host synchronisations per view of the reference become one per scene.
"""
from __future__ import annotations

import math
from typing import Dict, List, Tuple

import torch
from torch import Tensor

from . import _abi
from .render import render_views

BLOCK_WIDTH = 16  # /root/reference/utils/gs_utils.py:12


def viewmats_from_c2w(camera_to_worlds: Tensor) -> Tuple[Tensor, Tensor]:
    """(V,3|4,4) OpenGL camera-to-world -> ((V,3,4) OpenCV world-to-camera, (V,3) camera centres).
    The fused path does this inside the projection kernel (FLAG_C2W_OPENGL); this torch version is kept
    for callers of the per-view API."""
    c2w = camera_to_worlds.to(torch.float32)
    flip = torch.tensor([1.0, -1.0, -1.0], device=c2w.device, dtype=c2w.dtype)
    rot = c2w[:, :3, :3] * flip[None, None, :]          # R @ diag(1,-1,-1)
    trans = c2w[:, :3, 3:4]
    rot_inv = rot.transpose(1, 2)
    t_inv = -(rot_inv * trans.transpose(1, 2)).sum(dim=2, keepdim=True)      # -R^T T, no matmul
    viewmats = torch.cat([rot_inv, t_inv], dim=2)
    return viewmats.contiguous(), c2w[:, :3, 3].contiguous()


def _raw_inputs(gs_params: Dict[str, Tensor]):
    """Raw SplatFormer parameters -> kernel arguments + flags.  Nothing is activated or concatenated
    here: exp / normalise / sigmoid (gs_utils.py:45-56), the features_dc | features_rest split (:59-62)
    and the degree-0 sigmoid colour (:65) are applied inside the projection kernels, and the backward
    returns gradients w.r.t. these raw tensors directly."""
    gp = {k: (v.float() if v.dtype == torch.half else v) for k, v in gs_params.items()}
    flags = _abi.FLAG_RAW_SCALES | _abi.FLAG_RAW_QUATS | _abi.FLAG_C2W_OPENGL
    if "opacities" in gp:
        opacities, flags = gp["opacities"], flags | _abi.FLAG_RAW_OPACITIES
    elif "opacities_sigmoid" in gp:
        opacities = gp["opacities_sigmoid"]
    else:
        raise ValueError("No opacities found in gs_params")
    dc = gp["features_dc"]
    rest = gp.get("features_rest")
    k = 1 + (rest.shape[1] if rest is not None else 0)
    degree = int(math.sqrt(k) - 1)
    if degree == 0:
        rest, sh_degree, flags = None, -1, flags | _abi.FLAG_COLOR_SIGMOID
    else:
        if (degree + 1) ** 2 != k:
            raise ValueError(f"features_dc + features_rest hold {k} SH bases, not a square number")
        sh_degree = degree
    return gp["means"], gp["scales"], gp["quats"], opacities, dc, rest, sh_degree, flags


def _scalar(x) -> float:
    return float(x.item()) if isinstance(x, Tensor) else float(x)


def _intrinsics(fx, fy, cx, cy, v: int, device) -> Tensor:
    """(1,4) fx fy cx cy shared by all views (FLAG_INTRINS_SHARED): one small kernel, no .item()."""
    if all(isinstance(x, Tensor) for x in (fx, fy, cx, cy)):
        intr = torch.stack([fx.reshape(()), fy.reshape(()), cx.reshape(()), cy.reshape(())]).to(
            device=device, dtype=torch.float32)                                            # no .item()
    else:
        intr = torch.tensor([_scalar(fx), _scalar(fy), _scalar(cx), _scalar(cy)],
                            dtype=torch.float32).to(device, non_blocking=True)
    return intr.reshape(1, 4)


def render_scene(gs_params: Dict[str, Tensor], camera_to_worlds: Tensor, cx, cy, fx, fy, width, height,
                 background_color: Tensor, plan=None, **kwargs) -> Tuple[Tensor, Tensor]:
    """All views at once: rgb (V,H,W,3) clamped to <= 1 and alpha (V,H,W,1)."""
    means, scales, quats, opacities, dc, rest, sh_degree, flags = _raw_inputs(gs_params)
    c2w = camera_to_worlds.to(means.device)
    v = c2w.shape[0]
    if isinstance(height, Tensor) and isinstance(width, Tensor) and height.is_cuda:
        h, w = (int(x) for x in torch.stack([height, width]).tolist())     # one sync instead of two
    else:
        h, w = int(_scalar(height)), int(_scalar(width))
    intrins = _intrinsics(fx, fy, cx, cy, v, means.device)
    rgb, alpha = render_views(means, scales, quats, opacities, dc, sh_degree, c2w, None, intrins, h, w,
                              BLOCK_WIDTH, background=background_color.to(means.device), clamp_rgb=True,
                              color_rest=rest, flags=flags | _abi.FLAG_INTRINS_SHARED, plan=plan)
    return rgb, alpha.unsqueeze(-1)


def render_l1_loss(gs_params: Dict[str, Tensor], cameras: Dict[str, Tensor], images: Tensor,
                   return_images: bool = False, plan=None):
    """The render loss of SplatFormer's training step in one call (train.py:272-278):

        pred_imgs, _ = rasterize_gaussians_to_multiimgs(out_gs, cameras)
        for pred, gt in zip(pred_imgs, images): l1 += abs(pred - gt).mean(); psnr(pred, gt)

    images: (V,H,W,3) float32 targets in [0,1] on the device.  Returns (l1 (V,), mse (V,)) -- the
    per-view means of |pred - gt| (differentiable w.r.t. the raw Gaussian parameters) and of
    (pred - gt)^2 (for PSNR = -10 log10 mse; not differentiable) -- and the predicted images
    (V,H,W,3) if return_images; those are differentiable as well, so an image-space term (the
    reference's LPIPS loss, train.py:276-277) can be added on top of the fused L1: its gradient image
    enters the raster backward next to sign(pred - gt).  The image size comes from `images`, so no host
    synchronisation is spent on cameras['height'/'width']; the L1 loss and its gradient image are
    computed inside the raster kernels and never exist as tensors.  `plan`: render.RenderPlan of the
    call site (binning capacities learnt from earlier scenes); the module default if omitted."""
    means, scales, quats, opacities, dc, rest, sh_degree, flags = _raw_inputs(gs_params)
    if images.dim() != 4 or images.shape[-1] != 3:
        raise ValueError("images must be (V, H, W, 3)")
    c2w = cameras["camera_to_worlds"].to(means.device)
    v, h, w = c2w.shape[0], int(images.shape[1]), int(images.shape[2])
    if images.shape[0] != v:
        raise ValueError(f"{images.shape[0]} target images for {v} cameras")
    intrins = _intrinsics(cameras["fx"], cameras["fy"], cameras["cx"], cameras["cy"], v, means.device)
    l1_sum, sq_sum, rgb = render_views(means, scales, quats, opacities, dc, sh_degree, c2w, None, intrins, h, w,
                                       BLOCK_WIDTH, background=cameras["background_color"].to(means.device),
                                       clamp_rgb=True, color_rest=rest, flags=flags | _abi.FLAG_INTRINS_SHARED, gt=images,
                                       plan=plan)
    inv = 1.0 / (h * w * 3)
    out = (l1_sum * inv, sq_sum * inv)
    return out + (rgb,) if return_images else out


def rasterize_gaussians_to_multiimgs(gs_params: Dict[str, Tensor], cameras: Dict[str, Tensor]
                                     ) -> Tuple[List[Tensor], List[Tensor]]:
    """Same contract as the reference: two lists of V tensors, (H,W,3) and (H,W,1)."""
    rgb, alpha = render_scene(gs_params, **cameras)
    return list(rgb.unbind(0)), list(alpha.unbind(0))


def rasterize_gaussians_to_singleimg(gs_params: Dict[str, Tensor], camera_to_world: Tensor, cx, cy, fx, fy,
                                     width, height, background_color: Tensor, **kwargs
                                     ) -> Tuple[Tensor, Tensor]:
    rgb, alpha = render_scene(gs_params, camera_to_world[None], cx, cy, fx, fy, width, height,
                              background_color)
    return rgb[0], alpha[0]


# ------------------------------------------------------------------------------------------------
# Drop-in level 1: the reference's own per-view wrapper logic on the gsplat-compatible API.
# This is what SplatFormer runs when only `import gsplat` is redirected to this repository
# (three API calls and about nine host synchronisations per view); bench.py times it next to the
# fused path ("wrapper_path"), and the tests use it to tie the fused path to the per-view API.
# ------------------------------------------------------------------------------------------------
def per_view_singleimg(gs_params: Dict[str, Tensor], camera_to_world: Tensor, cx, cy, fx, fy, width, height,
                       background_color: Tensor, **kwargs) -> Tuple[Tensor, Tensor]:
    """One view through spherical_harmonics / project_gaussians / rasterize_gaussians with the
    conventions of /root/reference/utils/gs_utils.py:29-114."""
    from . import api
    gp = {k: (v.float() if v.dtype == torch.half else v) for k, v in gs_params.items()}
    viewmats, campos = viewmats_from_c2w(camera_to_world[None])                 # :32-42
    means = gp["means"]
    scales = torch.exp(gp["scales"])                                            # :45
    quats = gp["quats"] / torch.norm(gp["quats"], dim=-1, keepdim=True)         # :46
    if "opacities" in gp:                                                       # :53-58
        opacities = torch.sigmoid(gp["opacities"])
    elif "opacities_sigmoid" in gp:
        opacities = gp["opacities_sigmoid"]
    else:
        raise ValueError("No opacities found in gs_params")
    coeffs = gp["features_dc"].unsqueeze(1)
    if "features_rest" in gp:
        coeffs = torch.cat([coeffs, gp["features_rest"]], dim=1)                # :59-62
    degree = int(math.sqrt(coeffs.shape[1]) - 1)
    if degree == 0:
        rgbs = torch.sigmoid(coeffs[:, 0, :])                                   # :65
    else:
        dirs = means.detach() - campos[0]
        dirs = dirs / dirs.norm(dim=-1, keepdim=True)
        rgbs = torch.clamp(api.spherical_harmonics(degree, dirs, coeffs) + 0.5, min=0.0)   # :78-79
    h, w = int(_scalar(height)), int(_scalar(width))                            # host syncs, as upstream (:80)
    xys, depths, radii, conics, _comp, num_tiles_hit, _cov3d = api.project_gaussians(
        means, scales, 1, quats, viewmats[0], _scalar(fx), _scalar(fy), _scalar(cx), _scalar(cy), h, w, BLOCK_WIDTH)
    rgb, alpha = api.rasterize_gaussians(xys, depths, radii, conics, num_tiles_hit, rgbs, opacities, h, w,
                                         BLOCK_WIDTH, background=background_color, return_alpha=True)
    return torch.clamp(rgb, max=1.0), alpha.unsqueeze(-1)                       # :111-112


def per_view_multiimgs(gs_params: Dict[str, Tensor], cameras: Dict[str, Tensor]) -> Tuple[List[Tensor], List[Tensor]]:
    """The reference's Python loop over the views (/root/reference/utils/gs_utils.py:20-27)."""
    rgbs, alphas = [], []
    for c2w in cameras["camera_to_worlds"]:
        rgb, alpha = per_view_singleimg(gs_params, c2w, **cameras)
        rgbs.append(rgb)
        alphas.append(alpha)
    return rgbs, alphas
