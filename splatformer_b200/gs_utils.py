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

from .render import render_views

BLOCK_WIDTH = 16  # /root/reference/utils/gs_utils.py:12


def viewmats_from_c2w(camera_to_worlds: Tensor) -> Tuple[Tensor, Tensor]:
    """(V,3|4,4) OpenGL camera-to-world -> ((V,3,4) OpenCV world-to-camera, (V,3) camera centres)."""
    c2w = camera_to_worlds.to(torch.float32)
    flip = torch.tensor([1.0, -1.0, -1.0], device=c2w.device, dtype=c2w.dtype)
    rot = c2w[:, :3, :3] * flip[None, None, :]          # R @ diag(1,-1,-1)
    trans = c2w[:, :3, 3:4]
    rot_inv = rot.transpose(1, 2)
    t_inv = -(rot_inv * trans.transpose(1, 2)).sum(dim=2, keepdim=True)      # -R^T T, no matmul
    viewmats = torch.cat([rot_inv, t_inv], dim=2)
    return viewmats.contiguous(), c2w[:, :3, 3].contiguous()


def _activate(gs_params: Dict[str, Tensor]):
    gp = {k: (v.float() if v.dtype == torch.half else v) for k, v in gs_params.items()}
    means = gp["means"]
    scales = torch.exp(gp["scales"])
    quats = gp["quats"] / torch.norm(gp["quats"], dim=-1, keepdim=True)
    if "opacities" in gp:
        opacities = torch.sigmoid(gp["opacities"])
    elif "opacities_sigmoid" in gp:
        opacities = gp["opacities_sigmoid"]
    else:
        raise ValueError("No opacities found in gs_params")
    if "features_rest" in gp:
        colors = torch.cat([gp["features_dc"].unsqueeze(1), gp["features_rest"]], dim=1)
    else:
        colors = gp["features_dc"].unsqueeze(1)
    degree = int(math.sqrt(colors.shape[1]) - 1)
    if degree == 0:
        color_src, sh_degree = torch.sigmoid(colors[:, 0, :]), -1
    else:
        color_src, sh_degree = colors, degree
    return means, scales, quats, opacities, color_src, sh_degree


def _scalar(x) -> float:
    return float(x.item()) if isinstance(x, Tensor) else float(x)


def render_scene(gs_params: Dict[str, Tensor], camera_to_worlds: Tensor, cx, cy, fx, fy, width, height,
                 background_color: Tensor, **kwargs) -> Tuple[Tensor, Tensor]:
    """All views at once: rgb (V,H,W,3) clamped to <= 1 and alpha (V,H,W,1)."""
    means, scales, quats, opacities, color_src, sh_degree = _activate(gs_params)
    viewmats, campos = viewmats_from_c2w(camera_to_worlds.to(means.device))
    v = viewmats.shape[0]
    if isinstance(height, Tensor) and isinstance(width, Tensor) and height.is_cuda:
        h, w = (int(x) for x in torch.stack([height, width]).tolist())     # one sync instead of two
    else:
        h, w = int(_scalar(height)), int(_scalar(width))
    if all(isinstance(x, Tensor) for x in (fx, fy, cx, cy)):
        intr = torch.stack([fx, fy, cx, cy]).to(device=means.device, dtype=torch.float32)  # no .item()
    else:
        intr = torch.tensor([_scalar(fx), _scalar(fy), _scalar(cx), _scalar(cy)],
                            dtype=torch.float32).to(means.device, non_blocking=True)
    intrins = intr.reshape(1, 4).expand(v, 4).contiguous()
    rgb, alpha = render_views(means, scales, quats, opacities, color_src, sh_degree, viewmats, campos, intrins,
                              h, w, BLOCK_WIDTH, background=background_color.to(means.device), clamp_rgb=True)
    return rgb, alpha.unsqueeze(-1)


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
