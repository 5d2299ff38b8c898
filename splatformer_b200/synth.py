"""Seeded synthetic scenes and cameras of the BASELINE.json shapes.

Tensors follow what SplatFormer feeds the render path: raw (pre-activation)
Gaussian parameters as produced by `/root/reference/dataset/GS.py:163-193` /
`/root/reference/models/feature_predictor.py:227-235` (means in [0,1]^3, log
scales, unnormalised wxyz quaternions, opacity logits, SH coefficients split
into `features_dc` (N,3) and `features_rest` (N,K-1,3)) and a `cameras` dict
with OpenGL camera-to-world matrices plus shared intrinsics
(`/root/reference/dataset/GS.py:386-394`).

Everything is generated on the CPU from `torch.Generator().manual_seed(...)`
so that every rank, the oracle and the CUDA path see identical inputs.
"""
from __future__ import annotations

import math
from typing import Dict, Tuple

import torch

GOLDEN_ANGLE = math.pi * (3.0 - math.sqrt(5.0))


def _unit_sphere(n: int, gen: torch.Generator) -> torch.Tensor:
    v = torch.randn(n, 3, generator=gen)
    return v / v.norm(dim=1, keepdim=True).clamp_min(1e-12)


def shell_scene(num_gaussians: int, sh_degree: int, seed: int = 1000, *, scale_mult: float = 0.15,
                scale_sigma: float = 0.5) -> Dict[str, torch.Tensor]:
    """Object-like shell of Gaussians around (0.5,0.5,0.5) (configs 1-3, 5)."""
    gen = torch.Generator().manual_seed(seed)
    n = num_gaussians
    u = _unit_sphere(n, gen)
    rad = 0.30 * (0.8 + 0.2 * torch.rand(n, 1, generator=gen))
    means = 0.5 + u * rad
    base = math.log(scale_mult * math.sqrt(4.0 * math.pi / n))
    scales = base + scale_sigma * torch.randn(n, 3, generator=gen)
    quats = torch.randn(n, 4, generator=gen)
    opacities = 2.0 + 2.0 * torch.randn(n, 1, generator=gen)
    features_dc = 0.5 * torch.randn(n, 3, generator=gen)
    gs = dict(means=means, scales=scales, quats=quats, opacities=opacities, features_dc=features_dc)
    k = (sh_degree + 1) ** 2
    if k > 1:
        gs["features_rest"] = 0.1 * torch.randn(n, k - 1, 3, generator=gen)
    return gs


def room_scene(num_gaussians: int, sh_degree: int, seed: int = 1000) -> Dict[str, torch.Tensor]:
    """Room-like volume + ground plane (config 4)."""
    gen = torch.Generator().manual_seed(seed)
    n = num_gaussians
    n_vol = int(0.7 * n)
    vol = 0.05 + 0.9 * torch.rand(n_vol, 3, generator=gen)
    gnd = 0.05 + 0.9 * torch.rand(n - n_vol, 3, generator=gen)
    gnd[:, 2] = 0.1
    means = torch.cat([vol, gnd], dim=0)
    scales = (math.log(0.004) + torch.randn(n, 3, generator=gen)).clamp_max(math.log(0.05))
    quats = torch.randn(n, 4, generator=gen)
    opacities = 2.0 + 2.0 * torch.randn(n, 1, generator=gen)
    features_dc = 0.5 * torch.randn(n, 3, generator=gen)
    gs = dict(means=means, scales=scales, quats=quats, opacities=opacities, features_dc=features_dc)
    k = (sh_degree + 1) ** 2
    if k > 1:
        gs["features_rest"] = 0.1 * torch.randn(n, k - 1, 3, generator=gen)
    return gs


def look_at_c2w(eye: torch.Tensor, target: torch.Tensor, up=(0.0, 0.0, 1.0)) -> torch.Tensor:
    """OpenGL camera-to-world (x right, y up, z back), 4x4 float32."""
    eye = eye.to(torch.float64)
    fwd = target.to(torch.float64) - eye
    fwd = fwd / fwd.norm()
    upv = torch.tensor(up, dtype=torch.float64)
    right = torch.linalg.cross(fwd, upv)
    if right.norm() < 1e-8:
        right = torch.linalg.cross(fwd, torch.tensor((0.0, 1.0, 0.0), dtype=torch.float64))
    right = right / right.norm()
    true_up = torch.linalg.cross(right, fwd)
    c2w = torch.eye(4, dtype=torch.float64)
    c2w[:3, 0], c2w[:3, 1], c2w[:3, 2], c2w[:3, 3] = right, true_up, -fwd, eye
    return c2w.to(torch.float32)


def orbit_cameras(num_views: int, width: int, height: int, *, radius: float = 1.3,
                  elev_deg: Tuple[float, float] = (10.0, 80.0), fov_half_deg: float = 25.0,
                  centre=(0.5, 0.5, 0.5), background=(0.0, 0.0, 0.0), focal: float | None = None,
                  azim0_deg: float = 45.0) -> Dict[str, torch.Tensor]:
    """V cameras on a sphere looking at `centre` (Appendix B): elevations linear in
    `elev_deg`, azimuths in golden-angle steps; shared pinhole intrinsics."""
    tgt = torch.tensor(centre, dtype=torch.float64)
    c2ws = []
    for v in range(num_views):
        frac = 0.5 if num_views == 1 else v / (num_views - 1)
        elev = math.radians(elev_deg[0] + (elev_deg[1] - elev_deg[0]) * frac) if num_views > 1 \
            else math.radians(30.0)
        azim = math.radians(azim0_deg) + v * GOLDEN_ANGLE
        eye = tgt + radius * torch.tensor(
            [math.cos(elev) * math.cos(azim), math.cos(elev) * math.sin(azim), math.sin(elev)],
            dtype=torch.float64)
        c2ws.append(look_at_c2w(eye, tgt))
    f = focal if focal is not None else 0.5 * width / math.tan(math.radians(fov_half_deg))
    return dict(camera_to_worlds=torch.stack(c2ws, dim=0),
                fx=torch.tensor(f, dtype=torch.float32), fy=torch.tensor(f, dtype=torch.float32),
                cx=torch.tensor(width / 2.0, dtype=torch.float32),
                cy=torch.tensor(height / 2.0, dtype=torch.float32),
                width=torch.tensor(width), height=torch.tensor(height),
                background_color=torch.tensor(background, dtype=torch.float32))


def ring_cameras(num_views: int, width: int, height: int, *, radius: float = 0.35, z: float = 0.5,
                 focal: float = 1500.0, background=(0.0, 0.0, 0.0)) -> Dict[str, torch.Tensor]:
    """Eval views inside a room, on a ring looking inward (config 4)."""
    tgt = torch.tensor((0.5, 0.5, z), dtype=torch.float64)
    c2ws = []
    for v in range(num_views):
        a = 2.0 * math.pi * v / num_views
        eye = tgt + radius * torch.tensor([math.cos(a), math.sin(a), 0.0], dtype=torch.float64)
        c2ws.append(look_at_c2w(eye, tgt))
    return dict(camera_to_worlds=torch.stack(c2ws, dim=0),
                fx=torch.tensor(focal, dtype=torch.float32), fy=torch.tensor(focal, dtype=torch.float32),
                cx=torch.tensor(width / 2.0, dtype=torch.float32),
                cy=torch.tensor(height / 2.0, dtype=torch.float32),
                width=torch.tensor(width), height=torch.tensor(height),
                background_color=torch.tensor(background, dtype=torch.float32))


def target_images(num_views: int, height: int, width: int, seed: int = 3000) -> torch.Tensor:
    gen = torch.Generator().manual_seed(seed)
    return torch.rand(num_views, height, width, 3, generator=gen)


CONFIGS = {
    # name: (scene fn, N, sh degree, V, W, H, camera fn kwargs)
    "plumbing_4k_sh0_128": dict(n=4096, deg=0, views=1, width=128, height=128),
    "objaverse_100k_sh3_512": dict(n=100_000, deg=3, views=32, width=512, height=512),
    "real_1m_sh3_1080p": dict(n=1_000_000, deg=3, views=8, width=1920, height=1080),
    "stress_6m_4k": dict(n=6_000_000, deg=3, views=2, width=3840, height=2160),
}


def make_config(name: str, scene_index: int = 0, num_views: int | None = None):
    """Returns (gs_params, cameras, target images) for a named BASELINE.json config."""
    c = CONFIGS[name]
    v = num_views if num_views is not None else c["views"]
    s = scene_index
    if name == "real_1m_sh3_1080p":
        gs = room_scene(c["n"], c["deg"], seed=1000 + s)
        cams = ring_cameras(v, c["width"], c["height"])
    elif name == "stress_6m_4k":
        n = c["n"]
        gs = shell_scene(n, c["deg"], seed=1000 + s, scale_sigma=0.25)
        gs["scales"] = gs["scales"] - math.log(0.15 * math.sqrt(4.0 * math.pi / n)) + math.log(0.008)
        cams = orbit_cameras(v, c["width"], c["height"], radius=1.5, focal=3000.0)
    else:
        gs = shell_scene(c["n"], c["deg"], seed=1000 + s)
        cams = orbit_cameras(v, c["width"], c["height"])
    gt = target_images(v, c["height"], c["width"], seed=3000 + s)
    return gs, cams, gt
