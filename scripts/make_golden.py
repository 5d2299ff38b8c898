#!/usr/bin/env python
"""Freezes golden vectors for the render path under tests/golden/ from the ORACLE
(oracle/ref_torch.py, float32, CPU).  The reference owns no fixtures for this path and gsplat
v0.1.11 itself is not importable here, so the oracle's outputs -- pinned by
tests/test_oracle_known_answers.py and tests/test_oracle_crosscheck.py -- are what gets frozen.

    python scripts/make_golden.py          # rewrites tests/golden/*.npz

Each file holds the raw SplatFormer-style inputs, the cameras, and per view: the projection
outputs, intersection count, sorted ids, tile bins, rendered rgb / alpha, the per-pixel decision
margin, plus the gradients of sum_v mean(|rgb_v - gt_v| * solid_v) w.r.t. the raw parameters -- from the
float32 oracle (grad_*, what the CPU test re-derives) and from the FLOAT64 oracle with the same pixel mask
(grad64_*, the reference of the GPU test's 1e-3 tolerance)."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_torch as R          # noqa: E402
from splatformer_b200 import synth         # noqa: E402

CASES = {
    # name: (n, sh degree, views, W, H, scale_mult, seed, background)
    "plumbing_4k_sh0_128": (4096, 0, 1, 128, 128, 0.15, 1000, (0.0, 0.0, 0.0)),
    "small_sh3_ragged": (1500, 3, 2, 104, 72, 0.3, 41, (0.1, 0.2, 0.3)),
    "small_sh1_dense": (800, 1, 2, 64, 64, 0.6, 42, (1.0, 1.0, 1.0)),
}
MARGIN = 2e-5


def make(name):
    n, deg, views, w, h, mult, seed, bg = CASES[name]
    gs = synth.shell_scene(n, deg, seed=seed, scale_mult=mult)
    cams = synth.orbit_cameras(views, w, h, background=bg)
    gt = synth.target_images(views, h, w, seed=seed + 1)
    out = {f"in_{k}": v.numpy() for k, v in gs.items()}
    out.update({f"cam_{k}": v.numpy() for k, v in cams.items()})
    out["gt"] = gt.numpy()
    raw = {k: v.clone().requires_grad_() for k, v in gs.items()}
    loss = 0.0
    fx, fy, cx, cy = (float(cams[k]) for k in ("fx", "fy", "cx", "cy"))
    for v in range(views):
        c2w = cams["camera_to_worlds"][v]
        rgb, alpha, aux = R.render_view(raw, c2w, cx, cy, fx, fy, w, h, cams["background_color"], return_aux=True)
        solid = aux["margin"] > MARGIN
        loss = loss + ((rgb - gt[v]).abs() * solid[..., None]).mean()
        with torch.no_grad():
            vm = R.viewmat_from_c2w(c2w)[:3].contiguous()
            scales = torch.exp(gs["scales"])
            quats = gs["quats"] / gs["quats"].norm(dim=-1, keepdim=True)
            proj = R.project_gaussians(gs["means"], scales, 1, quats, vm, fx, fy, cx, cy, h, w, 16)
        for nm, t in zip(("xys", "depths", "radii", "conics", "compensation", "num_tiles_hit", "cov3d"), proj):
            out[f"v{v}_{nm}"] = t.numpy()
        out[f"v{v}_viewmat"] = vm.numpy()
        out[f"v{v}_num_intersects"] = np.int64(aux["num_intersects"])
        out[f"v{v}_isect_ids_sorted"] = aux["isect_ids_sorted"].numpy()
        out[f"v{v}_gaussian_ids_sorted"] = aux["gaussian_ids_sorted"].numpy()
        out[f"v{v}_tile_bins"] = aux["tile_bins"].numpy()
        out[f"v{v}_rgb"] = rgb.detach().numpy()
        out[f"v{v}_alpha"] = alpha.detach().numpy()
        out[f"v{v}_solid"] = solid.numpy()
    loss.backward()
    out["loss"] = np.float64(loss.detach())
    for k in gs:
        out[f"grad_{k}"] = raw[k].grad.numpy()
    raw64 = {k: v.double().requires_grad_() for k, v in gs.items()}
    loss64 = 0.0
    for v in range(views):
        rgb64, _ = R.render_view(raw64, cams["camera_to_worlds"][v].double(), cx, cy, fx, fy, w, h,
                                 cams["background_color"])
        solid = torch.from_numpy(out[f"v{v}_solid"])
        loss64 = loss64 + ((rgb64 - gt[v].double()).abs() * solid[..., None]).mean()
    loss64.backward()
    for k in gs:
        out[f"grad64_{k}"] = raw64[k].grad.numpy()
    path = os.path.join(ROOT, "tests", "golden", name + ".npz")
    np.savez_compressed(path, **out)
    print(name, os.path.getsize(path) // 1024, "KiB", "loss", float(loss))


if __name__ == "__main__":
    for name in CASES:
        make(name)
