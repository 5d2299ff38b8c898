import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
from splatformer_b200 import gs_utils, render, synth
dev = torch.device("cuda:0")
def run(name, fn):
    try:
        r = fn(); torch.cuda.synchronize(); print("OK  ", name, r)
    except Exception as e:
        print("FAIL", name, type(e).__name__, str(e)[:200])
def scene(n, deg, **kw):
    gs = synth.shell_scene(max(n, 1), deg, seed=1, **kw)
    return {k: v[:n].to(dev).requires_grad_() for k, v in gs.items()}
cams = {k: v.to(dev) for k, v in synth.orbit_cameras(2, 40, 24).items()}
gt = torch.rand(2, 24, 40, 3, device=dev)
def fused(n, deg, cams=cams, gt=gt):
    raw = scene(n, deg)
    l1, mse = gs_utils.render_l1_loss(raw, cams, gt)
    l1.sum().backward()
    return [round(float(x), 4) for x in l1], all(torch.isfinite(p.grad).all().item() for p in raw.values() if p.grad is not None)
run("N=0 fused", lambda: fused(0, 1))
run("N=1 fused", lambda: fused(1, 3))
run("N=5 deg0", lambda: fused(5, 0))
small = {k: v.to(dev) for k, v in synth.orbit_cameras(1, 7, 5).items()}
run("7x5 image", lambda: fused(300, 2, small, torch.rand(1, 5, 7, 3, device=dev)))
def half():
    raw = {k: v.half() for k, v in scene(200, 1).items()}
    rgbs, alphas = gs_utils.rasterize_gaussians_to_multiimgs(raw, cams)
    return rgbs[0].dtype, float(alphas[0].mean())
run("half inputs", half)
def nan_means():
    raw = scene(300, 1)
    with torch.no_grad(): raw["means"][:10] = float("nan"); raw["scales"][10:20] = float("inf")
    l1, _ = gs_utils.render_l1_loss(raw, cams, gt); l1.sum().backward()
    return [float(x) for x in l1], bool(torch.isfinite(raw["means"].grad[20:]).all())
run("nan/inf Gaussians", nan_means)
def huge():
    raw = scene(50, 1)
    with torch.no_grad(): raw["scales"] += 5.0
    l1, _ = gs_utils.render_l1_loss(raw, cams, gt); l1.sum().backward()
    return [round(float(x), 4) for x in l1], render.last_num_intersects()
run("huge Gaussians", huge)
def perview_empty():
    raw = scene(0, 1)
    r, a = gs_utils.per_view_multiimgs(raw, cams)
    return r[0].shape, float(a[0].mean())
run("per-view N=0", perview_empty)
