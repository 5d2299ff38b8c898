#!/usr/bin/env python
"""Host-side cost of one fused render step (tiny scene: the GPU work is negligible), cProfile top entries."""
import cProfile, os, pstats, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from splatformer_b200 import gs_utils, synth
dev = torch.device("cuda:0")
gs = synth.shell_scene(2000, 3, seed=1, scale_mult=0.3)
cams = {k: v.to(dev) for k, v in synth.orbit_cameras(4, 64, 64).items()}
gt = (synth.target_images(4, 64, 64) * 255).round().to(torch.uint8).to(dev)
raw = {k: v.to(dev).requires_grad_() for k, v in gs.items()}
def step():
    for p in raw.values(): p.grad = None
    l1, _ = gs_utils.render_l1_loss(raw, cams, gt)
    (l1.sum() * 0.25).backward()
for _ in range(20): step()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(300): step()
torch.cuda.synchronize()
print("wall per step: %.1f us" % ((time.perf_counter() - t0) / 300 * 1e6))
pr = cProfile.Profile(); pr.enable()
for _ in range(300): step()
torch.cuda.synchronize(); pr.disable()
st = pstats.Stats(pr); st.sort_stats("tottime").print_stats(22)
