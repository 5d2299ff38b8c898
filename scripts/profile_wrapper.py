#!/usr/bin/env python
"""torch.profiler view of the per-view wrapper path (gs_utils.per_view_multiimgs + L1 + backward)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import profile, ProfilerActivity
from splatformer_b200 import gs_utils, synth

dev = torch.device("cuda:0")
gs, cams, gt = synth.make_config("objaverse_100k_sh3_512", 0)
raw = {k: v.to(dev).requires_grad_() for k, v in gs.items()}
cams = {k: v.to(dev) for k, v in cams.items()}
gt = gt.to(dev)

def step():
    for p in raw.values():
        p.grad = None
    rgbs, _ = gs_utils.per_view_multiimgs(raw, cams)
    loss = 0.0
    for rgb, g in zip(rgbs, gt):
        loss = loss + (rgb - g).abs().mean()
    (loss / 32).backward()

for _ in range(2):
    step()
torch.cuda.synchronize()
t0 = time.perf_counter(); step(); torch.cuda.synchronize(); print("step wall ms", (time.perf_counter() - t0) * 1e3)
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    step()
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=25, max_name_column_width=60))
print(prof.key_averages().table(sort_by="self_cpu_time_total", row_limit=15, max_name_column_width=60))
