#!/usr/bin/env python
"""Per-step device times and stage shares of one BASELINE shape (default: config 5, 6 M Gaussians at 4K)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from splatformer_b200 import gs_utils, render, synth
name = sys.argv[1] if len(sys.argv) > 1 else "stress_6m_4k"
views = int(sys.argv[2]) if len(sys.argv) > 2 else 1
dev = torch.device("cuda:0")
gs, cams, gt = synth.make_config(name, 0, num_views=views)
raw = {k: v.to(dev).requires_grad_(True) for k, v in gs.items()}
cams = {k: v.to(dev) for k, v in cams.items()}
gt = (gt * 255).round().to(torch.uint8).to(dev)
def step():
    for p in raw.values(): p.grad = None
    l1, _ = gs_utils.render_l1_loss(raw, cams, gt)
    l1.mean().backward()
for i in range(8):
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); step(); b.record(); torch.cuda.synchronize()
    st = torch.cuda.memory_stats()
    print(f"step {i}: {a.elapsed_time(b):8.2f} ms  reserved {torch.cuda.memory_reserved() / 1e9:6.2f} GB  "
          f"cudaMalloc calls {st['num_device_alloc']}  frees {st['num_device_free']}  retries {st['num_alloc_retries']}")
with render.profile_stages() as t:
    step()
print({k: round(v, 3) for k, v in t.totals_ms().items()})
