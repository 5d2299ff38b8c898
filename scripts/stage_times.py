#!/usr/bin/env python
"""Per-stage CUDA-event times of one render step for a named config.  stage_times.py <config> <views> <fwd|fwdbwd>"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from splatformer_b200 import gs_utils, render, synth
name, views, mode = sys.argv[1], int(sys.argv[2]), sys.argv[3]
dev = torch.device("cuda:0")
if os.environ.get("SB_FORCE_GLOBAL"):
    render.DEFAULT_TILE_CAPACITY = 1 << 20
gs, cams, gt = synth.make_config(name, 0, num_views=views)
bw = mode == "fwdbwd"
raw = {k: v.to(dev).requires_grad_(bw) for k, v in gs.items()}
cams = {k: v.to(dev) for k, v in cams.items()}
gt = (gt * 255).round().to(torch.uint8).to(dev)
def step():
    if bw:
        for p in raw.values(): p.grad = None
        l1, _ = gs_utils.render_l1_loss(raw, cams, gt); l1.mean().backward()
    else:
        with torch.no_grad(): gs_utils.render_scene(raw, **cams)
for _ in range(3): step()
torch.cuda.synchronize()
with render.profile_stages() as t:
    for _ in range(3): step()
tot, cnt = t.totals_ms(), t.counts()
print(name, "isects", render.last_num_intersects(), "overflowed", render._default_plan.overflowed)
for k in tot: print(f"  {k:16s} {tot[k]/3:9.3f} ms/step ({cnt[k]//3} launches)")
print("  sum", sum(tot.values()) / 3)
