#!/usr/bin/env python
"""Work statistics of the rect-list rasteriser on the bench workload (run on the GPU box):
intersections, (rectangle, Gaussian) list entries, entries per intersection, list-length histogram."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from splatformer_b200 import gs_utils, render, synth

dev = torch.device("cuda:0")
gs, cams, gt = synth.make_config("objaverse_100k_sh3_512", 0)
raw = {k: v.to(dev).requires_grad_() for k, v in gs.items()}
l1, _ = gs_utils.render_l1_loss(raw, {k: v.to(dev) for k, v in cams.items()}, gt.to(dev))
fn = l1.grad_fn
while not hasattr(fn, "saved_tensors"):
    fn = fn.next_functions[0][0]
saved = fn.saved_tensors
rect_counts = saved[15].long()
isect = render.last_num_intersects()
entries = int(rect_counts.sum())
nz = rect_counts[rect_counts > 0]
print(f"intersections {isect}  rect entries {entries}  entries/intersection {entries / isect:.3f}")
print(f"non-empty rects {nz.numel()} of {rect_counts.numel()}  mean len {nz.float().mean():.1f}  "
      f"median {nz.float().median():.0f}  p90 {nz.float().quantile(0.9):.0f}  max {int(nz.max())}")
chunks = ((nz + 31) // 32).sum()
print(f"32-entry chunks {int(chunks)}  chunk fill {entries / (32 * int(chunks)):.3f}")
