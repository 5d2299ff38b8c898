#!/usr/bin/env python
"""Fraction of (view, Gaussian) gradient records the projection backward actually has to differentiate."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from splatformer_b200 import gs_utils, render, synth
dev = torch.device("cuda:0"); torch.cuda.set_device(dev)
gs, cams, gt = synth.make_config("objaverse_100k_sh3_512", 0)
gt = (gt * 255).round().to(torch.uint8).to(dev)
cams = {k: v.to(dev) for k, v in cams.items()}
raw = {k: v.to(dev).requires_grad_() for k, v in gs.items()}
render.COLLECT_BWD_STATS = True
l1, _, img = gs_utils.render_l1_loss(raw, cams, gt, return_images=True)
ps = img.grad_fn.ps
(l1.sum() / 32).backward()
scanned, live = render.bwd_stats()
drawn = int((ps.radii > 0).sum())
print(f"records scanned {scanned}, drawn (radius > 0) {drawn}, with a non-zero gradient {live} = {live/scanned:.3f} of all, {live/max(drawn,1):.3f} of drawn")
