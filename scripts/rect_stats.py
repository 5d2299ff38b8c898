#!/usr/bin/env python
"""Length distribution of the rect lists of the benchmark scene (what the rasterisers walk)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from splatformer_b200 import gs_utils, render, synth
dev = torch.device("cuda:0"); torch.cuda.set_device(dev)
gs, cams, gt = synth.make_config("objaverse_100k_sh3_512", 0)
gt = (gt * 255).round().to(torch.uint8).to(dev)
raw = {k: v.to(dev).requires_grad_() for k, v in gs.items()}
l1, _, img = gs_utils.render_l1_loss(raw, {k: v.to(dev) for k, v in cams.items()}, gt, return_images=True)
ps = img.grad_fn.ps
cnt = ps.rect_meta[:, 1].long()
walked = (ps.final_idx.view(32, 128, 4, 64, 8).permute(0, 1, 3, 2, 4).reshape(-1, 32).amax(dim=1) + 1).clamp_min(0)   # per rect
nz = cnt[cnt > 0]
q = torch.quantile(nz.float(), torch.tensor([0.1, 0.5, 0.9, 0.99], device=dev)).tolist()
print(f"rects {cnt.numel()}, non-empty {nz.numel()} ({nz.numel()/cnt.numel():.3f}); entries {int(cnt.sum())} "
      f"(pool reserved {ps.rect_ids.numel()} ints, intersections {render.last_num_intersects()})")
print(f"non-empty list length: mean {nz.float().mean():.1f}  p10/p50/p90/p99 {q}  max {int(nz.max())}")
print(f"lists longer than 64 entries: {(nz > 64).float().mean():.3f} of the non-empty ones, holding "
      f"{cnt[cnt > 64].sum().item() / cnt.sum().item():.3f} of all entries")
print(f"entries up to the last composited one (what the backward walks): {int(walked.sum())} = "
      f"{int(walked.sum()) / int(cnt.sum()):.3f} of all")
