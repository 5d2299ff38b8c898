#!/usr/bin/env python
"""What a split of every 8x4 rectangle into two 4x4 halves (walked by the two half-warps in lockstep,
each over its own sub-list) would save on the bench workload: steps = max(|L|, |R|) instead of |L u R|."""
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
packed, rect_ids, rect_counts = saved[7], saved[14], saved[15].long()
V, N = packed.shape[0], packed.shape[1]
R = rect_counts.numel()
cap = rect_ids.numel() // R
T = R // (8 * V)
tiles_x = 32
ids = rect_ids.view(R, cap)
tot = dict(entries=0, steps_lr=0, steps_tb=0, any_hit=0, lane_hits=0, both_lr=0, both_tb=0, steps_q=0)
lane = torch.arange(32, device=dev)
lx, ly = (lane & 7).float(), (lane >> 3).float()
left = (lane & 7) < 4
top = (lane >> 3) < 2
CH = 4096
for r0 in range(0, R, CH):
    cnt = rect_counts[r0:r0 + CH]
    m = int(cnt.max())
    if m == 0:
        continue
    pos = torch.arange(m, device=dev)[None, :]
    valid = pos < cnt[:, None]                                   # (ch, m)
    rid = torch.arange(r0, r0 + cnt.numel(), device=dev)
    v = (rid // 8) // T
    tile = (rid // 8) % T
    wt = rid % 8
    x0 = ((tile % tiles_x) * 16 + (wt & 1) * 8).float() + 0.5
    y0 = ((tile // tiles_x) * 16 + (wt >> 1) * 4).float() + 0.5
    g = ids[r0:r0 + cnt.numel(), :m].long().clamp_(0, N - 1)
    rec = packed[v[:, None].expand_as(g)[valid], g[valid]]       # (E, 12)
    rr = torch.nonzero(valid)[:, 0]
    px = x0[rr][:, None] + lx[None, :]
    py = y0[rr][:, None] + ly[None, :]
    dx, dy = rec[:, 0:1] - px, rec[:, 1:2] - py
    pw = dx * (rec[:, 5:6] * dy + rec[:, 4:5] * dx) + rec[:, 6:7] * dy * dy
    alpha = torch.exp2(pw + rec[:, 7:8]).clamp(max=0.999)
    hit = (pw <= 0) & (alpha >= 1.0 / 255.0)                     # (E, 32)
    hl, hr = hit[:, left].any(1), hit[:, ~left].any(1)
    ht, hb = hit[:, top].any(1), hit[:, ~top].any(1)
    quad = torch.stack([hit[:, left & top].any(1), hit[:, left & ~top].any(1), hit[:, ~left & top].any(1),
                        hit[:, ~left & ~top].any(1)], 1)
    nrect = cnt.numel()
    def per_rect(mask):
        return torch.zeros(nrect, device=dev, dtype=torch.long).index_add_(0, rr, mask.long())
    tot["entries"] += int(valid.sum())
    tot["any_hit"] += int((hl | hr).sum())
    tot["lane_hits"] += int(hit.sum())
    tot["both_lr"] += int((hl & hr).sum())
    tot["both_tb"] += int((ht & hb).sum())
    tot["steps_lr"] += int(torch.maximum(per_rect(hl), per_rect(hr)).sum())
    tot["steps_tb"] += int(torch.maximum(per_rect(ht), per_rect(hb)).sum())
    tot["steps_q"] += int(torch.stack([per_rect(quad[:, k]) for k in range(4)], 1).max(1).values.sum())
e = tot["entries"]
print(tot)
print(f"entries {e}; with a hit {tot['any_hit'] / e:.3f}; lane efficiency {tot['lane_hits'] / (32 * e):.3f}")
print(f"left/right halves: both hit {tot['both_lr'] / e:.3f}; lockstep steps {tot['steps_lr']} = {tot['steps_lr'] / e:.3f} of entries")
print(f"top/bottom halves: both hit {tot['both_tb'] / e:.3f}; lockstep steps {tot['steps_tb']} = {tot['steps_tb'] / e:.3f} of entries")
print(f"four 4x2 quarters: lockstep steps {tot['steps_q']} = {tot['steps_q'] / e:.3f} of entries")
