#!/usr/bin/env python
"""Bucket statistics of tile_sort.cu's per-tile bucket sort on the bench workload (run on the GPU box):
how full the depth buckets get under the kernel's linear map of the depth bits, how many tiles fall back
to the bitonic sort, and the insertion-sort work per warp (a warp pays for its fullest bucket)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from splatformer_b200 import gs_utils, render, synth

dev = torch.device("cuda:0")
gs, cams, gt = synth.make_config("objaverse_100k_sh3_512", 0)
raw = {k: v.to(dev).requires_grad_() for k, v in gs.items()}
cams = {k: v.to(dev) for k, v in cams.items()}
l1, _ = gs_utils.render_l1_loss(raw, cams, gt.to(dev))
fn = l1.grad_fn
while not hasattr(fn, "saved_tensors"):
    fn = fn.next_functions[0][0]
saved = fn.saved_tensors
gids, bins = saved[8], saved[9]
viewmats, _ = gs_utils.viewmats_from_c2w(cams["camera_to_worlds"])
T = bins.shape[0] // viewmats.shape[0]
tot = {"tiles": 0, "bitonic": 0, "entries": 0, "ins_steps": 0.0, "bitonic_entries": 0}
hist = torch.zeros(64, dtype=torch.long, device=dev)
for v in range(0, viewmats.shape[0], 8):
    depth = (raw["means"].detach() @ viewmats[v, :3, :3].T + viewmats[v, :3, 3])[:, 2].contiguous()
    b = bins[v * T:(v + 1) * T].long()
    lens = b[:, 1] - b[:, 0]
    occ = (lens > 32).nonzero().flatten()
    tile_of = torch.repeat_interleave(occ, lens[occ])
    start = torch.repeat_interleave(b[occ, 0], lens[occ])
    pos = torch.arange(tile_of.numel(), device=dev) - torch.repeat_interleave(
        torch.cumsum(lens[occ], 0) - lens[occ], lens[occ])
    g = gids[(start + pos)].long()
    bits = depth[g].view(torch.int32).long()
    lo = torch.full((T,), 2 ** 40, device=dev).scatter_reduce(0, tile_of, bits, "amin")
    hi = torch.zeros(T, dtype=torch.long, device=dev).scatter_reduce(0, tile_of, bits, "amax")
    nb = torch.pow(2, torch.ceil(torch.log2(lens.clamp(min=64).float()))).long()
    scale = nb.float() / ((hi - lo).float() + 1.0)
    bk = torch.minimum(((bits - lo[tile_of]).float() * scale[tile_of]).long(), nb[tile_of] - 1)
    cnt = torch.bincount(tile_of * 8192 + bk, minlength=T * 8192).view(T, 8192)
    mx = cnt.max(dim=1).values
    over = mx > 24
    hist += torch.bincount(cnt[occ][cnt[occ] > 0].clamp(max=63), minlength=64)
    # insertion work: thread t of 256 owns buckets t, t+256, ...; a warp pays the max over its 32 lanes
    work = (cnt.float() * (cnt.float() - 1) / 4).view(T, -1, 8, 32).max(dim=3).values.sum(dim=(1, 2))
    tot["tiles"] += occ.numel(); tot["bitonic"] += int(over[occ].sum()); tot["entries"] += int(lens[occ].sum())
    tot["bitonic_entries"] += int(lens[occ][over[occ]].sum())
    tot["ins_steps"] += float(work[occ][~over[occ]].sum())
print(tot)
print("entries in tiles that fall back to bitonic: %.1f %%" % (100.0 * tot["bitonic_entries"] / tot["entries"]))
print("insertion inner steps per entry (warp-level, non-bitonic tiles): %.2f" %
      (tot["ins_steps"] / max(1, tot["entries"] - tot["bitonic_entries"])))
h = hist.tolist()
print("bucket-size histogram (non-empty buckets):", {i: h[i] for i in range(1, 64) if h[i]})
