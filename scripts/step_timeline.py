#!/usr/bin/env python
"""Where the device idles inside the fused step: start / end of every stage (CUDA events) of a few consecutive
steady-state steps, relative to the first one, next to the host wall-clock per step."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from splatformer_b200 import gs_utils, render, synth
dev = torch.device("cuda:0"); torch.cuda.set_device(dev)
gs, cams, gt = synth.make_config("objaverse_100k_sh3_512", 0)
gt = (gt * 255).round().to(torch.uint8).to(dev)
cams = {k: v.to(dev) for k, v in cams.items()}
raw = {k: v.to(dev).requires_grad_() for k, v in gs.items()}
def step():
    for p in raw.values(): p.grad = None
    l1, _ = gs_utils.render_l1_loss(raw, cams, gt)
    (l1.sum() / 32).backward()
for _ in range(10): step()
torch.cuda.synchronize()
N = int(os.environ.get("STEPS", 50))
t0 = time.perf_counter()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(N): step()
b.record()
host = (time.perf_counter() - t0) / N * 1e3
torch.cuda.synchronize()
print(f"no timers: device {a.elapsed_time(b)/N:.3f} ms/step, host loop {host:.3f} ms/step (enqueue only); "
      f"reserved {torch.cuda.memory_reserved()/2**30:.2f} GiB, peak allocated {torch.cuda.max_memory_allocated()/2**30:.2f} GiB")
with render.profile_stages() as t:
    base = torch.cuda.Event(enable_timing=True); base.record()
    marks = []
    for i in range(6):
        h0 = time.perf_counter(); step(); marks.append((time.perf_counter() - h0) * 1e3)
    torch.cuda.synchronize()
    prev_end = 0.0
    for name, ea, eb in t.events[:14]:
        s, e = base.elapsed_time(ea), base.elapsed_time(eb)
        print(f"{name:16s} start {s:8.3f} end {e:8.3f}  dur {e-s:6.3f}  gap before {s-prev_end:6.3f}")
        prev_end = e
    print("host ms per step:", [round(x, 3) for x in marks])
