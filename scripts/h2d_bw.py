#!/usr/bin/env python
"""Host->device copy bandwidth of this box for the e2e scene upload (one pinned buffer of the scene's size)."""
import torch, subprocess
dev = torch.device("cuda:0")
for mb in (48.8, 256.0):
    n = int(mb * 1e6)
    h = torch.empty(n, dtype=torch.uint8).pin_memory()
    d = torch.empty(n, dtype=torch.uint8, device=dev)
    for _ in range(3): d.copy_(h, non_blocking=True)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(20): d.copy_(h, non_blocking=True)
    b.record(); torch.cuda.synchronize()
    ms = a.elapsed_time(b) / 20
    print(f"H2D {mb} MB pinned: {ms:.3f} ms per copy = {n / ms / 1e6:.1f} GB/s")
print(subprocess.run("nvidia-smi --query-gpu=pcie.link.gen.current,pcie.link.gen.max,pcie.link.width.current --format=csv",
                     shell=True, capture_output=True, text=True).stdout)
