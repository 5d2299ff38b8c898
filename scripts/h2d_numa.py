#!/usr/bin/env python
"""Simultaneous host->device upload bandwidth of all ranks, with the rank's pinned buffers allocated (a) wherever the
process happens to run and (b) after binding the process to the CPUs next to its GPU (loader.bind_host_to_gpu).
torchrun --nproc-per-node N scripts/h2d_numa.py"""
import json
import os
import subprocess
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from splatformer_b200 import loader  # noqa: E402

rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
local = int(os.environ.get("LOCAL_RANK", 0))
dev = torch.device("cuda", local)
torch.cuda.set_device(dev)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)


def measure(tag):
    n = int(48.8e6)
    host = [torch.empty(n, dtype=torch.uint8).pin_memory() for _ in range(2)]
    for h in host:
        h.fill_(1)                                   # first touch
    d = torch.empty(n, dtype=torch.uint8, device=dev)
    for i in range(3):
        d.copy_(host[i % 2], non_blocking=True)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(40):
        d.copy_(host[i % 2], non_blocking=True)
    b.record()
    torch.cuda.synchronize()
    gbs = n * 40 / (a.elapsed_time(b) * 1e-3) / 1e9
    t = torch.tensor([gbs], device=dev, dtype=torch.float64)
    if world > 1:
        lst = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(lst, t)
        vals = [float(x) for x in lst]
    else:
        vals = [gbs]
    if rank == 0:
        print(json.dumps({"mode": tag, "gbs_per_rank": [round(v, 1) for v in vals], "sum": round(sum(vals), 1)}), flush=True)


measure("unbound")
info = loader.bind_host_to_gpu(local)
infos = [None] * world
if world > 1:
    dist.all_gather_object(infos, info)
else:
    infos = [info]
if rank == 0:
    print(json.dumps({"bind": infos}), flush=True)
measure("bound")
if rank == 0:
    print(subprocess.run("nvidia-smi topo -m | head -14; lscpu | grep -i 'numa\\|socket\\|^CPU(s)'", shell=True,
                         capture_output=True, text=True).stdout)
