#!/usr/bin/env python
"""Where the end-to-end step (bench.py's `e2e`) spends its wall time beyond the resident step: host
time stamps at the phase boundaries of the e2e loop on the benchmark workload, plus the same loop with
the upload switched off (scene already resident) to separate launch latency from copy interference."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from splatformer_b200 import gs_utils, synth, loader

dev = torch.device("cuda:0")
WORKLOAD = "objaverse_100k_sh3_512"
scenes = []
for s in range(2):
    gs, cams, gt = synth.make_config(WORKLOAD, s)
    gt = (gt * 255.0).round().to(torch.uint8)
    scenes.append((gs, dict(cams), gt))
host = [({k: v.pin_memory() for k, v in gs.items()}, {k: v.pin_memory() for k, v in cams.items()}, gt.pin_memory())
        for gs, cams, gt in scenes]
resident = [({k: v.to(dev) for k, v in gs.items()}, {k: v.to(dev) for k, v in cams.items()}, gt.to(dev))
            for gs, cams, gt in scenes]
V = scenes[0][2].shape[0]
feeder = loader.ScenePrefetcher(dev)
NAMES = ["get", "fwd_host", "bwd_host", "submit", "item"]


def run(steps, upload, stamps):
    ticket = feeder.submit(host[0]) if upload else None
    for i in range(steps):
        t = [time.perf_counter()]
        if upload:
            gs_d, cams, gt_d = feeder.get(ticket)
        else:
            gs_d, cams, gt_d = resident[i % 2]
        raw = {k: v.detach().requires_grad_() for k, v in gs_d.items()}
        t.append(time.perf_counter())
        l1, _ = gs_utils.render_l1_loss(raw, cams, gt_d)
        loss = l1.sum() * (1.0 / V)
        t.append(time.perf_counter())
        loss.backward()
        t.append(time.perf_counter())
        if upload and i + 1 < steps:
            ticket = feeder.submit(host[(i + 1) % 2])
        t.append(time.perf_counter())
        loss.item()
        t.append(time.perf_counter())
        stamps.append(t)


for upload in (True, False):
    run(5, upload, [])
    torch.cuda.synchronize()
    stamps = []
    t0 = time.perf_counter()
    run(40, upload, stamps)
    torch.cuda.synchronize()
    wall = (time.perf_counter() - t0) / 40
    print(f"upload={upload}: wall per step {wall * 1e3:.3f} ms")
    for j, name in enumerate(NAMES):
        d = sorted(s[j + 1] - s[j] for s in stamps)
        print(f"   {name:9s} median {d[len(d) // 2] * 1e6:8.1f} us   min {d[0] * 1e6:8.1f} us")
