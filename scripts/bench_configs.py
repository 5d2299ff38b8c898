#!/usr/bin/env python
"""Throughput of the other BASELINE.json shapes (configs 4 and 5; config 2 is bench.py's line) on one
B200: Mpix/s, Gaussians/s and the step's algorithmic HBM fraction, CUDA-event timed.
    python scripts/bench_configs.py [> profiles/rX_configs.json]"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from splatformer_b200 import gs_utils, render, synth

dev = torch.device("cuda:0")
try:
    PEAK = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    PEAK = 6650.0


def timed(fn, warm, steps):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(steps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / steps


def run(name, views, backward, warm, steps):
    c = synth.CONFIGS[name]
    gs, cams, gt = synth.make_config(name, 0, num_views=views)
    raw = {k: v.to(dev).requires_grad_(backward) for k, v in gs.items()}
    cams = {k: v.to(dev) for k, v in cams.items()}
    gt = (gt * 255).round().to(torch.uint8).to(dev)

    def step():
        if backward:
            for p in raw.values():
                p.grad = None
            l1, _ = gs_utils.render_l1_loss(raw, cams, gt)
            l1.mean().backward()
        else:
            with torch.no_grad():
                gs_utils.render_scene(raw, **cams)

    ms = timed(step, warm, steps)
    n, k, P = c["n"], (c["deg"] + 1) ** 2, c["width"] * c["height"]
    T = ((c["width"] + 15) // 16) * ((c["height"] + 15) // 16)
    I = render.last_num_intersects() / views
    b_fwd = (12 * k + 188) * n + 48 * I + 20 * P + 8 * T
    b_bwd = (12 * k + 248) * n + 4 * I + 24 * P
    alg = (b_fwd + (b_bwd if backward else 0)) * views
    out = {"config": name, "gaussians": n, "sh_degree": c["deg"], "views": views, "image": [c["height"], c["width"]],
           "pass": "fwd+bwd" if backward else "fwd", "ms_per_step": ms, "mpix_per_s": views * P / ms / 1e3,
           "gaussians_per_s": views * n / (ms * 1e-3), "intersections_per_view": I,
           "alg_bytes_per_step": alg, "hbm_gbs": alg / (ms * 1e-3) / 1e9, "hbm_frac": alg / (ms * 1e-3) / 1e9 / PEAK,
           "steps": steps, "warmup": warm}
    print(json.dumps(out), flush=True)
    del raw, gt
    torch.cuda.empty_cache()


if __name__ == "__main__":
    run("objaverse_100k_sh3_512", 32, True, 5, 30)
    run("real_1m_sh3_1080p", 8, False, 3, 10)
    run("stress_6m_4k", 1, True, 4, 5)     # the caching allocator needs 3 calls to settle (scripts/stress_probe.py)
