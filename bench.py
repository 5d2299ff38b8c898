#!/usr/bin/env python
"""bench.py -- forward+backward splat-render throughput on B200 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A *step* is one pass of the hot path over one scene: forward + backward of the render loss
`mean(abs(rgb - gt))` (/root/reference/train.py:273-275,289) over all V views of a
100k-Gaussian / SH3 / 512x512 scene (BASELINE.json configs[1]), through the public call
`splatformer_b200.gs_utils.render_l1_loss` = `rasterize_gaussians_to_multiimgs`
(/root/reference/utils/gs_utils.py:20) + the L1 loop of train.py:274-275 in one autograd node.

  value   whole-job Mpix/s with the scene, cameras and targets already resident in HBM
          (CUDA events, barrier + synchronize on both sides, max over ranks)
  e2e     the same metric with HOST (pinned) inputs: every step copies the raw Gaussian
          parameters, cameras and target images host->device, renders fwd+bwd through the
          same public call and reads the loss back
  roofline   dominant kernel: algorithmic bytes per launch / its mean CUDA-event
             duration inside the timed region, against MEASURED_PEAKS.json
  cpu_baseline   the oracle (pure-PyTorch port of the reference path) on the host cores,
                 on a bounded sample (a few views of the same scene)

N > 1 (torchrun, one rank per GPU): scene-sharded as SplatFormer's DDP does (one scene per GPU
per step, /root/reference/dataset/Loader.py:15-17): no data-path collective, weak scaling.
`--shard view` splits every scene's views over the ranks instead and all-reduces the packed
per-Gaussian gradients over NCCL (strong scaling per scene).

`--impl reference` times the reference's CPU implementation of the path (the oracle port: gsplat
v0.1.11 is not on disk) on rank 0's host cores.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402

METRIC = "Mpix/s fwd+bwd splat render (100k Gauss, SH3, 512^2)"
UNIT = "Mpix/s"
WORKLOAD = "objaverse_100k_sh3_512"


def _env_int(name: str, default: int) -> int:
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


def stage_bytes_per_view(n: int, k: int, isect: float, pix: int, tiles: int) -> dict:
    """Algorithmic (compulsory) bytes per view and stage, SURVEY.md 8(d)."""
    return {
        "project_sh_fwd": (12 * k + 24) * n + 100 * n,
        "cumsum": 8 * n,
        "map_isect": 20 * n + 12 * isect,
        "sort": 24 * isect,
        "bin_edges": 8 * isect + 8 * tiles,
        "raster_fwd": 4 * isect + 36 * n + 20 * pix,
        "raster_bwd": 4 * isect + 72 * n + 24 * pix,
        "project_sh_bwd": 152 * n + (12 * k + 24) * n,
    }


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled while the timed regions run."""
    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,utilization.gpu,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, device_index: int):
        self.proc = None
        self.path = None
        try:
            uuid = str(torch.cuda.get_device_properties(device_index).uuid)
            ident = uuid if uuid.startswith("GPU-") else "GPU-" + uuid
        except Exception:
            ident = str(device_index)
        fd, self.path = tempfile.mkstemp(prefix="clocks_", suffix=".csv")
        os.close(fd)
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={ident}", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits",
                 "-lms", "50"], stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except OSError:
            self.proc = None

    def stop(self) -> dict:
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, smax, reasons, busy = [], [], set(), []
        try:
            for line in open(self.path):
                f = [x.strip() for x in line.split(",")]
                if len(f) < 8:
                    continue
                try:
                    clk, mx = float(f[0]), float(f[1])
                except ValueError:
                    continue
                try:
                    util = float(f[3])
                except ValueError:
                    util = 100.0
                sm.append(clk)
                smax.append(mx)
                if util > 0:
                    busy.append(clk)
                for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"),
                                     f[4:8]):
                    if val.lower().startswith("active"):
                        reasons.add(name)
        finally:
            try:
                os.unlink(self.path)
            except OSError:
                pass
        use = busy or sm
        if use:
            out.update(sm_mhz=statistics.median(use), sm_max_mhz=max(smax), reasons=sorted(reasons),
                       samples=len(use))
        return out


def run_reference(args, rank: int) -> None:
    """Reference arm: the reference's CPU implementation of the path (oracle port), rank 0 only."""
    if rank != 0:
        return
    from oracle import ref_torch as R
    from splatformer_b200 import synth
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    views = max(1, args.ref_views)
    gs, cams, gt = synth.make_config(WORKLOAD, 0)
    c = synth.CONFIGS[WORKLOAD]
    sample_cams = dict(cams)
    # evenly spaced subset of the 32 views = the bounded sample
    idx = torch.linspace(0, c["views"] - 1, views).round().long()
    sample_cams["camera_to_worlds"] = cams["camera_to_worlds"][idx]
    gt_s = gt[idx]

    def step():
        raw = {k: v.clone().requires_grad_() for k, v in gs.items()}
        rgbs, _ = R.render_views(raw, sample_cams)
        loss = sum((rgb - g).abs().mean() for rgb, g in zip(rgbs, gt_s)) / views
        loss.backward()
        return float(loss.detach())

    for _ in range(args.warmup_ref):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps_ref):
        step()
    dt = (time.perf_counter() - t0) / args.steps_ref
    pix = views * c["height"] * c["width"]
    val = pix / dt / 1e6
    sample = f"{views} of {c['views']} views of the scene per step (fwd+bwd), {args.steps_ref} steps"
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps_ref, "warmup": args.warmup_ref, "ms_per_step": dt * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "gaussians": c["n"], "sh_degree": c["deg"], "views": c["views"],
                   "image": [c["height"], c["width"]], "sample_views_per_step": views},
        "gaussians_per_s": views * c["n"] / dt,
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def cpu_baseline(views: int) -> dict:
    from oracle import ref_torch as R
    from splatformer_b200 import synth
    cores = os.cpu_count() or 1
    prev = torch.get_num_threads()
    torch.set_num_threads(cores)
    try:
        gs, cams, gt = synth.make_config(WORKLOAD, 0)
        c = synth.CONFIGS[WORKLOAD]
        idx = torch.linspace(0, c["views"] - 1, views).round().long()
        sample_cams = dict(cams)
        sample_cams["camera_to_worlds"] = cams["camera_to_worlds"][idx]
        raw = {k: v.clone().requires_grad_() for k, v in gs.items()}
        t0 = time.perf_counter()
        rgbs, _ = R.render_views(raw, sample_cams)
        loss = sum((rgb - g).abs().mean() for rgb, g in zip(rgbs, gt[idx])) / views
        loss.backward()
        dt = time.perf_counter() - t0
    finally:
        torch.set_num_threads(prev)
    return {"value": views * c["height"] * c["width"] / dt / 1e6, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"{views} of {c['views']} views of the same scene, fwd+bwd once, oracle/ref_torch.py",
            "seconds": dt}


def run_config3(args, steps, rank, world, dev, timed, V, H, W, n):
    """8 scenes x 32 views per step on `world` GPUs.
    scene-sharded: rank r renders scenes r, r + G, ... (SplatFormer's DDP layout, no collective);
    view-sharded : every scene's views are split over the ranks, Gaussians replicated; the per-Gaussian
                   gradients of scene i are all-reduced (NCCL, in place on the backward's flat buffer) while
                   scene i + 1 renders; the step ends when every reduce has landed."""
    import torch.distributed as dist
    from splatformer_b200 import dist as sbdist
    from splatformer_b200 import gs_utils, render, synth
    from splatformer_b200.graphs import capture_step
    S = args.config3_scenes
    inv_views = 1.0 / V
    pix_step = S * V * H * W
    out = {"workload": f"{S} scenes x {V} views of {n} Gaussians per step (BASELINE config 3)", "steps": steps,
           "unit": UNIT}

    def load(sid, lo, hi):
        gs, cams, gt = synth.make_config(WORKLOAD, 100 + sid, num_views=V)
        gt = (gt * 255.0).round().to(torch.uint8)
        cams = dict(cams)
        cams["camera_to_worlds"] = cams["camera_to_worlds"][lo:hi].contiguous()
        return ({k: v.to(dev).requires_grad_() for k, v in gs.items()}, {k: v.to(dev) for k, v in cams.items()},
                gt[lo:hi].contiguous().to(dev))

    # ---- scene-sharded ----
    mine = sbdist.scene_slice(S, rank, world)
    scenes = [load(sid, 0, V) for sid in mine]
    plan = render.RenderPlan()

    def step_scene(i):
        for raw, cams, gt in scenes:
            for p in raw.values():
                p.grad = None
            l1, _ = gs_utils.render_l1_loss(raw, cams, gt, plan=plan)
            (l1.sum() * inv_views).backward()

    for i in range(2):
        step_scene(i)
    ms, _, _, _ = timed(step_scene, steps)
    out["scene_sharded"] = {"value": pix_step / (ms / steps * 1e-3) / 1e6, "ms_per_step": ms / steps,
                            "scenes_per_rank": len(mine), "collective": "none"}
    if args.graphs:
        g, why = capture_step(step_scene, plan, len(mine))
        if g is not None:
            ms_g, _, _, _ = timed(lambda i: g.replay(), steps)
            ok = plan.check()
            out["scene_sharded"]["graph_replay"] = {"value": pix_step / (ms_g / steps * 1e-3) / 1e6,
                                                    "ms_per_step": ms_g / steps, "no_overflow": ok}
        else:
            out["scene_sharded"]["graph_replay"] = {"unavailable": why}
        del g
    del scenes
    if world == 1:
        out["view_sharded"] = None          # one GPU: the two layouts are the same loop
        return out

    # ---- view-sharded ----
    lo, hi = sbdist.view_slice(V, rank, world)
    scenes = [load(sid, lo, hi) for sid in range(S)]
    plan_v = render.RenderPlan()

    def step_view(sync, overlap):
        def fn(i):
            for raw, cams, gt in scenes:
                for p in raw.values():
                    p.grad = None
                routed = sbdist.sync_gaussian_grads(raw, overlap=overlap) if sync else raw
                l1, _ = gs_utils.render_l1_loss(routed, cams, gt, plan=plan_v)
                (l1.sum() * inv_views).backward()
            sbdist.wait_pending()
        return fn

    for i in range(2):
        step_view(True, True)(i)
    ms_ov, _, _, _ = timed(step_view(True, True), steps)
    ms_blk, _, _, _ = timed(step_view(True, False), steps)
    ms_none, _, _, _ = timed(step_view(False, False), steps)
    # the exchange on its own: one all-reduce of a scene's flat gradient buffer
    raw0 = scenes[0][0]
    nfloat = sum(p.numel() for p in raw0.values())
    buf = torch.zeros(nfloat, device=dev)
    for _ in range(3):
        dist.all_reduce(buf)

    def ar(i):
        dist.all_reduce(buf)
    ms_ar, _, _, _ = timed(ar, 20)
    ar_ms = ms_ar / 20
    hidden = (ms_blk - ms_ov) / steps
    out["view_sharded"] = {
        "value": pix_step / (ms_ov / steps * 1e-3) / 1e6, "ms_per_step": ms_ov / steps,
        "views_per_rank": hi - lo, "allreduces_per_step": S, "allreduce_bytes": nfloat * 4, "allreduce_ms": ar_ms,
        "ms_per_step_blocking_allreduce": ms_blk / steps, "ms_per_step_without_allreduce": ms_none / steps,
        "overlap_frac": max(0.0, min(1.0, 1.0 - (ms_ov - ms_none) / steps / max(S * ar_ms, 1e-9))),
        "collective": "NCCL all-reduce(sum) of the flat (N, 11+3K) gradient buffer, in place, async under the next scene",
        "nccl_ranks": dist.get_world_size(), "hidden_ms_per_step_vs_blocking": hidden,
    }
    # correctness on the device: the all-reduced gradients of scene 0 == the gradients of all its views on one GPU
    raw, cams, gt = scenes[0]
    for p in raw.values():
        p.grad = None
    routed = sbdist.sync_gaussian_grads(raw, overlap=False)
    l1, _ = gs_utils.render_l1_loss(routed, cams, gt, plan=plan_v)
    (l1.sum() * inv_views).backward()
    full_raw, full_cams, full_gt = load(0, 0, V)
    l1f, _ = gs_utils.render_l1_loss(full_raw, full_cams, full_gt)
    (l1f.sum() * inv_views).backward()
    per = {}
    for k in raw:
        a, b = raw[k].grad.double(), full_raw[k].grad.double()
        finite = bool(torch.isfinite(a).all()) and bool(torch.isfinite(b).all())
        per[k] = float((a - b).norm() / b.norm().clamp_min(1e-30)) if finite else float("inf")
    worst = max(per.values())
    t = torch.tensor([worst], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    worst = float(t.item())
    out["view_sharded"]["grad_rel_l2_vs_single_gpu"] = worst
    out["view_sharded"]["grad_rel_l2_per_tensor"] = per
    if not (worst <= 1e-4):          # also true for NaN / inf
        raise SystemExit(f"view-sharded gradients differ from the single-GPU ones: rel L2 {worst:.3e} ({per})")
    # (the autograd graphs of the check above must be gone before anything is captured: an AccumulateGrad node kept
    # alive from an eager pass on the default stream makes the captured backward synchronise with that stream, which
    # invalidates the capture)
    del routed, l1, l1f
    if args.graphs:
        # One graph per scene holds the rank's render + backward; the all-reduce of the scene's flat gradient buffer
        # is issued between replays (NCCL's stream, under the next scene's replay).  At 4-8 ranks the eager loop is
        # host-bound (a pass renders 4-8 views: 0.3-0.5 ms of kernels behind ~0.6 ms of Python and launches).
        plans_g = [render.RenderPlan() for _ in range(S)]

        def scene_fn(i):
            def fn(_):
                raw_i, cams_i, gt_i = scenes[i]
                for p in raw_i.values():
                    p.grad = None
                l1_i, _ = gs_utils.render_l1_loss(raw_i, cams_i, gt_i, plan=plans_g[i])
                (l1_i.sum() * inv_views).backward()
            return fn
        graphs_v, why_v = [], None
        for i in range(S):
            fn = scene_fn(i)
            for _ in range(2):
                fn(0)
            g_i, why_v = capture_step(fn, plans_g[i], 1)
            if g_i is None:
                break
            graphs_v.append(g_i)
        ok_all = torch.tensor([1 if len(graphs_v) == S else 0], device=dev)
        dist.all_reduce(ok_all, op=dist.ReduceOp.MIN)
        if int(ok_all.item()) == 1:
            for g_i in graphs_v:
                g_i.replay()
            flats = [sbdist.flat_grad_buffer(scenes[i][0]) for i in range(S)]
            if all(f is not None for f in flats):
                def step_view_graph(_):
                    for i in range(S):
                        graphs_v[i].replay()
                        sbdist.allreduce_flat(flats[i], overlap=True)
                    sbdist.wait_pending()
                for i in range(2):
                    step_view_graph(i)
                ms_vg, _, _, _ = timed(step_view_graph, steps)
                ok = all(p.check() for p in plans_g)
                # the graph path's gradients of scene 0 against all its views on one GPU
                per_g = {}
                for k in raw:
                    a, b = scenes[0][0][k].grad.double(), full_raw[k].grad.double()
                    fin = bool(torch.isfinite(a).all())
                    per_g[k] = float((a - b).norm() / b.norm().clamp_min(1e-30)) if fin else float("inf")
                tg = torch.tensor([max(per_g.values())], device=dev, dtype=torch.float64)
                dist.all_reduce(tg, op=dist.ReduceOp.MAX)
                if not (float(tg.item()) <= 1e-4):
                    raise SystemExit(f"view-sharded graph replay: gradients differ from the single-GPU ones: {per_g}")
                out["view_sharded"]["graph_replay"] = {
                    "value": pix_step / (ms_vg / steps * 1e-3) / 1e6, "ms_per_step": ms_vg / steps, "no_overflow": ok,
                    "grad_rel_l2_vs_single_gpu": float(tg.item()),
                    "how": "one CUDA graph per scene (render + backward of the rank's views); the in-place NCCL "
                           "all-reduce of the scene's flat gradient buffer is issued between replays, on NCCL's "
                           "stream under the next scene's replay (dist.allreduce_flat), joined at the end of the step"}
            else:
                out["view_sharded"]["graph_replay"] = {"unavailable": "gradients are not one flat buffer"}
        else:
            out["view_sharded"]["graph_replay"] = {"unavailable": why_v or "capture failed on another rank"}
        del graphs_v
        # (An NCCL all-reduce INSIDE the capture is not attempted: with the async work handle it invalidates the
        # capture -- cudaErrorStreamCaptureInvalidated -- and the blocking form hung a 2-rank capture on this stack,
        # torch 2.11 / NCCL 2.28.9, until the box's limit.)
    return out


def run_other_configs(dev):
    """BASELINE configs 4 (1M Gaussians, 8 views 1080p, forward only) and 5 (6M Gaussians, 2 views 4K, fwd+bwd)."""
    from splatformer_b200 import gs_utils, render, synth
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    res = []
    for name, views, backward, steps in (("real_1m_sh3_1080p", 8, False, 5), ("stress_6m_4k", 2, True, 4)):
        c = synth.CONFIGS[name]
        gs, cams, gt = synth.make_config(name, 0, num_views=views)
        gt = (gt * 255.0).round().to(torch.uint8).to(dev)
        cams = {k: v.to(dev) for k, v in cams.items()}
        raw = {k: v.to(dev).requires_grad_(backward) for k, v in gs.items()}
        plan = render.RenderPlan()

        def step():
            if backward:
                for p in raw.values():
                    p.grad = None
                l1, _ = gs_utils.render_l1_loss(raw, cams, gt, plan=plan)
                (l1.sum() / views).backward()
            else:
                with torch.no_grad():
                    gs_utils.render_scene(raw, **cams, plan=plan)

        for _ in range(3):
            step()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(steps):
            step()
        b.record()
        torch.cuda.synchronize()
        ms = a.elapsed_time(b) / steps
        isect = render.last_num_intersects() / views
        k = (c["deg"] + 1) ** 2
        tiles = ((c["width"] + 15) // 16) * ((c["height"] + 15) // 16)
        per_view = stage_bytes_per_view(c["n"], k, isect, c["height"] * c["width"], tiles)
        fwd_keys = ("project_sh_fwd", "cumsum", "map_isect", "sort", "bin_edges", "raster_fwd")
        b_alg = sum(v for kk, v in per_view.items() if backward or kk in fwd_keys) * views
        res.append({"workload": name, "gaussians": c["n"], "views": views, "image": [c["height"], c["width"]],
                    "pass": "fwd+bwd" if backward else "fwd", "steps": steps, "ms_per_step": ms,
                    "mpix_per_s": views * c["height"] * c["width"] / (ms * 1e-3) / 1e6,
                    "intersections_per_view": isect, "alg_bytes_per_step": b_alg,
                    "hbm_frac": b_alg / (ms * 1e-3) / 1e9 / peak, "binning": "global sort (hand-written one-sweep radix)"})
        del raw, gs, gt
        torch.cuda.empty_cache()
    return res


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--shard", default="scene", choices=["scene", "view"])
    ap.add_argument("--views", type=int, default=None, help="override the views per scene (default: config's 32)")
    ap.add_argument("--cpu-views", type=int, default=3, help="views of the CPU-baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--e2e-packed", action="store_true",
                    help="e2e loop uploads the scene as one packed pinned buffer (loader.pack_scene) instead of "
                         "tensor by tensor; measured equal within box variance (DESIGN.md section 5)")
    ap.add_argument("--e2e-submit", choices=["before_fwd", "after_fwd", "after_bwd"], default="after_bwd",
                    help="where the e2e loop enqueues the next scene's host->device upload")
    ap.add_argument("--wrapper-steps", type=int, default=5,
                    help="timed steps of the unmodified per-view wrapper path (0 = skip)")
    ap.add_argument("--config3-steps", type=int, default=None,
                    help="timed steps of the 8-scene train-step shape (default: steps // 5, at least 4; 0 = skip)")
    ap.add_argument("--config3-scenes", type=int, default=8)
    ap.add_argument("--graphs", type=int, default=1,
                    help="1: also time config 3's step replayed from one captured CUDA graph; 0: skip")
    ap.add_argument("--headline", choices=["graph", "eager"], default="graph",
                    help="graph: value / e2e are the step replayed from a captured CUDA graph when the capture "
                         "succeeds (the eager loop is reported under `eager`); eager: the uncaptured calls")
    ap.add_argument("--other-configs", type=int, default=1,
                    help="1: also time BASELINE configs 4 and 5 (single GPU) after the headline; 0: skip")
    ap.add_argument("--ref-views", type=int, default=2)
    ap.add_argument("--steps-ref", type=int, default=None)
    ap.add_argument("--warmup-ref", type=int, default=None)
    args = ap.parse_args()

    rank = _env_int("RANK", 0)
    local_rank = _env_int("LOCAL_RANK", 0)
    world = _env_int("WORLD_SIZE", 1)

    if args.impl == "reference":
        # the CPU path costs ~5 s per view: bound the run irrespective of the GPU arm's K / W
        args.steps_ref = args.steps_ref if args.steps_ref is not None else max(1, min(args.steps, 3))
        args.warmup_ref = args.warmup_ref if args.warmup_ref is not None else min(args.warmup, 1)
        run_reference(args, rank)
        return

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the renderer has no CPU path "
                         "(use --impl reference for the CPU reference arm)")
    if args.warmup < 3:
        args.warmup = 3
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    import torch.distributed as dist
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    from splatformer_b200 import _abi, gs_utils, render, synth
    from splatformer_b200 import dist as sbdist
    _abi.load()

    c = synth.CONFIGS[WORKLOAD]
    n, deg, H, W = c["n"], c["deg"], c["height"], c["width"]
    V = args.views or c["views"]
    K = (deg + 1) ** 2
    tiles = ((W + 15) // 16) * ((H + 15) // 16)

    # ---- inputs: NSCENE distinct scenes per rank, rotated between steps (inputs > L2) --------
    NSCENE = 2
    scenes = []
    view_lo, view_hi = 0, V
    if args.shard == "view" and world > 1:
        view_lo, view_hi = sbdist.view_slice(V, rank, world)
    for s in range(NSCENE):
        sid = (s if args.shard == "view" else rank * NSCENE + s)
        gs, cams, gt = synth.make_config(WORKLOAD, sid, num_views=V)
        gt = (gt * 255.0).round().to(torch.uint8)        # 8-bit target images, as they are on disk
        cams = dict(cams)
        cams["camera_to_worlds"] = cams["camera_to_worlds"][view_lo:view_hi].contiguous()
        gt = gt[view_lo:view_hi].contiguous()
        scenes.append((gs, cams, gt))
    v_local = view_hi - view_lo
    inv_views = 1.0 / V                     # mean over all V views of the scene, also when view-sharded

    host = [({k: v.pin_memory() for k, v in gs.items()}, {k: v.pin_memory() for k, v in cams.items()},
             gt.pin_memory()) for gs, cams, gt in scenes]
    resident = [({k: v.to(dev).requires_grad_() for k, v in gs.items()}, {k: v.to(dev) for k, v in cams.items()},
                 gt.to(dev)) for gs, cams, gt in scenes]
    h2d_bytes = sum(v.numel() * v.element_size() for v in host[0][0].values()) \
        + sum(v.numel() * v.element_size() for v in host[0][1].values()) \
        + host[0][2].numel() * host[0][2].element_size()

    host_scenes = list(host)              # (tensor by tensor; `host` may be re-packed for --e2e-packed below)

    def render_loss(raw, cams, gt):
        # train.py:272-275 in one public call: per-view mean |pred - gt| of all views of the scene
        l1, _ = gs_utils.render_l1_loss(raw, cams, gt)
        return l1.sum() * inv_views

    def sync_grads(raw):
        if args.shard == "view" and world > 1:
            sbdist.allreduce_gaussian_grads(raw)

    def step_resident(i):
        raw, cams, gt = resident[i % NSCENE]
        for p in raw.values():
            p.grad = None
        loss = render_loss(raw, cams, gt)
        loss.backward()
        sync_grads(raw)
        return loss

    # e2e: every step's scene comes from pinned host memory.  splatformer_b200.loader.ScenePrefetcher uploads
    # scene i+1 on a copy stream while scene i renders (every scene is still copied on every step); the
    # upload is submitted while the device is busy with the step's kernels, so its host-side cost hides too.
    from splatformer_b200 import loader as sbloader
    feeder = sbloader.ScenePrefetcher(dev)
    if args.e2e_packed:
        # each host scene lives in ONE pinned buffer (loader.pack_scene, done once where a DataLoader worker
        # would collate it): the per-step upload is a single allocation + copy of the same bytes
        host = [sbloader.pack_scene(h) for h in host]
        h2d_bytes = host[0].blob.numel()          # what is copied per step (tensors + < 3 kB of alignment)
    e2e_feed = {"ticket": None, "steps": 0}

    def step_e2e(i):
        if i == 0:      # the loop's first upload is issued here, inside the timed region, and not overlapped
            e2e_feed["ticket"] = feeder.submit(host[0])
        gs_d, cams, gt_d = feeder.get(e2e_feed["ticket"])
        if args.e2e_submit == "before_fwd" and i + 1 < e2e_feed["steps"]:
            e2e_feed["ticket"] = feeder.submit(host[(i + 1) % NSCENE])
        raw = {k: v.requires_grad_() for k, v in gs_d.items()}
        loss = render_loss(raw, cams, gt_d)
        # (measured: enqueueing the upload before the forward or between forward and backward is 1.5-3 %
        #  slower than after the backward -- the host is the critical resource in those windows)
        if args.e2e_submit == "after_fwd" and i + 1 < e2e_feed["steps"]:
            e2e_feed["ticket"] = feeder.submit(host[(i + 1) % NSCENE])
        loss.backward()
        sync_grads(raw)
        if args.e2e_submit == "after_bwd" and i + 1 < e2e_feed["steps"]:
            e2e_feed["ticket"] = feeder.submit(host[(i + 1) % NSCENE])
        # device->host read of the step's result: the loss goes to pinned memory asynchronously and is read
        # on the host one step later (as a training loop that logs the loss does), so the host never stalls
        # the launch queue; every step's loss is read inside the timed region (e2e_drain)
        slot = e2e_loss_ring[i % len(e2e_loss_ring)]
        slot.copy_(loss.detach().reshape(1), non_blocking=True)
        ev = torch.cuda.Event()
        ev.record()
        e2e_pending.append((ev, slot))
        if len(e2e_pending) > 1:
            ev0, s0 = e2e_pending.pop(0)
            ev0.synchronize()
            e2e_losses.append(float(s0[0]))

    e2e_loss_ring = [torch.zeros(1, dtype=torch.float32).pin_memory() for _ in range(4)]
    e2e_pending, e2e_losses = [], []

    def e2e_drain():
        while e2e_pending:
            ev0, s0 = e2e_pending.pop(0)
            ev0.synchronize()
            e2e_losses.append(float(s0[0]))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, with_stages=False, finish=None):
        barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        launches0 = _abi.LAUNCHES
        stage_ms, stage_calls = {}, {}
        ctx = render.profile_stages() if with_stages else None
        timer = ctx.__enter__() if ctx else None
        a.record()
        for i in range(steps):
            fn(i)
        if finish is not None:
            finish()
        b.record()
        barrier()
        if ctx:
            stage_ms, stage_calls = timer.totals_ms(), timer.counts()
            ctx.__exit__(None, None, None)
        ms = a.elapsed_time(b)
        if world > 1:
            t = torch.tensor([ms], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, _abi.LAUNCHES - launches0, stage_ms, stage_calls

    for i in range(args.warmup):
        step_resident(i)
    sampler = ClockSampler(local_rank) if rank == 0 else None
    ms_total, launches, stage_ms, stage_calls = timed(step_resident, args.steps, with_stages=True)
    isect_per_view = render.last_num_intersects() / max(v_local, 1)

    # ---- the same step replayed from a CUDA graph (splatformer_b200.graphs.capture_step): one graph per resident
    # scene, replayed in the same rotation; what a training loop that captures its render step runs ----
    graph_replay = None
    if args.graphs and not (args.shard == "view" and world > 1):
        from splatformer_b200.graphs import capture_step
        plans = [render.RenderPlan() for _ in range(NSCENE)]

        graphs, why = [], None
        graph_loss = [None] * NSCENE

        def scene_step_keep_loss(sidx):
            def fn(i):
                raw, cams, gt = resident[sidx]
                for p in raw.values():
                    p.grad = None
                l1, _ = gs_utils.render_l1_loss(raw, cams, gt, plan=plans[sidx])
                loss = l1.sum() * inv_views
                loss.backward()
                graph_loss[sidx] = loss.detach()          # static: the graph rewrites it on every replay
            return fn
        for sidx in range(NSCENE):
            fn = scene_step_keep_loss(sidx)
            for _ in range(2):
                fn(0)
            g, why = capture_step(fn, plans[sidx], 1)
            if g is None:
                break
            graphs.append(g)
        if len(graphs) == NSCENE:
            for i in range(3):
                graphs[i % NSCENE].replay()
            ms_g, _, _, _ = timed(lambda i: graphs[i % NSCENE].replay(), args.steps)
            graph_replay = {"ms_per_step": ms_g / args.steps, "no_overflow": all(p.check() for p in plans),
                            "api": "splatformer_b200.graphs.capture_step(step, plan, passes=1): forward + backward "
                                   "of one scene as one CUDA graph per resident scene, same rotation"}
            if not args.no_e2e:
                # end to end through the graphs: every step's scene is copied from pinned host memory INTO the
                # captured input tensors of the graph that will run one step later (copy stream, overlapped with
                # the running replay), the loss is read from pinned memory one step later -- the same bytes and the
                # same reads as the eager e2e loop below
                copy_stream = torch.cuda.Stream()
                uploaded = [torch.cuda.Event() for _ in range(NSCENE)]
                consumed = [torch.cuda.Event() for _ in range(NSCENE)]
                ge = {"steps": 0, "pending": [], "losses": []}
                ring = [torch.zeros(1, dtype=torch.float32).pin_memory() for _ in range(4)]

                def upload(sidx):
                    gs_h, cams_h, gt_h = host_scenes[sidx]
                    raw, cams, gt = resident[sidx]
                    with torch.cuda.stream(copy_stream):
                        copy_stream.wait_event(consumed[sidx])
                        for k, v in gs_h.items():
                            raw[k].data.copy_(v, non_blocking=True)
                        for k, v in cams_h.items():
                            cams[k].copy_(v, non_blocking=True)
                        gt.copy_(gt_h, non_blocking=True)
                        uploaded[sidx].record(copy_stream)

                def step_e2e_graph(i):
                    sidx = i % NSCENE
                    if i == 0:
                        upload(0)
                    cur = torch.cuda.current_stream()
                    cur.wait_event(uploaded[sidx])
                    graphs[sidx].replay()
                    consumed[sidx].record(cur)
                    if i + 1 < ge["steps"]:
                        upload((i + 1) % NSCENE)
                    slot = ring[i % len(ring)]
                    slot.copy_(graph_loss[sidx].reshape(1), non_blocking=True)
                    ev = torch.cuda.Event()
                    ev.record()
                    ge["pending"].append((ev, slot))
                    if len(ge["pending"]) > 1:
                        ev0, s0 = ge["pending"].pop(0)
                        ev0.synchronize()
                        ge["losses"].append(float(s0[0]))

                def drain():
                    while ge["pending"]:
                        ev0, s0 = ge["pending"].pop(0)
                        ev0.synchronize()
                        ge["losses"].append(float(s0[0]))
                for ev in consumed:
                    ev.record()
                ge["steps"] = 3
                for i in range(3):
                    step_e2e_graph(i)
                drain()
                ge["losses"].clear()
                ge["steps"] = args.steps
                ms_ge, _, _, _ = timed(step_e2e_graph, args.steps, finish=drain)
                assert len(ge["losses"]) == args.steps and all(x == x for x in ge["losses"])
                graph_replay["e2e_ms_per_step"] = ms_ge / args.steps
                graph_replay["no_overflow"] = graph_replay["no_overflow"] and all(p.check() for p in plans)
        else:
            graph_replay = {"unavailable": why}
        del graphs

    e2e = None
    if not args.no_e2e:
        e2e_feed["steps"] = 3
        for i in range(3):
            step_e2e(i)
        e2e_drain()
        e2e_losses.clear()
        e2e_feed["steps"] = args.steps
        e2e_ms, _, _, _ = timed(step_e2e, args.steps, finish=e2e_drain)
        assert len(e2e_losses) == args.steps and all(x == x for x in e2e_losses)
    # ---- the unmodified wrapper path: the reference's per-view Python loop on the
    # gsplat-compatible API -- what SplatFormer runs when only `import gsplat` is redirected here
    wrapper_ms = None
    if args.wrapper_steps > 0 and world == 1:
        def step_wrapper(i):
            raw, cams, gt = resident[i % NSCENE]
            for p in raw.values():
                p.grad = None
            rgbs, _ = gs_utils.per_view_multiimgs(raw, cams)
            loss = 0.0
            for rgb, g in zip(rgbs, gt):                     # train.py:274-275
                loss = loss + (rgb - g.float() / 255.0).abs().mean()
            (loss * inv_views).backward()
        for i in range(2):
            step_wrapper(i)
        wrapper_ms, wrapper_launches, _, _ = timed(step_wrapper, args.wrapper_steps)
    # ---- BASELINE config 3: the train-step render loss, 8 scenes x 32 views per step, in BOTH layouts ----
    config3 = None
    c3_steps = args.config3_steps if args.config3_steps is not None else max(4, args.steps // 5)
    if c3_steps > 0:
        config3 = run_config3(args, c3_steps, rank, world, dev, timed, V, H, W, n)
    # ---- BASELINE configs 4 and 5 (one GPU): driver-visible numbers for the other named shapes ----
    other = None
    if args.other_configs and world == 1:
        other = run_other_configs(dev)
    clocks = sampler.stop() if sampler else None

    if rank == 0:
        ms_step = ms_total / args.steps
        scenes_per_step = world if args.shard == "scene" else 1
        pix_step = scenes_per_step * V * H * W
        value = pix_step / (ms_step * 1e-3) / 1e6
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if peaks else "fallback 6650 GB/s"
        per_view = stage_bytes_per_view(n, K, isect_per_view, H * W, tiles)
        # DRAM bytes per launch of each kernel from the committed ncu --set full capture of this same
        # command (profiles/ncu_traffic.json); only valid for the captured shape (32 views per launch)
        traffic = {}
        try:
            traffic = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
        except Exception:
            pass
        top = max(stage_ms, key=stage_ms.get) if stage_ms else None
        roofline = None
        stages = {}
        for name, tot in stage_ms.items():
            calls = max(stage_calls.get(name, 1), 1)
            avg_ms = tot / calls
            entry = {"ms_per_launch": avg_ms, "share_of_step": tot / ms_total}
            if name in per_view:
                gbs = per_view[name] * v_local / (avg_ms * 1e-3) / 1e9
                entry.update(alg_bytes_per_launch=per_view[name] * v_local, gbs=gbs, frac=gbs / peak)
            if name in traffic and v_local == 32 and V == 32:
                entry["dram_bytes_per_launch_ncu"] = traffic[name]
            stages[name] = entry
        if top and top in per_view:
            roofline = {"kernel": top, "bound": "hbm", "achieved": stages[top]["gbs"], "peak": peak, "unit": "GB/s",
                        "frac": stages[top]["frac"], "traffic": stages[top].get("dram_bytes_per_launch_ncu"),
                        "traffic_source": traffic.get("_source") if stages[top].get("dram_bytes_per_launch_ncu") else None,
                        "peak_source": peak_src,
                        "alg_bytes_per_launch": stages[top]["alg_bytes_per_launch"],
                        "ms_per_launch": stages[top]["ms_per_launch"], "views_per_launch": v_local}
        b_alg_view = sum(per_view.values())
        step_gbs = b_alg_view * v_local / (ms_step * 1e-3) / 1e9
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
            "scaling": "weak" if args.shard == "scene" else "strong", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic (seeded shell scene, orbit cameras, uint8 random target images)",
            "config": {"workload": WORKLOAD, "gaussians": n, "sh_degree": deg, "views": V, "image": [H, W],
                       "scenes_per_step": scenes_per_step, "sharding": args.shard if world > 1 else "single",
                       "l2": f"inputs rotate over {NSCENE} scenes per rank ({NSCENE * h2d_bytes / 1e6:.0f} MB) and "
                             "the per-step working set (~0.6 GB) exceeds the 126 MB L2",
                       "intersections_per_view": isect_per_view},
            "gaussians_per_s": scenes_per_step * V * n / (ms_step * 1e-3),
            "hbm_frac_step": {"alg_bytes_per_view": b_alg_view, "gbs": step_gbs, "frac": step_gbs / peak,
                              "peak": peak, "peak_source": peak_src},
            "roofline": roofline, "stages": stages, "gpu_launches": launches, "clocks": clocks,
        }
        if not args.no_e2e:
            e2e_step = e2e_ms / args.steps
            line["e2e"] = {"value": pix_step / (e2e_step * 1e-3) / 1e6, "unit": UNIT, "ms_per_step": e2e_step,
                           "h2d_bytes_per_step": int(h2d_bytes), "d2h_bytes_per_step": 4 + 8,
                           "d2h": "the step's loss (4 B, read on the host one step later, all steps inside the timed "
                                  "region) + the binning totals (8 B, read when the backward starts)",
                           "api": "splatformer_b200.loader.ScenePrefetcher(" + ("loader.pack_scene(host scene)" if args.e2e_packed else "host scenes") + ") -> gs_utils.render_l1_loss(gs_params, "
                                  "cameras, images) + backward + pinned read of the loss"}
        if graph_replay is not None:
            if "ms_per_step" in graph_replay:
                graph_replay["value"] = pix_step / (graph_replay["ms_per_step"] * 1e-3) / 1e6
                graph_replay["unit"] = UNIT
                if "e2e_ms_per_step" in graph_replay:
                    graph_replay["e2e_value"] = pix_step / (graph_replay["e2e_ms_per_step"] * 1e-3) / 1e6
            line["graph_replay"] = graph_replay
            if args.headline == "graph" and graph_replay.get("no_overflow") and "value" in graph_replay:
                # the headline is the captured step (the same kernels, launched by one graph replay per step); the
                # eager loop's numbers stay next to it; per-kernel stage times / roofline come from the eager loop
                # (CUDA events around every launch)
                line["eager"] = {"value": line["value"], "ms_per_step": line["ms_per_step"]}
                g_ms = graph_replay["ms_per_step"]
                line["value"], line["ms_per_step"] = graph_replay["value"], g_ms
                line["gaussians_per_s"] = scenes_per_step * V * n / (g_ms * 1e-3)
                g_gbs = b_alg_view * v_local / (g_ms * 1e-3) / 1e9
                line["hbm_frac_step"].update(gbs=g_gbs, frac=g_gbs / peak)
                line["config"]["launch"] = ("one CUDA graph replay per step (splatformer_b200.graphs.capture_step "
                                            "of render_l1_loss + backward); `eager` = the same calls uncaptured")
                if "e2e" in line and "e2e_value" in graph_replay:
                    line["eager"].update(e2e_value=line["e2e"]["value"], e2e_ms_per_step=line["e2e"]["ms_per_step"])
                    line["e2e"].update(
                        value=graph_replay["e2e_value"], ms_per_step=graph_replay["e2e_ms_per_step"],
                        d2h_bytes_per_step=4 + 12,
                        d2h="the step's loss (4 B, read on the host one step later, all steps inside the timed region) "
                            "+ the binning totals (12 B into a pinned slot, checked after the loop: plan.check())",
                        api="pinned host scene -> copy_ into the captured input tensors on a copy stream, one step "
                            "ahead -> graph.replay() (capture_step of gs_utils.render_l1_loss + backward) -> pinned "
                            "read of the loss")
            else:
                line["config"]["launch"] = "eager"
        if config3 is not None:
            line["config3"] = config3
        if other is not None:
            line["configs"] = other
        if wrapper_ms is not None:
            w_step = wrapper_ms / args.wrapper_steps
            line["wrapper_path"] = {"value": pix_step / (w_step * 1e-3) / 1e6, "unit": UNIT, "ms_per_step": w_step,
                                    "steps": args.wrapper_steps, "abi_launches_per_step": wrapper_launches / args.wrapper_steps,
                                    "api": "splatformer_b200.gs_utils.per_view_multiimgs = the loop of "
                                           "utils/gs_utils.py:20-114 on gsplat.spherical_harmonics / project_gaussians / "
                                           "rasterize_gaussians, + torch L1 loss + backward"}
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(args.cpu_views)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
