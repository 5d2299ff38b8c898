"""splatformer_b200.graphs.capture_step: a whole step (forward + backward through a graph-safe RenderPlan) replayed
from one CUDA graph gives what the eager calls give -- also after new inputs have been copied into the captured
tensors -- and plan.check() reports a replay whose tile segments overflowed."""
import pytest
import torch

from splatformer_b200 import gs_utils, render, synth
from splatformer_b200.graphs import capture_step
import helpers as H

pytestmark = pytest.mark.gpu


def _scene(cuda, n, deg, size, views, seed):
    gs = synth.shell_scene(n, deg, seed=seed, scale_mult=0.3)
    cams = synth.orbit_cameras(views, size[0], size[1], background=(0.2, 0.3, 0.1))
    gt = (synth.target_images(views, size[1], size[0], seed=seed + 1) * 255.0).round().to(torch.uint8)
    return gs, {k: v.to(cuda) for k, v in cams.items()}, gt.to(cuda)


def _eager(gs, cams, gt, cuda):
    raw = {k: v.to(cuda).requires_grad_() for k, v in gs.items()}
    l1, _ = gs_utils.render_l1_loss(raw, cams, gt, plan=render.RenderPlan())
    loss = l1.sum() / gt.shape[0]
    loss.backward()
    return float(loss.detach()), {k: v.grad.clone() for k, v in raw.items()}


def test_captured_step_replays_like_the_eager_step(cuda, sb_lib):
    views, size = 3, (144, 112)
    gs_a, cams, gt_a = _scene(cuda, 2500, 3, size, views, seed=31)
    gs_b, _, gt_b = _scene(cuda, 2500, 3, size, views, seed=47)
    raw = {k: v.to(cuda).requires_grad_() for k, v in gs_a.items()}
    gt = gt_a.clone()
    plan = render.RenderPlan()
    out = {}

    def step(i):
        for p in raw.values():
            p.grad = None
        l1, _ = gs_utils.render_l1_loss(raw, cams, gt, plan=plan)
        loss = l1.sum() / views
        loss.backward()
        out["loss"] = loss.detach()

    for _ in range(2):                  # eager warm-up: the plan learns the shape's capacities
        step(0)
    graph, why = capture_step(step, plan, passes=1)
    assert graph is not None, why
    # replay on the captured inputs (scene a), twice
    ref_loss, ref = _eager(gs_a, cams, gt_a, cuda)
    for _ in range(2):
        graph.replay()
        torch.cuda.synchronize()
        assert plan.check()
        assert abs(float(out["loss"]) - ref_loss) <= 1e-6 * max(1.0, abs(ref_loss))
        for k in raw:
            H.assert_grad_close(raw[k].grad, ref[k], f"replay/{k}")
    # new inputs are copied INTO the captured tensors (scene b), the gradients come out of the same .grad tensors
    with torch.no_grad():
        for k in raw:
            raw[k].copy_(gs_b[k].to(cuda))
        gt.copy_(gt_b)
    graph.replay()
    torch.cuda.synchronize()
    assert plan.check()
    ref_loss_b, ref_b = _eager(gs_b, cams, gt_b, cuda)
    assert H.rel_l2(ref_b["means"], ref["means"]) > 0.1     # the two scenes do differ
    assert abs(float(out["loss"]) - ref_loss_b) <= 1e-6 * max(1.0, abs(ref_loss_b))
    for k in raw:
        H.assert_grad_close(raw[k].grad, ref_b[k], f"replay-new-inputs/{k}")


def test_check_reports_a_replay_that_overflowed(cuda, sb_lib):
    """Captured with a sparse scene under a small segment capacity; then a much denser scene is copied into the
    captured tensors: its tiles run over the segments, which a graph cannot repair -- check() must say so."""
    views, size, n = 2, (96, 96), 800
    gs_sparse = synth.shell_scene(n, 0, seed=5, scale_mult=0.05)
    gs_dense = synth.shell_scene(n, 0, seed=5, scale_mult=1.5)
    cams = {k: v.to(cuda) for k, v in synth.orbit_cameras(views, size[0], size[1]).items()}
    gt = (synth.target_images(views, size[1], size[0], seed=2) * 255.0).round().to(torch.uint8).to(cuda)
    raw = {k: v.to(cuda).requires_grad_() for k, v in gs_sparse.items()}
    plan = render.RenderPlan(tile_capacity=64)

    def step(i):
        for p in raw.values():
            p.grad = None
        l1, _ = gs_utils.render_l1_loss(raw, cams, gt, plan=plan)
        (l1.sum() / views).backward()

    for _ in range(4):                  # the plan doubles the capacity until the sparse scene fits
        step(0)
    torch.cuda.synchronize()
    cap = plan.capacity(sb_lib, views, 36, n)
    assert 0 < cap < n, "the dense scene (every Gaussian in every tile) must not fit what the sparse one needed"
    graph, why = capture_step(step, plan, passes=1)
    assert graph is not None, why
    graph.replay()
    torch.cuda.synchronize()
    assert plan.check()
    with torch.no_grad():
        for k in raw:
            raw[k].copy_(gs_dense[k].to(cuda))
    graph.replay()
    torch.cuda.synchronize()
    assert not plan.check()
