"""splatformer_b200.loader: the prefetching upload in front of the render loss."""
import pytest
import torch

from splatformer_b200 import loader


def test_prefetcher_refuses_cpu_devices():
    with pytest.raises(RuntimeError):
        loader.ScenePrefetcher(torch.device("cpu"))


@pytest.mark.gpu
def test_prefetch_yields_every_item_in_order(cuda):
    items = [({"a": torch.full((1000, 3), float(i)).pin_memory()}, {"k": torch.tensor(i)}, torch.full((8,), i, dtype=torch.uint8))
             for i in range(7)]
    out = list(loader.prefetch(iter(items), cuda, ahead=2))
    assert len(out) == 7
    for i, (gs, cams, gt) in enumerate(out):
        assert gs["a"].is_cuda and float(gs["a"].mean()) == float(i)
        assert int(cams["k"]) == i and gt.dtype == torch.uint8 and int(gt[0]) == i
    assert list(loader.prefetch(iter([]), cuda)) == []


@pytest.mark.gpu
def test_prefetched_scene_renders_like_a_resident_one(cuda, sb_lib):
    from splatformer_b200 import gs_utils, synth
    scenes = []
    for s in range(3):
        gs = synth.shell_scene(1200, 1, seed=s, scale_mult=0.3)
        cams = synth.orbit_cameras(2, 64, 48)
        gt = (synth.target_images(2, 48, 64, seed=s) * 255).round().to(torch.uint8)
        scenes.append(({k: v.pin_memory() for k, v in gs.items()}, cams, gt.pin_memory()))
    ref = []
    for gs, cams, gt in scenes:
        raw = {k: v.to(cuda).requires_grad_() for k, v in gs.items()}
        l1, _ = gs_utils.render_l1_loss(raw, {k: v.to(cuda) for k, v in cams.items()}, gt.to(cuda))
        l1.sum().backward()
        ref.append((l1.detach().clone(), raw["means"].grad.clone()))
    for (gs_d, cams_d, gt_d), (l1_ref, g_ref) in zip(loader.prefetch(iter(scenes), cuda), ref):
        raw = {k: v.requires_grad_() for k, v in gs_d.items()}
        l1, _ = gs_utils.render_l1_loss(raw, cams_d, gt_d)
        l1.sum().backward()
        assert torch.equal(l1.detach(), l1_ref)
        assert float((raw["means"].grad - g_ref).norm() / g_ref.norm()) <= 1e-5
