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


def test_pack_scene_round_trip_on_the_host():
    """One-buffer layout: every tensor comes back with its dtype, shape and bytes (0-dim, empty, uint8,
    int64 and nested containers included), at 256-byte aligned offsets."""
    g = torch.Generator().manual_seed(3)
    item = ({"means": torch.randn(37, 3, generator=g), "empty": torch.zeros(0, 4), "ids": torch.arange(5)},
            {"fx": torch.tensor(311.5), "c2w": torch.randn(2, 3, 4, generator=g), "name": "scene-7"},
            [torch.randint(0, 255, (2, 5, 7, 3), generator=g).to(torch.uint8), None])
    packed = loader.pack_scene(item, pin=False)
    assert all(e[2] % 256 == 0 for e in packed.entries)
    assert packed.nbytes == sum(t.numel() * t.element_size() for t in
                                (*item[0].values(), item[1]["fx"], item[1]["c2w"], item[2][0]))
    out = loader._unpack(packed.spec, packed.blob, packed.entries)
    assert isinstance(out, tuple) and isinstance(out[2], list) and out[2][1] is None and out[1]["name"] == "scene-7"
    for a, b in ((item[0]["means"], out[0]["means"]), (item[0]["empty"], out[0]["empty"]),
                 (item[0]["ids"], out[0]["ids"]), (item[1]["fx"], out[1]["fx"]), (item[1]["c2w"], out[1]["c2w"]),
                 (item[2][0], out[2][0])):
        assert a.dtype == b.dtype and a.shape == b.shape and torch.equal(a, b)


@pytest.mark.gpu
def test_packed_scene_uploads_in_one_copy_and_renders_like_a_resident_one(cuda, sb_lib):
    from splatformer_b200 import gs_utils, synth
    gs = synth.shell_scene(1500, 2, seed=5, scale_mult=0.3)
    cams = synth.orbit_cameras(3, 64, 48)
    gt = (synth.target_images(3, 48, 64, seed=5) * 255).round().to(torch.uint8)
    raw = {k: v.to(cuda).requires_grad_() for k, v in gs.items()}
    l1_ref, _ = gs_utils.render_l1_loss(raw, {k: v.to(cuda) for k, v in cams.items()}, gt.to(cuda))
    l1_ref.sum().backward()
    packed = loader.pack_scene((gs, cams, gt))
    assert packed.blob.is_pinned()
    pf = loader.ScenePrefetcher(cuda)
    for _ in range(2):                                   # the same packed scene can be submitted again
        gs_d, cams_d, gt_d = pf.get(pf.submit(packed))
        assert gt_d.dtype == torch.uint8 and gs_d["means"].data_ptr() % 16 == 0
        raw_d = {k: v.requires_grad_() for k, v in gs_d.items()}
        l1, _ = gs_utils.render_l1_loss(raw_d, cams_d, gt_d)
        l1.sum().backward()
        assert torch.equal(l1.detach(), l1_ref.detach())
        for k in raw:
            assert torch.equal(raw_d[k].grad, raw[k].grad) or \
                float((raw_d[k].grad - raw[k].grad).norm() / raw[k].grad.norm()) <= 1e-5
