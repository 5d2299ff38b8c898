"""Multi-process logic of the view-/scene-sharded render loss on CPU (gloo, world_size 2):
partitioning, the packed per-Gaussian gradient all-reduce, and the property the view-sharded
layout rests on -- the sum over ranks of the gradients of each rank's views equals the gradient of
the whole scene.  The render itself is the oracle here (no GPU on this box); the GPU path runs the
same host logic through bench.py --shard view."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from splatformer_b200 import dist as sbdist
from splatformer_b200 import synth


def test_view_and_scene_slices_partition_exactly():
    for v in (1, 4, 5, 32, 33):
        for world in (1, 2, 3, 4, 8):
            spans = [sbdist.view_slice(v, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == v
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1
    assert sorted(sum((sbdist.scene_slice(8, r, 3) for r in range(3)), [])) == list(range(8))
    with pytest.raises(ValueError):
        sbdist.view_slice(4, 2, 2)


def test_pack_unpack_round_trip():
    gs = synth.shell_scene(50, 2, seed=1)
    params = {k: v.clone().requires_grad_() for k, v in gs.items()}
    torch.manual_seed(0)
    for p in params.values():
        p.grad = torch.randn_like(p)
    want = {k: p.grad.clone() for k, p in params.items()}
    buf, layout = sbdist.pack_grads(params)
    assert buf.shape == (50, 3 + 3 + 4 + 1 + 3 + 8 * 3)          # 11 + 3K floats per Gaussian, K = 9
    for p in params.values():
        p.grad = None
    sbdist.unpack_grads(buf, layout, params)
    for k in params:
        assert torch.equal(params[k].grad, want[k])


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import ref_torch as R
        torch.set_num_threads(2)
        views, w, h = 4, 48, 40
        gs = synth.shell_scene(300, 1, seed=3, scale_mult=0.6)
        cams = synth.orbit_cameras(views, w, h)
        gt = synth.target_images(views, h, w, seed=4)
        lo, hi = sbdist.view_slice(views, rank, world)
        raw = {k: v.clone().requires_grad_() for k, v in gs.items()}
        my = dict(cams)
        my["camera_to_worlds"] = cams["camera_to_worlds"][lo:hi]
        rgbs, _ = R.render_views(raw, my)
        loss = sum((rgb - gt[lo + i]).abs().sum() for i, rgb in enumerate(rgbs)) / (views * h * w * 3)
        loss.backward()
        sbdist.allreduce_gaussian_grads(raw)
        total = torch.tensor([float(loss.detach())], dtype=torch.float64)
        dist.all_reduce(total)
        torch.save({"grads": {k: p.grad for k, p in raw.items()}, "loss": float(total)},
                   os.path.join(out_dir, f"rank{rank}.pt"))
    finally:
        dist.destroy_process_group()


def test_view_sharded_gradients_sum_to_whole_scene(tmp_path):
    from oracle import ref_torch as R
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    views, w, h = 4, 48, 40
    gs = synth.shell_scene(300, 1, seed=3, scale_mult=0.6)
    cams = synth.orbit_cameras(views, w, h)
    gt = synth.target_images(views, h, w, seed=4)
    raw = {k: v.clone().requires_grad_() for k, v in gs.items()}
    rgbs, _ = R.render_views(raw, cams)
    loss = sum((rgb - gt[i]).abs().sum() for i, rgb in enumerate(rgbs)) / (views * h * w * 3)
    loss.backward()
    r0 = torch.load(os.path.join(tmp_path, "rank0.pt"))
    r1 = torch.load(os.path.join(tmp_path, "rank1.pt"))
    assert abs(r0["loss"] - float(loss.detach())) < 1e-6
    for k in raw:
        assert torch.equal(r0["grads"][k], r1["grads"][k]), "ranks disagree after the all-reduce"
        assert torch.allclose(r0["grads"][k], raw[k].grad, rtol=1e-4, atol=1e-7), k


def _sync_worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from splatformer_b200 import render
        torch.manual_seed(rank)
        n, k = 40, 4
        params = {"means": torch.randn(n, 3), "scales": torch.randn(n, 3), "quats": torch.randn(n, 4),
                  "opacities": torch.randn(n, 1), "features_dc": torch.randn(n, 3),
                  "features_rest": torch.randn(n, k - 1, 3)}
        params = {kk: v.requires_grad_() for kk, v in params.items()}

        class FlatGrads(torch.autograd.Function):
            """Stands in for the render: its backward hands out blocks of ONE flat buffer (render.grad_layout)."""
            @staticmethod
            def forward(ctx, *xs):
                ctx.shapes = [tuple(x.shape) for x in xs]
                return sum((x * (i + 1)).sum() for i, x in enumerate(xs))

            @staticmethod
            def backward(ctx, g):
                blocks, size = render.grad_layout(n, ctx.shapes[4], ctx.shapes[5])
                flat = torch.empty(size)
                outs = []
                for i, (name, off, shape) in enumerate(blocks):
                    cnt = 1
                    for d in shape:
                        cnt *= d
                    view = flat[off:off + cnt].view(shape)
                    view.fill_(float(i + 1) * (rank + 1))
                    outs.append(view)
                FlatGrads.flat = flat
                return tuple(outs)

        for overlap in (False, True):
            for p in params.values():
                p.grad = None
            routed = sbdist.sync_gaussian_grads(params, overlap=overlap)
            keys = ("means", "scales", "quats", "opacities", "features_dc", "features_rest")
            FlatGrads.apply(*[routed[kk] for kk in keys]).backward()
            assert sbdist.wait_pending() == (1 if overlap else 0)
            for i, kk in enumerate(keys):
                want = float(i + 1) * sum(r + 1 for r in range(world))
                assert torch.all(params[kk].grad == want), (kk, overlap)
                # in place on the render's buffer: the leaf gradient is a block of it, nothing was packed
                assert params[kk].grad.untyped_storage().data_ptr() == FlatGrads.flat.untyped_storage().data_ptr()
        # the exchange issued OUTSIDE autograd (steps replayed from a CUDA graph): the flat buffer behind the
        # .grad fields is found again and reduced in place
        for overlap in (False, True):
            for p in params.values():
                p.grad = None
            keys = ("means", "scales", "quats", "opacities", "features_dc", "features_rest")
            FlatGrads.apply(*[params[kk] for kk in keys]).backward()
            flat = sbdist.flat_grad_buffer(params)
            assert flat is not None and flat.untyped_storage().data_ptr() == FlatGrads.flat.untyped_storage().data_ptr()
            assert flat.numel() == sum(p.numel() for p in params.values())
            sbdist.allreduce_flat(flat, overlap=overlap)
            assert sbdist.wait_pending() == (1 if overlap else 0)
            for i, kk in enumerate(keys):
                assert torch.all(params[kk].grad == float(i + 1) * sum(r + 1 for r in range(world))), (kk, overlap)
        # separate allocations have no flat buffer
        for p in params.values():
            p.grad = torch.ones_like(p)
        assert sbdist.flat_grad_buffer(params) is None
        # generic fallback: gradients that are separate allocations
        for p in params.values():
            p.grad = None
        routed = sbdist.sync_gaussian_grads(params)
        sum((routed[kk] * (rank + 2)).sum() for kk in params).backward()
        for kk in params:
            assert torch.all(params[kk].grad == float(sum(r + 2 for r in range(world)))), kk
        torch.save({"ok": True}, os.path.join(out_dir, f"sync{rank}.pt"))
    finally:
        dist.destroy_process_group()


def test_sync_gaussian_grads_reduces_in_place_on_the_flat_buffer(tmp_path):
    world = 2
    mp.spawn(_sync_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    assert all(os.path.exists(os.path.join(tmp_path, f"sync{r}.pt")) for r in range(world))
