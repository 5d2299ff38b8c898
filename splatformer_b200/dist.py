"""Multi-GPU partitioning of the render loss: one process per GPU under torchrun.

* by scene -- SplatFormer's own layout (one scene per GPU per step,
  /root/reference/dataset/Loader.py:15-17, per-rank scene split /root/reference/dataset/GS.py:92-118):
  independent units, no collective on the render path.
* by view  -- the Gaussians of a scene are replicated, its V views are split contiguously over the
  ranks; the forward needs no exchange, the backward needs exactly one: the sum of the per-Gaussian
  gradients.  They are packed into ONE (N, 11 + 3K) float32 buffer (means 3, scales 3, quats 4,
  opacity 1, SH 3K) and reduced with a single NCCL all-reduce over NVLink / NVSwitch.

  `SyncGaussianGrads` puts that exchange INTO the autograd graph, between the scene's parameters and
  the render: the render's backward writes all per-Gaussian gradients into one flat allocation
  (render.grad_layout), so the all-reduce runs in place on exactly that buffer -- no pack, no unpack --
  and, launched asynchronously, proceeds on NCCL's stream while the next scene renders on the compute
  stream (`wait_pending()` joins them before the gradients are used).

Works with any torch.distributed backend (the CPU tests run it on gloo with world_size 2).
"""
from __future__ import annotations

from typing import Dict, List, Sequence, Tuple

import torch
import torch.distributed as dist
from torch import Tensor

GRAD_KEYS = ("means", "scales", "quats", "opacities", "opacities_sigmoid", "features_dc", "features_rest")


def view_slice(num_views: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous [lo, hi) block of views of rank `rank`; the first (V mod world) ranks get one more."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError(f"bad rank/world {rank}/{world}")
    base, extra = divmod(num_views, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def scene_slice(num_scenes: int, rank: int, world: int) -> List[int]:
    """Scenes of rank `rank` (round-robin, as the reference's per-rank split of its scene list)."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError(f"bad rank/world {rank}/{world}")
    return list(range(rank, num_scenes, world))


def pack_grads(params: Dict[str, Tensor]) -> Tuple[Tensor, List[Tuple[str, torch.Size]]]:
    """Concatenate the .grad of every per-Gaussian parameter into one (N, C) buffer."""
    cols, layout = [], []
    n = None
    for key in GRAD_KEYS:
        p = params.get(key)
        if p is None or not p.requires_grad:
            continue
        g = p.grad if p.grad is not None else torch.zeros_like(p)
        n = p.shape[0] if n is None else n
        if p.shape[0] != n:
            raise ValueError(f"{key}: leading dimension {p.shape[0]} != {n}")
        cols.append(g.reshape(n, -1).to(torch.float32))
        layout.append((key, p.shape))
    if not cols:
        raise ValueError("no per-Gaussian parameter requires grad")
    return torch.cat(cols, dim=1).contiguous(), layout


def unpack_grads(buf: Tensor, layout: Sequence[Tuple[str, torch.Size]], params: Dict[str, Tensor]) -> None:
    col = 0
    for key, shape in layout:
        width = 1
        for d in shape[1:]:
            width *= d
        params[key].grad = buf[:, col:col + width].reshape(shape).to(params[key].dtype)
        col += width


def allreduce_gaussian_grads(params: Dict[str, Tensor], group=None, async_op: bool = False):
    """Sum the per-Gaussian gradients of a view-sharded scene over the ranks (one all-reduce)."""
    buf, layout = pack_grads(params)
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return None
    work = dist.all_reduce(buf, op=dist.ReduceOp.SUM, group=group, async_op=async_op)
    if async_op:
        return work, (lambda: unpack_grads(buf, layout, params))
    unpack_grads(buf, layout, params)
    return None


# ---------------------------------------------------------------------------------------------
# The exchange inside the autograd graph
# ---------------------------------------------------------------------------------------------
_pending: list = []


def wait_pending() -> int:
    """Make the current stream wait for every all-reduce launched by SyncGaussianGrads(overlap=True) so far
    (a stream-side wait for NCCL, a blocking one for gloo).  Returns how many were outstanding."""
    n = len(_pending)
    for work, keep in _pending:
        work.wait()
    _pending.clear()
    return n


def _as_one_buffer(grads: Sequence[Tensor]):
    """If the gradients are disjoint contiguous blocks of ONE allocation (what render._RenderViews.backward
    returns), the flat tensor spanning them; else None."""
    if any(g is None or not g.is_contiguous() for g in grads):
        return None
    st = grads[0].untyped_storage()
    if any(g.untyped_storage().data_ptr() != st.data_ptr() or g.dtype != grads[0].dtype for g in grads):
        return None
    spans = sorted((g.storage_offset(), g.storage_offset() + g.numel()) for g in grads)
    for (a0, a1), (b0, b1) in zip(spans, spans[1:]):
        if b0 < a1 or b0 - a1 > 3:            # overlap, or a hole larger than the 16-byte block alignment
            return None
    lo, hi = spans[0][0], spans[-1][1]
    flat = grads[0].new_empty(0).set_(st, lo, (hi - lo,))
    return flat


def flat_grad_buffer(params: Dict[str, Tensor]):
    """The one flat tensor that spans the `.grad` fields of the per-Gaussian parameters when they are blocks of a
    single allocation (what the fused render's backward produces, render.grad_layout); None otherwise."""
    grads = [params[k].grad for k in GRAD_KEYS if k in params and params[k].requires_grad]
    if not grads or any(g is None for g in grads):
        return None
    return _as_one_buffer(grads)


def allreduce_flat(flat: Tensor, group=None, overlap: bool = False) -> None:
    """In-place sum over the ranks of a flat gradient buffer (flat_grad_buffer).  For steps whose render + backward
    is replayed from a CUDA graph (graphs.capture_step): a collective cannot live inside the capture, so the graph
    holds the render alone and the exchange is issued between replays -- on NCCL's stream, under the next scene's
    replay when overlap=True (join with wait_pending() before the gradients are read or the same graph is replayed
    again: its backward rewrites this buffer)."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return
    work = dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group, async_op=True)
    if overlap:
        _pending.append((work, flat))
    else:
        work.wait()


class SyncGaussianGrads(torch.autograd.Function):
    """y_i = x_i in the forward; in the backward the incoming gradients are summed over the ranks of `group`
    with ONE all-reduce.  Usage (view-sharded scene): params = SyncGaussianGrads.apply(group, overlap, *params)
    before handing them to the render.  overlap=True launches the all-reduce asynchronously and returns at
    once; the gradients must then reach fresh `.grad` fields (no accumulation into an existing one) and
    `wait_pending()` must be called before they are read."""

    @staticmethod
    def forward(ctx, group, overlap, *params):
        ctx.group, ctx.overlap = group, bool(overlap)
        return tuple(p.view_as(p) for p in params)

    @staticmethod
    def backward(ctx, *grads):
        world = dist.get_world_size(ctx.group) if (dist.is_available() and dist.is_initialized()) else 1
        live = [g for g in grads if g is not None]
        if world > 1 and live:
            flat = _as_one_buffer(live)
            packed = None
            if flat is None:                   # generic fallback: gather the pieces, reduce, scatter back
                packed = torch.cat([g.reshape(-1).to(torch.float32) for g in live])
                flat = packed
            work = dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=ctx.group, async_op=True)
            if packed is None and ctx.overlap:
                _pending.append((work, flat))
            else:
                work.wait()
                if packed is not None:
                    off = 0
                    for g in live:
                        g.copy_(packed[off:off + g.numel()].view_as(g))
                        off += g.numel()
        return (None, None) + tuple(grads)


def sync_gaussian_grads(params: Dict[str, Tensor], group=None, overlap: bool = False) -> Dict[str, Tensor]:
    """The per-Gaussian tensors of `params` routed through SyncGaussianGrads (others passed through)."""
    keys = [k for k in GRAD_KEYS if k in params and params[k].requires_grad]
    outs = SyncGaussianGrads.apply(group, overlap, *[params[k] for k in keys])
    res = dict(params)
    res.update(dict(zip(keys, outs)))
    return res
