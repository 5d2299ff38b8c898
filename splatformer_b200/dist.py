"""Multi-GPU partitioning of the render loss: one process per GPU under torchrun.

* by scene -- SplatFormer's own layout (one scene per GPU per step,
  /root/reference/dataset/Loader.py:15-17, per-rank scene split /root/reference/dataset/GS.py:92-118):
  independent units, no collective on the render path.
* by view  -- the Gaussians of a scene are replicated, its V views are split contiguously over the
  ranks; the forward needs no exchange, the backward needs exactly one: the sum of the per-Gaussian
  gradients.  They are packed into ONE (N, 11 + 3K) float32 buffer (means 3, scales 3, quats 4,
  opacity 1, SH 3K) and reduced with a single NCCL all-reduce over NVLink / NVSwitch.

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
