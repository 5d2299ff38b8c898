"""Host -> device staging of scenes for the render loss (the data path right in front of the hot path).

SplatFormer's training loop receives, per step, one scene from its DataLoader -- raw Gaussian
parameters, cameras and target images in host memory (/root/reference/dataset/Loader.py:10-20) --
and moves it to the GPU with `gpu_utils.move_to_device` right before the forward
(/root/reference/train.py:234-236, /root/reference/utils/gpu_utils.py:3-9).  Done on the compute stream, that copy
(48 MB for a 100k-Gaussian SH3 scene with 32 uint8 512x512 targets) sits in front of every step:
0.45 ms of PCIe time during which the SMs idle.  `prefetch` runs it on a copy stream one scene
ahead, so the upload of scene i+1 overlaps the render of scene i; every scene is still copied from
host memory on every step.

Plumbing only (streams, events, the caching allocator's `record_stream`): no kernels.
"""
from __future__ import annotations

from typing import Any, Dict, Iterable, Iterator, Optional, Tuple

import torch
from torch import Tensor

Scene = Tuple[Dict[str, Tensor], Dict[str, Tensor], Optional[Tensor]]      # (gs_params, cameras, target images)


def _upload(x: Any, device: torch.device):
    if isinstance(x, Tensor):
        return x.to(device, non_blocking=True)
    if isinstance(x, dict):
        return {k: _upload(v, device) for k, v in x.items()}
    if isinstance(x, (tuple, list)):
        return type(x)(_upload(v, device) for v in x)
    return x


def _record(x: Any, stream: torch.cuda.Stream) -> None:
    if isinstance(x, Tensor):
        if x.is_cuda:
            x.record_stream(stream)
    elif isinstance(x, dict):
        for v in x.values():
            _record(v, stream)
    elif isinstance(x, (tuple, list)):
        for v in x:
            _record(v, stream)


class ScenePrefetcher:
    """Uploads items (nested dicts / tuples of host tensors, ideally pinned) on a private copy stream.

        pf = ScenePrefetcher(device)
        ticket = pf.submit(host_scene)      # enqueue the copies, returns immediately
        scene = pf.get(ticket)              # device tensors; the current stream waits for the copies
    """

    def __init__(self, device: torch.device):
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("ScenePrefetcher needs a CUDA device (the renderer has no CPU path)")
        self.stream = torch.cuda.Stream(device=self.device)

    def submit(self, item: Any):
        with torch.cuda.stream(self.stream):
            dev_item = _upload(item, self.device)
            ready = torch.cuda.Event()
            ready.record(self.stream)
        return dev_item, ready

    def get(self, ticket):
        dev_item, ready = ticket
        cur = torch.cuda.current_stream(self.device)
        cur.wait_event(ready)
        _record(dev_item, cur)              # the tensors were allocated on the copy stream
        return dev_item


def prefetch(items: Iterable[Any], device: torch.device, ahead: int = 1) -> Iterator[Any]:
    """Yields the items of `items` as device tensors, uploading `ahead` items in advance."""
    pf = ScenePrefetcher(device)
    it = iter(items)
    queue = []
    try:
        for _ in range(max(1, ahead)):
            queue.append(pf.submit(next(it)))
    except StopIteration:
        pass
    while queue:
        ticket = queue.pop(0)
        try:
            queue.append(pf.submit(next(it)))
        except StopIteration:
            pass
        yield pf.get(ticket)
