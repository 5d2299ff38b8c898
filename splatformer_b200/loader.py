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


def _parse_cpulist(text: str):
    cpus = set()
    for part in text.strip().split(","):
        if not part:
            continue
        lo, _, hi = part.partition("-")
        cpus.update(range(int(lo), int(hi or lo) + 1))
    return cpus


def bind_host_to_gpu(device_index: int) -> Dict[str, Any]:
    """Binds the calling process to the CPUs next to GPU `device_index` (the PCI device's `local_cpulist` in sysfs),
    so that the pinned staging buffers it allocates AFTERWARDS are first-touched on that GPU's NUMA node and the
    uploads do not cross the socket interconnect.  With one process per GPU and every rank uploading a scene per
    step, all ranks otherwise tend to draw from one socket's memory.  Call it before the first pinned allocation.
    Best effort: returns what it did ({"bound": False, "why": ...} where sysfs says nothing useful)."""
    import os
    try:
        props = torch.cuda.get_device_properties(device_index)
        addr = f"{getattr(props, 'pci_domain_id', 0):04x}:{props.pci_bus_id:02x}:{props.pci_device_id:02x}.0"
        base = f"/sys/bus/pci/devices/{addr}"
        cpus = _parse_cpulist(open(f"{base}/local_cpulist").read())
        try:
            node = int(open(f"{base}/numa_node").read())
        except Exception:                              # noqa: BLE001
            node = -1
        allowed = os.sched_getaffinity(0)
        want = cpus & allowed
        if not want or want == allowed:
            return {"bound": False, "why": "the GPU's local CPU list is the whole allowed set", "pci": addr,
                    "numa_node": node}
        os.sched_setaffinity(0, want)
        return {"bound": True, "pci": addr, "numa_node": node, "cpus": len(want)}
    except Exception as e:                             # noqa: BLE001
        return {"bound": False, "why": f"{type(e).__name__}: {e}"}


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


class PackedScene:
    """A scene flattened into ONE pinned host buffer (`pack_scene`): its upload is a single copy and a
    single allocation instead of one per tensor, which is what the submit / get calls cost on the host."""

    __slots__ = ("blob", "spec", "entries")

    def __init__(self, blob: Tensor, spec: Any, entries):
        self.blob, self.spec, self.entries = blob, spec, entries

    @property
    def nbytes(self) -> int:
        return sum(e[3] for e in self.entries)


_ALIGN = 256


def pack_scene(item: Any, pin: bool = True) -> PackedScene:
    """Flattens a nested dict / tuple / list of host tensors into one (pinned) uint8 buffer, every tensor
    at a 256-byte aligned offset.  Done once per scene where it is collated (a DataLoader worker)."""
    entries = []            # (dtype, shape, offset, nbytes)
    tensors = []
    total = 0

    def walk(x):
        nonlocal total
        if isinstance(x, Tensor):
            if x.is_cuda:
                raise ValueError("pack_scene takes host tensors")
            t = x.detach().contiguous()
            nbytes = t.numel() * t.element_size()
            entries.append((t.dtype, tuple(t.shape), total, nbytes))
            tensors.append(t)
            total += (nbytes + _ALIGN - 1) // _ALIGN * _ALIGN
            return ("t", len(entries) - 1)
        if isinstance(x, dict):
            return ("d", {k: walk(v) for k, v in x.items()})
        if isinstance(x, (tuple, list)):
            return ("l" if isinstance(x, list) else "u", [walk(v) for v in x])
        return ("o", x)

    spec = walk(item)
    blob = torch.empty(max(total, 1), dtype=torch.uint8)
    if pin:
        blob = blob.pin_memory()
    for t, (_, _, off, nbytes) in zip(tensors, entries):
        if nbytes:
            blob[off:off + nbytes].copy_(t.reshape(-1).view(torch.uint8))
    return PackedScene(blob, spec, entries)


def _unpack(spec: Any, dev_blob: Tensor, entries):
    kind, val = spec
    if kind == "t":
        dtype, shape, off, nbytes = entries[val]
        return dev_blob[off:off + nbytes].view(dtype).view(shape)
    if kind == "d":
        return {k: _unpack(v, dev_blob, entries) for k, v in val.items()}
    if kind == "l":
        return [_unpack(v, dev_blob, entries) for v in val]
    if kind == "u":
        return tuple(_unpack(v, dev_blob, entries) for v in val)
    return val


class ScenePrefetcher:
    """Uploads items (nested dicts / tuples of host tensors, ideally pinned) on a private copy stream.

        pf = ScenePrefetcher(device)
        ticket = pf.submit(host_scene)      # enqueue the copies, returns immediately
        scene = pf.get(ticket)              # device tensors; the current stream waits for the copies

    `host_scene` may be a `PackedScene` (see `pack_scene`): one copy, and the device tensors handed out
    are views of one device buffer.
    """

    def __init__(self, device: torch.device):
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("ScenePrefetcher needs a CUDA device (the renderer has no CPU path)")
        self.stream = torch.cuda.Stream(device=self.device)

    def submit(self, item: Any):
        with torch.cuda.stream(self.stream):
            if isinstance(item, PackedScene):
                dev_item = (item, item.blob.to(self.device, non_blocking=True))
            else:
                dev_item = _upload(item, self.device)
            ready = torch.cuda.Event()
            ready.record(self.stream)
        return dev_item, ready

    def get(self, ticket):
        dev_item, ready = ticket
        cur = torch.cuda.current_stream(self.device)
        cur.wait_event(ready)
        if isinstance(dev_item, tuple) and len(dev_item) == 2 and isinstance(dev_item[0], PackedScene):
            packed, dev_blob = dev_item
            dev_blob.record_stream(cur)     # allocated on the copy stream; the views share its storage
            return _unpack(packed.spec, dev_blob, packed.entries)
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
