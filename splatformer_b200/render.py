"""Fused multi-view render: all V views of a scene in one pass per stage.

This is the B200-native form of the loop SplatFormer runs in
`rasterize_gaussians_to_multiimgs` (/root/reference/utils/gs_utils.py:20-27): instead of V times
(SH, project, cumsum + .item(), emit, sort, bins, rasterise) -- about ten launches and nine host
synchronisations per view -- every stage is launched once for the whole (V, N) / (V, H, W) batch
through the C ABI, view-independent data (SH coefficients, Gaussian parameters) is read from HBM
once per scene, and the per-Gaussian gradients are summed over the views on chip.
Results are the ones the per-view gsplat-compatible API (`splatformer_b200.api`) produces: the
same device functions run underneath.

Host synchronisation.  A training step (some input requires grad) is enqueued WITHOUT any host
synchronisation: the tile-segment binning reports "a tile overflowed its segment" through a flag the
device writes into pinned host memory, and that flag is only looked at when the backward starts (by
then the device is long past the point that produced it; the forward rasteriser is still running, so
the GPU never waits for the host).  On the rare overflow the forward is replayed right there with the
global sort, overwriting the image / loss-sum buffers in place, and the backward continues on the
corrected state (tensors the caller DERIVED from the outputs before calling backward keep the values of
the truncated pass -- for that one step the logged loss is off, the gradients are not).  The flag is read
right after the forward -- one synchronisation -- when nothing will come back to repair the pass
(no autograd: evaluation) and on the FIRST call of a (views, tiles, Gaussians) shape through a plan, which
is where a shape that does not suit the tile segments at all (configs 4 and 5) is found out.
"""
from __future__ import annotations

import contextlib
import os
from typing import Dict, List, Optional, Tuple

import torch
from torch import Tensor
from torch.autograd import Function

from . import _abi, utils

F32, I32 = torch.float32, torch.int32


class StageTimer:
    """CUDA-event timing of the individual ABI calls (used by bench.py for the roofline of the
    dominant kernel).  Events are recorded on the current stream around each stage."""

    def __init__(self):
        self.events: List[Tuple[str, torch.cuda.Event, torch.cuda.Event]] = []

    @contextlib.contextmanager
    def stage(self, name: str):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        yield
        b.record()
        self.events.append((name, a, b))

    def totals_ms(self) -> Dict[str, float]:
        torch.cuda.synchronize()
        out: Dict[str, float] = {}
        for name, a, b in self.events:
            out[name] = out.get(name, 0.0) + a.elapsed_time(b)
        return out

    def counts(self) -> Dict[str, int]:
        out: Dict[str, int] = {}
        for name, _, _ in self.events:
            out[name] = out.get(name, 0) + 1
        return out


_timer: Optional[StageTimer] = None


@contextlib.contextmanager
def profile_stages():
    """with profile_stages() as t: ...render...; t.totals_ms()"""
    global _timer
    prev, _timer = _timer, StageTimer()
    try:
        yield _timer
    finally:
        _timer = prev


_NVTX = bool(os.environ.get("SB_NVTX"))      # SB_NVTX=1: an NVTX range per stage (nsys / ncu --nvtx timelines)


def _stage(name: str):
    if _timer is not None:
        return _timer.stage(name)
    if _NVTX:
        return torch.cuda.nvtx.range("sb." + name)
    return contextlib.nullcontext()


def _c(t: Tensor) -> Tensor:
    return _abi.aligned16(t.contiguous())


USE_RECT_LISTS = True               # fused path, block width 16: rect-list rasteriser (csrc/raster_rect.cu)
DEFAULT_TILE_CAPACITY = 2048        # doubled per overflow of a (views, tiles, Gaussians) shape, up to the library's limit
# Scenes whose Gaussians cover many tiles each (large screen radii: a room seen from inside, the 4K stress
# shape) are binned faster by the depth-ordered global sort than by per-tile slot claims -- hundreds of
# millions of contended atomics and scattered 8-byte stores.  A shape that averaged more tiles per
# (view, Gaussian) than this goes to the global sort from its next call on.
GLOBAL_SORT_TILES_PER_GAUSSIAN = 16.0


class _Totals:
    """The three ints the tile sort reports -- intersections, overflow flags (bit 0: a tile segment, bit 1: the
    rect-list pool), rect-list entries -- on their way to the host: a pinned buffer, the event after which it is
    valid, and the plan entry they belong to."""

    __slots__ = ("host", "event", "key", "plan", "value", "folded")

    def __init__(self, host: Optional[Tensor], event, key, plan, value: Optional[Tuple[int, int]] = None):
        self.host, self.event, self.key, self.plan, self.value, self.folded = host, event, key, plan, value, False

    def ready(self) -> bool:
        return self.value is not None or (self.event is not None and self.event.query())

    def read(self) -> Tuple[int, int, int]:
        """(intersections, overflow flags, rect-list entries); waits for the device to have produced them.
        A pass enqueued under a graph-safe plan carries no event (none may be waited for inside a captured
        region): its totals are read after a stream synchronisation and are not cached -- every replay of the
        graph rewrites the slot."""
        if self.event is None and self.host is not None:
            torch.cuda.current_stream().synchronize()
            vals = (int(self.host[0]), int(self.host[1]), int(self.host[2]))
            self.plan._fold(self.key, *vals)
            return vals
        if self.value is None:
            self.event.synchronize()
            self.value = (int(self.host[0]), int(self.host[1]), int(self.host[2]))
        if not self.folded:
            self.folded = True
            self.plan._fold(self.key, *self.value)
        return self.value


class RenderPlan:
    """What a call site has learnt about its scenes' binning: the tile-segment capacity per
    (views, tiles, Gaussians) shape (doubled after an overflow) and whether the shape is better served by
    the global sort.  `render_views(..., plan=...)` lets every caller keep its own; the module-level
    default is only used when none is passed."""

    def __init__(self, tile_capacity: Optional[int] = None):
        self.tile_capacity = tile_capacity
        # graph_safe: enqueue passes with nothing a CUDA graph capture could not hold -- no event that the host
        # waits for, no flag read in the backward, no replay.  For call sites that capture whole steps
        # (torch.cuda.graph) after an eager warm-up has taught the plan its shapes; `check()` after a replay
        # tells whether any captured pass overflowed (then the step has to be redone eagerly).
        self.graph_safe = False
        self.graph_totals: List[_Totals] = []
        self._graph_slot_pool: List[Tensor] = []
        self.overflowed: Dict[Tuple[int, int, int], int] = {}
        self.prefer_global: Dict[Tuple[int, int, int], bool] = {}
        self.pending: List[_Totals] = []
        self.last: Optional[_Totals] = None
        self.replays = 0
        self.seen: set = set()               # shapes whose first call (read synchronously) is behind us
        self.pool_used: Dict[Tuple[int, int, int], int] = {}     # rect-list entries the shape's scenes needed so far
        self._slots: List[Tensor] = []       # pinned (2,) int32 landing buffers of the totals, recycled
        self._slot_users: List[Optional[_Totals]] = []

    def host_slot(self, totals_factory) -> "_Totals":
        """A pinned landing buffer for a pass's totals.  Allocating pinned memory per step would cost a
        cudaHostAlloc whenever the host runs ahead of the device (the caching host allocator cannot recycle a
        block whose copy is still in flight), so the plan keeps a small ring and reuses a slot once its totals
        have been read; if every slot is still in flight the oldest one is waited for."""
        idx = None
        for i, user in enumerate(self._slot_users):
            if user is None or user.ready():
                idx = i
                break
        if idx is None and len(self._slots) < 16:
            self._slots.append(torch.zeros(3, dtype=I32).pin_memory())
            self._slot_users.append(None)
            idx = len(self._slots) - 1
        if idx is None:
            idx = 0
        prev = self._slot_users[idx]
        if prev is not None:
            prev.read()                      # (copies the two ints out of the slot before it is overwritten)
        t = totals_factory(self._slots[idx])
        self._slot_users[idx] = t
        return t

    def capacity(self, lib, v: int, num_tiles: int, n: int) -> int:
        """Segment capacity for the tile sort, or 0 to use the global sort."""
        self.poll()
        key = (v, num_tiles, n)
        if self.prefer_global.get(key):
            return 0
        base = self.tile_capacity if self.tile_capacity is not None else DEFAULT_TILE_CAPACITY
        cap = base << self.overflowed.get(key, 0)
        if cap > lib.sb_tile_sort_capacity() or v * num_tiles * cap >= 2 ** 31:
            return 0
        return cap

    def pool_capacity(self, key, cap: int) -> int:
        """Ints to reserve for the rect lists of a pass: 1.3 x what the shape's scenes needed so far, or, before
        anything is known, 12 per (view, Gaussian) -- never more than the fixed-capacity layout would take."""
        v, num_tiles, n = key
        ceiling = v * num_tiles * 8 * cap
        used = self.pool_used.get(key)
        want = int(used * 1.3) + (1 << 20) if used else max(12 * v * n, 1 << 20)
        return max(1024, min(ceiling, (want + (1 << 20) - 1) >> 20 << 20))

    def reserve_graph_slots(self, count: int) -> None:
        """Pinned landing buffers for `count` passes about to be captured (pinned memory cannot be allocated
        while a stream is capturing)."""
        while len(self._graph_slot_pool) < count:
            self._graph_slot_pool.append(torch.zeros(3, dtype=I32).pin_memory())

    def graph_slot(self) -> Tensor:
        if not self._graph_slot_pool:
            self.reserve_graph_slots(1)          # (raises under capture: call reserve_graph_slots before)
        return self._graph_slot_pool.pop()

    def check(self) -> bool:
        """Graph-safe plans: True if no pass enqueued (captured) through this plan overflowed in its latest
        execution.  Synchronises the stream."""
        ok = True
        for t in self.graph_totals:
            ok = ok and t.read()[1] == 0
        return ok

    def poll(self) -> None:
        """Fold in the totals of earlier calls that have reached the host by now (never blocks)."""
        keep = []
        for t in self.pending:
            if t.ready():
                t.read()
            else:
                keep.append(t)
        self.pending = keep

    def _fold(self, key, num_intersects: int, overflow: int, pool_used: int = 0) -> None:
        if overflow & 1:
            self.overflowed[key] = self.overflowed.get(key, 0) + 1
        if pool_used > 0:
            self.pool_used[key] = max(self.pool_used.get(key, 0), pool_used)
        v, _, n = key
        if num_intersects > GLOBAL_SORT_TILES_PER_GAUSSIAN * v * n:
            self.prefer_global[key] = True

    def last_num_intersects(self) -> int:
        return self.last.read()[0] if self.last is not None else 0


_default_plan = RenderPlan()
_last_plan: RenderPlan = _default_plan
# Tests that inspect the binning of a pass made under torch.no_grad() (no grad_fn to reach it through) set
# KEEP_LAST_PASS; the pass object (device buffers of the whole step) then stays alive until the next call.
KEEP_LAST_PASS = False
_last_pass = None


def last_num_intersects() -> int:
    """Intersections (over all views) of the most recent render_views call (waits for that call's
    binning to finish on the device; bench.py and the tests read it outside the timed loops)."""
    return _last_plan.last_num_intersects()


_side_streams: Dict[int, torch.cuda.Stream] = {}


def _side_stream(device) -> torch.cuda.Stream:
    idx = device.index if device.index is not None else torch.cuda.current_device()
    if idx not in _side_streams:
        _side_streams[idx] = torch.cuda.Stream(device=device)
    return _side_streams[idx]


class _Pass:
    """The device work of one forward pass.  Kept as an object so that the backward can replay it (global
    sort) over the same output buffers when the optimistic tile-segment binning turns out to have overflowed."""

    def __init__(self, lib, st, inputs, meta, gt, gt_ready):
        self.lib, self.st = lib, st
        (self.means, self.scales, self.quats, self.opacities, self.color_src, self.color_rest, self.viewmats,
         self.campos, self.intrins, self.background) = inputs
        (self.v, self.n, self.sh_degree, self.flags, self.h, self.w, self.bw, self.clip, self.vm_stride,
         self.clamp_rgb) = meta
        self.gt, self.gt_ready = gt, gt_ready
        self.gt_dtype = None if gt is None else gt.dtype
        self.tiles_x = (self.w + self.bw - 1) // self.bw
        self.tiles_y = (self.h + self.bw - 1) // self.bw
        self.num_tiles = self.tiles_x * self.tiles_y
        dev = self.means.device
        v, n = self.v, self.n
        self.packed = torch.empty(v, n, 12, dtype=F32, device=dev)
        self.depths = torch.empty(v, n, dtype=F32, device=dev)
        self.radii = torch.empty(v, n, dtype=I32, device=dev)
        self.tiles_hit = torch.empty(v, n, dtype=I32, device=dev)
        self.out_img = self.final_ts = self.final_idx = self.out_alpha = None
        self.loss_sums = None           # (V, 2) per-view sums of |rgb - gt| and (rgb - gt)^2, updated in place on a replay
        self.gids_sorted = self.tile_bins = self.rect_ids = self.rect_meta = None
        self.cap = 0

    # ---- buffers -------------------------------------------------------------------------------
    def alloc_images(self):
        # after the projection kernel is on its way: the host reaches the first launch sooner
        if self.out_img is not None:
            return
        dev, v, h, w = self.means.device, self.v, self.h, self.w
        self.out_img = torch.empty(v, h, w, 3, dtype=F32, device=dev)
        self.final_ts = torch.empty(v, h, w, dtype=F32, device=dev)
        self.final_idx = torch.empty(v, h, w, dtype=I32, device=dev)
        self.out_alpha = None if self.gt is not None else torch.empty(v, h, w, dtype=F32, device=dev)

    def _loss_partial(self, per_tile: int):
        if self.gt is None:
            return None
        return torch.empty(self.v, self.num_tiles * per_tile, 2, dtype=F32, device=self.means.device)

    def _finish_loss(self, partial):
        if partial is None:
            return
        sums = partial.sum(dim=1)                           # (V, 2)
        if self.loss_sums is None:
            self.loss_sums = sums
        else:
            self.loss_sums.data.copy_(sums)                 # replay: late readers see the corrected values

    # ---- stages --------------------------------------------------------------------------------
    def project(self, tile_counts, tile_keys, cap):
        lib, p = self.lib, _abi.ptr
        with _stage("project_sh_fwd"):
            _abi.check(lib.sb_views_project_fwd(
                self.v, self.n, p(self.means, F32, "means"), p(self.scales, F32, "scales"), 1.0,
                p(self.quats, F32, "quats"), p(self.opacities, F32, "opacities"),
                p(self.color_src, F32, "color_src"), p(self.color_rest, F32, "color_rest"), self.sh_degree,
                self.flags, p(self.viewmats, F32, "viewmats"), self.vm_stride, p(self.intrins, F32, "intrins"), 0.0,
                0.0, 0.0, 0.0, p(self.campos, F32, "campos"), self.h, self.w, self.bw,
                self.clip, p(self.packed, F32, "packed"), p(self.depths, F32, "depths"),
                p(self.radii, I32, "radii"), p(self.tiles_hit, I32, "num_tiles_hit"),
                p(tile_counts, I32, "tile_counts"), p(tile_keys, torch.int64, "tile_keys"), cap,
                self.st), "sb_views_project_fwd")

    def _wait_gt(self):
        if self.gt_ready is not None:
            torch.cuda.current_stream().wait_event(self.gt_ready)

    def raster_rect(self, view_has, work):
        lib, p = self.lib, _abi.ptr
        self._wait_gt()
        partial = self._loss_partial(8)
        with _stage("raster_fwd"):
            _abi.check(lib.sb_rasterize_rect_fwd(
                self.v, self.n, p(self.packed, F32, "packed"), p(self.rect_ids, I32, "rect_ids"),
                p(self.rect_meta, I32, "rect_meta"), p(self.background, F32, "background"),
                self.h, self.w, p(view_has, I32, "view_has_isects"),
                p(self.out_img, F32, "out_img"), p(self.final_ts, F32, "final_Ts"),
                p(self.final_idx, I32, "final_idx"), p(self.out_alpha, F32, "out_alpha"),
                1 if self.clamp_rgb else 0, p(self.gt, self.gt_dtype, "gt"), 1 if self.gt_dtype == torch.uint8 else 0,
                p(partial, F32, "loss_partial"), p(work, I32, "work_counter"), self.st),
                "sb_rasterize_rect_fwd")
        self._finish_loss(partial)

    def raster_tiles(self, view_has):
        lib, p = self.lib, _abi.ptr
        self._wait_gt()
        partial = self._loss_partial(1)
        with _stage("raster_fwd"):
            _abi.check(lib.sb_rasterize_fwd(
                self.v, self.n, p(self.packed, F32, "packed"), p(self.gids_sorted, I32, "gaussian_ids_sorted"),
                p(self.tile_bins, I32, "tile_bins"), p(self.background, F32, "background"), self.h,
                self.w, self.bw, p(view_has, I32, "view_has_isects"),
                p(self.out_img, F32, "out_img"), p(self.final_ts, F32, "final_Ts"),
                p(self.final_idx, I32, "final_idx"), p(self.out_alpha, F32, "out_alpha"),
                1 if self.clamp_rgb else 0, p(self.gt, self.gt_dtype, "gt"), 1 if self.gt_dtype == torch.uint8 else 0,
                p(partial, F32, "loss_partial"), self.st), "sb_rasterize_fwd")
        self._finish_loss(partial)

    # ---- the two binning strategies --------------------------------------------------------------
    def run_segments(self, cap: int, plan: RenderPlan) -> _Totals:
        """Fixed-capacity tile segments + per-tile sort (csrc/tile_sort.cu) + rect-list rasteriser, all
        enqueued without waiting for anything; returns the totals on their way to the host."""
        lib, p = self.lib, _abi.ptr
        dev, v, num_tiles = self.means.device, self.v, self.num_tiles
        self.cap = cap
        with _stage("zero_counts"):
            state = torch.zeros(v * num_tiles + v + 4, dtype=I32, device=dev)
        tile_counts, view_has = state[:v * num_tiles], state[v * num_tiles:-4]
        totals = state[-4:-1]           # intersections, overflow flags, rect-list pool head (= entries needed)
        work = state[-1:]               # rectangle queue head of the persistent raster kernel
        tile_keys = torch.empty(v * num_tiles * cap, dtype=torch.int64, device=dev)
        self.project(tile_counts, tile_keys, cap)
        self.alloc_images()
        self.gids_sorted = torch.empty(v * num_tiles * cap, dtype=I32, device=dev)
        self.tile_bins = torch.empty(v * num_tiles, 2, dtype=I32, device=dev)
        use_rect = self.bw == 16 and USE_RECT_LISTS
        pool_cap = 0
        if use_rect:
            # per-rectangle lists for the rect-list rasteriser (one warp per 8x4 rectangle), packed into one pool
            pool_cap = plan.pool_capacity((v, num_tiles, self.n), cap)
            self.rect_ids = torch.empty(pool_cap + 32, dtype=I32, device=dev)
            self.rect_meta = torch.empty(v * num_tiles * 8, 2, dtype=I32, device=dev)
        with _stage("sort_tiles"):
            _abi.check(lib.sb_sort_tiles(
                v, num_tiles, cap, p(tile_counts, I32, "tile_counts"),
                p(tile_keys, torch.int64, "tile_keys"), p(self.gids_sorted, I32, "gaussian_ids_sorted"),
                p(self.tile_bins, I32, "tile_bins"), p(view_has, I32, "view_has_isects"),
                p(totals, I32, "totals"), self.n, self.tiles_x, p(self.packed, F32, "packed"),
                p(self.rect_ids, I32, "rect_ids"), p(self.rect_meta, I32, "rect_meta"),
                p(totals[2:] if use_rect else None, I32, "pool_head"), pool_cap, self.st), "sb_sort_tiles")

        def landed(host_totals):
            host_totals.copy_(totals, non_blocking=True)
            ready = torch.cuda.Event()
            ready.record()
            return _Totals(host_totals, ready, (v, num_tiles, self.n), plan)
        if plan.graph_safe:
            # a slot of its own (a captured copy keeps writing to it on every replay), no event
            slot = plan.graph_slot()
            slot.copy_(totals, non_blocking=True)
            result = _Totals(slot, None, (v, num_tiles, self.n), plan)
            plan.graph_totals.append(result)
        else:
            result = plan.host_slot(landed)
        # the raster kernel is enqueued optimistically: the device produced the flag BEFORE it starts
        if use_rect:
            self.raster_rect(view_has, work)
        else:
            self.raster_tiles(view_has)
        return result

    def run_global(self, plan: RenderPlan, projected: bool) -> _Totals:
        """Very deep tiles / many tiles per Gaussian: depth pre-order of the Gaussians, emission in that
        order, radix sort over the tile bits only, per-tile rasteriser (one host read of the count)."""
        if not projected:
            self.project(None, None, 0)
        self.alloc_images()
        self.cap = 0
        self.rect_ids = self.rect_meta = None
        with _stage("global_sort"):
            num_intersects, _, self.gids_sorted, self.tile_bins, view_ends = utils.bin_depth_ordered(
                self.v, self.n, self.packed, 12, self.depths.view(-1), self.radii.view(-1), self.tiles_hit,
                (self.tiles_x, self.tiles_y, 1), self.bw, compact=True)
        view_has = (torch.diff(view_ends, prepend=view_ends.new_zeros(1)) > 0).to(I32)
        self.raster_tiles(view_has)
        return _Totals(None, None, (self.v, self.num_tiles, self.n), plan, value=(num_intersects, 0, 0))


def grad_layout(n: int, color_shape, rest_shape) -> Tuple[List[Tuple[str, int, Tuple[int, ...]]], int]:
    """Blocks of the flat per-Gaussian gradient buffer the backward writes, [(name, offset, shape)], and the
    buffer's size in floats.  One allocation, so that a view-sharded scene can all-reduce it in place
    (dist.SyncGaussianGrads)."""
    blocks, off = [], 0
    for name, shape in (("means", (n, 3)), ("scales", (n, 3)), ("quats", (n, 4)), ("opacities", (n, 1)),
                        ("color_src", tuple(color_shape)), ("color_rest", None if rest_shape is None else tuple(rest_shape))):
        if shape is None:
            continue
        size = 1
        for d in shape:
            size *= d
        blocks.append((name, off, shape))
        off += (size + 3) // 4 * 4           # every block starts 16-byte aligned
    return blocks, off


class _RenderViews(Function):
    """One autograd node for the whole scene.

    `flags` (see _abi.FLAG_*) tell the kernels which of the wrapper's activations they apply
    themselves; `color_rest` is features_rest when the SH coefficients arrive split as SplatFormer
    stores them.  With `gt` given the node returns the per-view L1 and squared-error sums next to the
    images (train.py:275,278) and the backward derives the L1 part of the image gradient inside the
    raster kernel; a gradient arriving on the returned images (LPIPS / SSIM terms, train.py:276-277)
    is added to it."""

    @staticmethod
    def forward(ctx, means, scales, quats, opacities, color_src, color_rest, viewmats, campos, intrins, gt,
                sh_degree, flags, img_height, img_width, block_width, background, clip_thresh, clamp_rgb, plan):
        global _last_plan
        dev = means.device
        with torch.cuda.device(dev):
            lib = _abi.load()
            st = _abi.stream()                      # one current-stream lookup per pass (it costs ~20 us)
            n = means.shape[0]
            v = viewmats.shape[0]
            opac_shape = tuple(opacities.shape)
            means, scales, quats = _c(means), _c(scales), _c(quats)
            opacities = _c(opacities).view(-1)
            color_src = _c(color_src)
            color_rest = None if color_rest is None else _c(color_rest)
            viewmats = _c(viewmats.to(F32))
            vm_stride = viewmats.shape[1] * viewmats.shape[2]
            campos = None if campos is None else _c(campos.to(F32))
            intrins = _c(intrins.to(F32))
            background = _c(background.to(F32))
            gt_ready = None
            if gt is not None:
                if gt.dtype not in (torch.uint8, F32):
                    gt = gt.to(F32)
                if not gt.is_cuda:
                    # targets in (pinned) host memory: the upload runs on a side stream underneath the
                    # projection / sort stages and is awaited right before the raster kernel needs it
                    side = _side_stream(dev)
                    with torch.cuda.stream(side):
                        gt = gt.contiguous().to(dev, non_blocking=True)
                        gt_ready = torch.cuda.Event()
                        gt_ready.record(side)
                    gt.record_stream(torch.cuda.current_stream())
                gt = _c(gt)
            meta = (v, n, sh_degree, flags, img_height, img_width, block_width, clip_thresh, vm_stride, clamp_rgb)
            ps = _Pass(lib, st, (means, scales, quats, opacities, color_src, color_rest, viewmats, campos, intrins,
                                 background), meta, gt, gt_ready)
            training = any(ctx.needs_input_grad[:6])       # False under torch.no_grad() / for constant inputs
            cap = plan.capacity(lib, v, ps.num_tiles, n)
            if cap:
                totals = ps.run_segments(cap, plan)
                key = (v, ps.num_tiles, n)
                first = key not in plan.seen
                plan.seen.add(key)
                if plan.graph_safe:
                    pass                    # nothing may be read here: plan.check() after the (replayed) step
                elif not training or first:
                    # evaluation (nothing will come back to repair an overflowed pass) or a shape this plan has
                    # not seen yet: look right away
                    if totals.read()[1]:
                        plan.replays += 1
                        totals = ps.run_global(plan, projected=True)
                else:
                    plan.pending.append(totals)
            else:
                totals = ps.run_global(plan, projected=False)
                totals.read()
            plan.last = totals
            _last_plan = plan
            if KEEP_LAST_PASS:
                global _last_pass
                _last_pass = ps
            ctx.ps, ctx.totals, ctx.plan = ps, totals, plan
            ctx.opac_shape = opac_shape
            ctx.color_shapes = (tuple(color_src.shape), None if color_rest is None else tuple(color_rest.shape))
            ctx.set_materialize_grads(False)
            # OUTPUTS must not hang off ctx as plain attributes: output -> grad_fn (= ctx) -> ps -> output is a
            # reference cycle that only Python's cyclic collector breaks, and until it runs every step's buffers
            # (> 1 GB) stay allocated.  save_for_backward handles outputs without the cycle.
            out_img, out_alpha = ps.out_img, ps.out_alpha
            ps.out_img = ps.out_alpha = None
            ctx.save_for_backward(out_img, out_alpha)
            if gt is not None:
                l1_sum, sq_sum = ps.loss_sums[:, 0], ps.loss_sums[:, 1]
                ctx.mark_non_differentiable(sq_sum)
                return l1_sum, sq_sum, out_img
            return out_img, out_alpha

    @staticmethod
    def backward(ctx, *grads):
        ps: _Pass = ctx.ps
        nin = 19
        dev = ps.means.device
        if ps.gt is not None:
            v_l1, _, v_out_img = grads
            v_out_alpha = None
            if v_l1 is None and v_out_img is None:
                return (None,) * nin
        else:
            v_out_img, v_out_alpha = grads
            v_l1 = None
            if v_out_img is None and v_out_alpha is None:
                return (None,) * nin
        ps.out_img, ps.out_alpha = ctx.saved_tensors
        try:
            return _RenderViews._backward(ctx, ps, dev, nin, v_l1, v_out_img, v_out_alpha)
        finally:
            ps.out_img = ps.out_alpha = None

    @staticmethod
    def _backward(ctx, ps, dev, nin, v_l1, v_out_img, v_out_alpha):
        with torch.cuda.device(dev):
            lib, p = _abi.load(), _abi.ptr
            st = _abi.stream()
            ps.st = st
            # the binning flag of the forward: the device wrote it before the forward rasteriser started
            overflow = 0 if ctx.plan.graph_safe else ctx.totals.read()[1]
            if overflow:
                # a tile ran over its segment: the lists the forward walked were truncated.  Redo the pass with
                # the global sort over the same buffers (images and loss sums are corrected in place), then
                # differentiate that.
                ctx.plan.replays += 1
                stale = ctx.totals
                ctx.totals = ps.run_global(ctx.plan, projected=True)
                if ctx.plan.last is stale:
                    ctx.plan.last = ctx.totals
            v, n = ps.v, ps.n
            v_loss = None
            if ps.gt is not None:
                v_loss = torch.zeros(v, dtype=F32, device=dev) if v_l1 is None else _c(v_l1.to(F32))
            v_out_img = None if v_out_img is None else _c(v_out_img.to(F32))
            v_out_alpha = None if v_out_alpha is None else _c(v_out_alpha.to(F32))
            with _stage("zero_grads"):
                v_packed = torch.zeros(v, n, 12, dtype=F32, device=dev)
            gt_dt = ps.gt_dtype
            out_img = ps.out_img if (ps.clamp_rgb or ps.gt is not None) else None
            if ps.rect_ids is not None:
                with _stage("raster_bwd"):
                    _abi.check(lib.sb_rasterize_rect_bwd(
                        v, n, p(ps.packed, F32, "packed"), p(ps.rect_ids, I32, "rect_ids"),
                        p(ps.rect_meta, I32, "rect_meta"), p(ps.background, F32, "background"),
                        ps.h, ps.w, p(ps.final_ts, F32, "final_Ts"),
                        p(ps.final_idx, I32, "final_idx"), p(v_out_img, F32, "v_out_img"),
                        p(v_out_alpha, F32, "v_out_alpha"), p(out_img, F32, "out_img"),
                        1 if ps.clamp_rgb else 0, p(ps.gt, gt_dt, "gt"), 1 if gt_dt == torch.uint8 else 0,
                        p(v_loss, F32, "v_loss"), p(v_packed, F32, "v_packed"), st),
                        "sb_rasterize_rect_bwd")
            else:
                with _stage("raster_bwd"):
                    _abi.check(lib.sb_rasterize_bwd(
                        v, n, p(ps.packed, F32, "packed"), p(ps.gids_sorted, I32, "gaussian_ids_sorted"),
                        p(ps.tile_bins, I32, "tile_bins"), p(ps.background, F32, "background"), ps.h,
                        ps.w, ps.bw, p(ps.final_ts, F32, "final_Ts"),
                        p(ps.final_idx, I32, "final_idx"), p(v_out_img, F32, "v_out_img"),
                        p(v_out_alpha, F32, "v_out_alpha"), p(out_img, F32, "out_img"),
                        1 if ps.clamp_rgb else 0, p(ps.gt, gt_dt, "gt"), 1 if gt_dt == torch.uint8 else 0,
                        p(v_loss, F32, "v_loss"), p(v_packed, F32, "v_packed"), st), "sb_rasterize_bwd")
            # all per-Gaussian gradients in ONE allocation (dist.SyncGaussianGrads all-reduces it in place)
            color_shape, rest_shape = ctx.color_shapes
            blocks, size = grad_layout(n, color_shape, rest_shape)
            flat = torch.empty(size, dtype=F32, device=dev)
            g = {}
            for name, off, shape in blocks:
                cnt = 1
                for x in shape:
                    cnt *= x
                g[name] = flat[off:off + cnt].view(shape)
            stats = _bwd_stats(dev) if COLLECT_BWD_STATS else None
            with _stage("project_sh_bwd"):
                _abi.check(lib.sb_views_project_bwd(
                    v, n, p(ps.means, F32, "means"), p(ps.scales, F32, "scales"), 1.0,
                    p(ps.quats, F32, "quats"), ps.sh_degree, ps.flags, p(ps.viewmats, F32, "viewmats"), ps.vm_stride,
                    p(ps.intrins, F32, "intrins"), 0.0, 0.0, 0.0, 0.0, p(ps.campos, F32, "campos"),
                    ps.h, ps.w, ps.bw, ps.clip, p(ps.packed, F32, "packed"),
                    p(v_packed, F32, "v_packed"), p(g["means"], F32, "v_means"),
                    p(g["scales"], F32, "v_scales"), p(g["quats"], F32, "v_quats"),
                    p(g["opacities"], F32, "v_opacities"), p(g["color_src"], F32, "v_color_src"),
                    p(g.get("color_rest"), F32, "v_color_rest"), p(stats, I32, "stats"), st), "sb_views_project_bwd")
            return (g["means"], g["scales"], g["quats"], g["opacities"].view(ctx.opac_shape), g["color_src"],
                    g.get("color_rest")) + (None,) * (nin - 6)


# Debug counter of the projection backward: (records scanned, records with a non-zero gradient), summed
# over the calls made while COLLECT_BWD_STATS is on.  Off by default (two extra atomics per warp).
COLLECT_BWD_STATS = False
_bwd_stats_buf: Dict[int, Tensor] = {}


def _bwd_stats(device) -> Tensor:
    idx = device.index if device.index is not None else torch.cuda.current_device()
    if idx not in _bwd_stats_buf:
        _bwd_stats_buf[idx] = torch.zeros(2, dtype=I32, device=device)
    return _bwd_stats_buf[idx]


def bwd_stats(device=None, reset: bool = True) -> Tuple[int, int]:
    """(records scanned, records with a non-zero gradient) since the last reset (COLLECT_BWD_STATS)."""
    device = torch.device("cuda", torch.cuda.current_device()) if device is None else device
    buf = _bwd_stats(device)
    a, b = (int(x) for x in buf.tolist())
    if reset:
        buf.zero_()
    return a, b


def _campos_from_viewmats(viewmats: Tensor) -> Tensor:
    """Camera centres -R^T t of (V,3|4,4) world->camera matrices."""
    rot, t = viewmats[:, :3, :3].to(F32), viewmats[:, :3, 3].to(F32)
    return -(rot.transpose(1, 2) * t[:, None, :]).sum(dim=2)


def render_views(means: Tensor, scales: Tensor, quats: Tensor, opacities: Tensor, color_src: Tensor,
                 sh_degree: int, viewmats: Tensor, campos: Optional[Tensor], intrins: Tensor,
                 img_height: int, img_width: int, block_width: int = 16,
                 background: Optional[Tensor] = None, clip_thresh: float = 0.01, clamp_rgb: bool = False,
                 color_rest: Optional[Tensor] = None, flags: int = 0, gt: Optional[Tensor] = None,
                 plan: Optional[RenderPlan] = None):
    """Render all V views of one scene.

    means (N,3), scales (N,3), quats (N,4), opacities (N,1|N); activated values unless `flags`
    (_abi.FLAG_RAW_SCALES | FLAG_RAW_OPACITIES | FLAG_RAW_QUATS | FLAG_COLOR_SIGMOID) say the kernels
    apply exp / sigmoid / the wrapper's quaternion normalisation themselves.  color_src: (N,K,3) SH
    coefficients with sh_degree in 0..4 (colour = max(SH + 0.5, 0) as gs_utils.py:78-79; or features_dc
    (N,3) with color_rest = features_rest (N,K-1,3)), or (N,3) rgb with sh_degree = -1.  viewmats
    (V,3|4,4) world->camera (campos (V,3) = camera centres, derived from the matrices when omitted), or
    OpenGL camera-to-world matrices with FLAG_C2W_OPENGL; intrins (V,4) = fx, fy, cx, cy, or (1,4) shared
    by all views with FLAG_INTRINS_SHARED.

    Returns rgb (V,H,W,3) and alpha (V,H,W); rgb is not clamped unless clamp_rgb=True, which folds the
    wrapper's torch.clamp(rgb, max=1) (gs_utils.py:111) into the kernels.  With gt (V,H,W,3) given it
    returns instead (l1_sum (V,), sq_sum (V,), rgb): the per-view sums of |rgb - gt| (differentiable)
    and (rgb - gt)^2 (for PSNR, not differentiable) computed inside the raster kernel, and the images,
    which are differentiable too (an image-space loss such as LPIPS can be added on top of the fused L1).
    Differentiable in means, scales, quats, opacities, color_src and color_rest.
    `plan` keeps what the call site has learnt about its scenes' binning (RenderPlan)."""
    assert 2 <= block_width <= 16, "block_width must be between 2 and 16"
    if not means.is_cuda:
        raise RuntimeError("render_views needs CUDA tensors (there is no CPU path)")
    if background is None:
        background = torch.ones(3, dtype=F32, device=means.device)
    want = (1, 4) if flags & _abi.FLAG_INTRINS_SHARED else (viewmats.shape[0], 4)
    if tuple(intrins.shape) != want:
        raise ValueError(f"intrins must be {want}, got {tuple(intrins.shape)}")
    if campos is None and not (flags & _abi.FLAG_C2W_OPENGL):
        # SH view directions need the camera centres; sh_degree = -1 never reads them
        campos = _campos_from_viewmats(viewmats) if sh_degree >= 0 else \
            torch.zeros(viewmats.shape[0], 3, dtype=F32, device=means.device)
    return _RenderViews.apply(means, scales, quats, opacities, color_src, color_rest, viewmats, campos, intrins,
                              gt, int(sh_degree), int(flags), int(img_height), int(img_width), int(block_width),
                              background, float(clip_thresh), bool(clamp_rgb), plan if plan is not None else _default_plan)
