"""Tiling helpers with the names and return values of gsplat.utils in v0.1.11, which
`gsplat.rasterize_gaussians` (call site /root/reference/utils/gs_utils.py:96-109) is built from.
Each one is a thin wrapper over the C ABI (include/splat_b200.h rows a5-a9)."""
from __future__ import annotations

from typing import Tuple

import torch
from torch import Tensor

from . import _abi


def _scratch(nbytes: int, device) -> Tensor:
    return torch.empty(max(int(nbytes), 1), dtype=torch.uint8, device=device)


def compute_cov2d_bounds(cov2d: Tensor) -> Tuple[Tensor, Tensor]:
    """gsplat.utils.compute_cov2d_bounds: (N,3) upper triangles of 2x2 covariances -> conics (N,3)
    (the inverses) and radii (N,) int32 = ceil(3 sqrt(largest eigenvalue)).  Not differentiable,
    as upstream."""
    if cov2d.shape[-1] != 3:
        raise ValueError("cov2d must have shape (N, 3)")
    if not cov2d.is_cuda:
        raise RuntimeError("compute_cov2d_bounds needs a CUDA tensor (there is no CPU path)")
    lib = _abi.load()
    flat = cov2d.detach().to(torch.float32).contiguous().view(-1, 3)
    n = flat.shape[0]
    conics = torch.empty(n, 3, dtype=torch.float32, device=flat.device)
    radii = torch.empty(n, dtype=torch.int32, device=flat.device)
    _abi.check(lib.sb_cov2d_bounds(n, _abi.ptr(flat, torch.float32, "cov2d"), _abi.ptr(conics, torch.float32, "conics"),
                                   _abi.ptr(radii, torch.int32, "radii"), _abi.stream()), "sb_cov2d_bounds")
    return conics, radii


def compute_cumulative_intersects(num_tiles_hit: Tensor, sync: bool = True) -> Tuple[int, Tensor]:
    """Inclusive int32 prefix sum of the per-Gaussian tile counts and its last element.
    The `.item()` is the one host synchronisation upstream has here as well (sync=False skips it and
    returns -1 for callers that already know the count)."""
    lib = _abi.load()
    flat = num_tiles_hit.contiguous().view(-1)
    cum = torch.empty_like(flat)
    n = flat.numel()
    if n == 0:
        return 0, cum
    tmp = _scratch(lib.sb_cumsum_temp_bytes(n), flat.device)
    _abi.check(lib.sb_cumsum_i32(n, _abi.ptr(flat, torch.int32, "num_tiles_hit"),
                                 _abi.ptr(cum, torch.int32, "cum_tiles_hit"),
                                 _abi.ptr(tmp, torch.uint8, "temp"), tmp.numel(), _abi.stream()),
               "sb_cumsum_i32")
    return (int(cum[-1].item()) if sync else -1), cum


def map_gaussian_to_intersects(num_points: int, num_intersects: int, xys: Tensor, depths: Tensor,
                               radii: Tensor, cum_tiles_hit: Tensor, tile_bounds, block_width: int,
                               num_views: int = 1, xys_stride: int = 2) -> Tuple[Tensor, Tensor]:
    """`xys_stride` = 12 lets the fused path pass its packed raster record (whose first two floats are
    the pixel centre) instead of a dense (V*N, 2) tensor."""
    lib = _abi.load()
    dev = xys.device
    isect_ids = torch.empty(num_intersects, dtype=torch.int64, device=dev)
    gaussian_ids = torch.empty(num_intersects, dtype=torch.int32, device=dev)
    _abi.check(lib.sb_map_gaussian_to_intersects(
        num_views, num_points, _abi.ptr(xys.contiguous(), torch.float32, "xys"), xys_stride,
        _abi.ptr(depths.contiguous(), torch.float32, "depths"),
        _abi.ptr(radii.contiguous(), torch.int32, "radii"),
        _abi.ptr(cum_tiles_hit.contiguous(), torch.int32, "cum_tiles_hit"),
        int(tile_bounds[0]), int(tile_bounds[1]), block_width,
        _abi.ptr(isect_ids, torch.int64, "isect_ids"), _abi.ptr(gaussian_ids, torch.int32, "gaussian_ids"),
        _abi.stream()), "sb_map_gaussian_to_intersects")
    return isect_ids, gaussian_ids


def sort_isects(isect_ids: Tensor, gaussian_ids: Tensor, num_tiles_total: int) -> Tuple[Tensor, Tensor]:
    """Ascending stable sort of the 64-bit (tile|depth) keys carrying the Gaussian ids."""
    lib = _abi.load()
    n = isect_ids.numel()
    keys_out = torch.empty_like(isect_ids)
    vals_out = torch.empty_like(gaussian_ids)
    if n == 0:
        return keys_out, vals_out
    end_bit = 32 + max(1, (max(num_tiles_total, 1) - 1).bit_length())
    tmp = _scratch(lib.sb_sort_temp_bytes(n), isect_ids.device)
    _abi.check(lib.sb_sort_pairs(n, end_bit, _abi.ptr(isect_ids, torch.int64, "isect_ids"),
                                 _abi.ptr(gaussian_ids, torch.int32, "gaussian_ids"),
                                 _abi.ptr(keys_out, torch.int64, "isect_ids_sorted"),
                                 _abi.ptr(vals_out, torch.int32, "gaussian_ids_sorted"),
                                 _abi.ptr(tmp, torch.uint8, "temp"), tmp.numel(), _abi.stream()),
               "sb_sort_pairs")
    return keys_out, vals_out


def sort_pairs32(keys: Tensor, vals: Tensor, end_bit: int = 32) -> Tuple[Tensor, Tensor]:
    """Stable sort of (uint32 key stored as int32, int32 value) pairs on the low `end_bit` key bits."""
    lib = _abi.load()
    n = keys.numel()
    keys_out, vals_out = torch.empty_like(keys), torch.empty_like(vals)
    if n == 0:
        return keys_out, vals_out
    tmp = _scratch(lib.sb_sort32_temp_bytes(n), keys.device)
    _abi.check(lib.sb_sort_pairs32(n, end_bit, _abi.ptr(keys, torch.int32, "keys"), _abi.ptr(vals, torch.int32, "vals"),
                                   _abi.ptr(keys_out, torch.int32, "keys_out"), _abi.ptr(vals_out, torch.int32, "vals_out"),
                                   _abi.ptr(tmp, torch.uint8, "temp"), tmp.numel(), _abi.stream()), "sb_sort_pairs32")
    return keys_out, vals_out


def sort_tile_pairs(tile_ids: Tensor, gaussian_ids: Tensor, num_tiles_total: int) -> Tuple[Tensor, Tensor]:
    """The fused path's tile-index sort on its own: (tile index, id) pairs -> ids sorted stably by tile, and
    tile_bins (num_tiles_total, 2).  The sort reads the keys once for its digit histograms here (inside
    bin_depth_ordered the emission kernel supplies them)."""
    lib = _abi.load()
    n = tile_ids.numel()
    dev = tile_ids.device
    out = torch.empty_like(gaussian_ids)
    bins = torch.empty(num_tiles_total, 2, dtype=torch.int32, device=dev)
    tmp = _scratch(lib.sb_sort_tiles_global_temp_bytes(n, num_tiles_total), dev)
    _abi.check(lib.sb_sort_tiles_global(n, num_tiles_total, _abi.ptr(tile_ids, torch.int32, "tile_ids"),
                                        _abi.ptr(gaussian_ids, torch.int32, "gaussian_ids"),
                                        _abi.ptr(out, torch.int32, "gaussian_ids_sorted"),
                                        _abi.ptr(bins, torch.int32, "tile_bins"), None,
                                        _abi.ptr(tmp, torch.uint8, "temp"), tmp.numel(), _abi.stream()),
               "sb_sort_tiles_global")
    return out, bins


def get_tile_bin_edges(num_intersects: int, isect_ids_sorted: Tensor, tile_bounds,
                       num_views: int = 1) -> Tensor:
    lib = _abi.load()
    tiles = int(tile_bounds[0]) * int(tile_bounds[1]) * num_views
    tile_bins = torch.zeros(tiles, 2, dtype=torch.int32, device=isect_ids_sorted.device)
    _abi.check(lib.sb_tile_bin_edges(num_intersects, _abi.ptr(isect_ids_sorted, torch.int64, "isect_ids_sorted"),
                                     tiles, _abi.ptr(tile_bins, torch.int32, "tile_bins"), _abi.stream()),
               "sb_tile_bin_edges")
    return tile_bins


def bin_and_sort_gaussians(num_points: int, num_intersects: int, xys: Tensor, depths: Tensor,
                           radii: Tensor, cum_tiles_hit: Tensor, tile_bounds, block_width: int,
                           num_views: int = 1):
    """-> (isect_ids_unsorted, gaussian_ids_unsorted, isect_ids_sorted, gaussian_ids_sorted, tile_bins)"""
    isect_ids, gaussian_ids = map_gaussian_to_intersects(
        num_points, num_intersects, xys, depths, radii, cum_tiles_hit, tile_bounds, block_width, num_views)
    tiles = int(tile_bounds[0]) * int(tile_bounds[1]) * num_views
    isect_ids_sorted, gaussian_ids_sorted = sort_isects(isect_ids, gaussian_ids, tiles)
    tile_bins = get_tile_bin_edges(num_intersects, isect_ids_sorted, tile_bounds, num_views)
    return isect_ids, gaussian_ids, isect_ids_sorted, gaussian_ids_sorted, tile_bins


def bin_depth_ordered(num_views: int, n: int, xys: Tensor, xys_stride: int, depths: Tensor, radii: Tensor,
                      num_tiles_hit: Tensor, tile_bounds, block_width: int, compact: bool = False):
    """The fused path's global sort (render.py): same per-tile lists as bin_and_sort_gaussians, built as
    depth pre-order of the V*N Gaussians (a small sort) -> emission in that order -> stable radix sort
    over the TILE bits only.  The emitted keys are already ascending in depth inside every tile, so the
    big sort needs 2 radix passes over bits [32, 32 + bits(V*T)) instead of 6 over the whole key.
    compact=True carries only the uint32 tile index through the big sort (8 bytes per pair instead of 12),
    whose last sweep keeps just the sorted ids and counts the pairs per tile (no sorted keys are returned:
    None in their place).
    Returns (num_intersects, isect_ids_sorted, gaussian_ids_sorted, tile_bins, last cumulative count per view)."""
    lib = _abi.load()
    dev = depths.device
    total = num_views * n
    I64, I32, F32 = torch.int64, torch.int32, torch.float32
    dkeys = torch.empty(total, dtype=I64, device=dev)
    index = torch.empty(total, dtype=I32, device=dev)
    _abi.check(lib.sb_depth_order_keys(num_views, n, _abi.ptr(depths, F32, "depths"), _abi.ptr(radii, I32, "radii"),
                                       _abi.ptr(dkeys, I64, "depth_keys"), _abi.ptr(index, I32, "index"),
                                       _abi.stream()), "sb_depth_order_keys")
    dkeys_sorted, order = torch.empty_like(dkeys), torch.empty_like(index)
    tmp = _scratch(lib.sb_sort_temp_bytes(total), dev)
    view_bits = max(1, (max(num_views, 1) - 1).bit_length())
    _abi.check(lib.sb_sort_pairs(total, 32 + view_bits, _abi.ptr(dkeys, I64, "depth_keys"),
                                 _abi.ptr(index, I32, "index"), _abi.ptr(dkeys_sorted, I64, "depth_keys_sorted"),
                                 _abi.ptr(order, I32, "order"), _abi.ptr(tmp, torch.uint8, "temp"), tmp.numel(),
                                 _abi.stream()), "sb_sort_pairs")
    del dkeys, dkeys_sorted, index
    hit_ordered = torch.empty(total, dtype=I32, device=dev)
    _abi.check(lib.sb_gather_i32(total, _abi.ptr(num_tiles_hit.view(-1), I32, "num_tiles_hit"),
                                 _abi.ptr(order, I32, "order"), _abi.ptr(hit_ordered, I32, "hit_ordered"),
                                 _abi.stream()), "sb_gather_i32")
    num_intersects, cum = compute_cumulative_intersects(hit_ordered)
    tiles_x, tiles_y = int(tile_bounds[0]), int(tile_bounds[1])
    tiles_total = tiles_x * tiles_y * num_views
    tile_bits = max(1, (max(tiles_total, 1) - 1).bit_length())
    kdt = torch.int32 if compact else I64              # int32 storage for the uint32 tile indices
    isect_ids = torch.empty(num_intersects, dtype=kdt, device=dev)
    gaussian_ids = torch.empty(num_intersects, dtype=I32, device=dev)
    vals_out = torch.empty_like(gaussian_ids)
    # emission is view-major (the view is the top of the pre-order key): offsets at the view boundaries
    view_ends = cum.view(num_views, n)[:, -1] if total > 0 else cum
    if compact:
        # tile-index sort that keeps only the sorted ids: the last sweep counts the pairs per tile, a single-pass
        # scan turns the counts into bins
        tile_bins = torch.empty(tiles_total, 2, dtype=I32, device=dev)
        if num_intersects > 0:
            _abi.check(lib.sb_map_gaussian_to_tiles_ordered(
                num_views, n, _abi.ptr(xys, F32, "xys"), xys_stride, _abi.ptr(depths, F32, "depths"),
                _abi.ptr(radii, I32, "radii"), _abi.ptr(order, I32, "order"), _abi.ptr(cum, I32, "cum_tiles_hit"),
                tiles_x, tiles_y, block_width, _abi.ptr(isect_ids, kdt, "tile_ids"),
                _abi.ptr(gaussian_ids, I32, "gaussian_ids"), _abi.stream()), "sb_map_gaussian_to_tiles_ordered")
        # (the sort takes its digit histograms with one read of the keys: counting them during the emission was
        #  measured slower -- 2.2 ms of shared atomics at 376 M pairs against a 0.5 ms histogram kernel)
        tmp = _scratch(lib.sb_sort_tiles_global_temp_bytes(num_intersects, tiles_total), dev)
        _abi.check(lib.sb_sort_tiles_global(
            num_intersects, tiles_total, _abi.ptr(isect_ids, kdt, "tile_ids"),
            _abi.ptr(gaussian_ids, I32, "gaussian_ids"), _abi.ptr(vals_out, I32, "gaussian_ids_sorted"),
            _abi.ptr(tile_bins, I32, "tile_bins"), None,
            _abi.ptr(tmp, torch.uint8, "temp"), tmp.numel(), _abi.stream()), "sb_sort_tiles_global")
        return num_intersects, None, vals_out, tile_bins, view_ends
    keys_out = torch.empty_like(isect_ids)
    if num_intersects > 0:
        _abi.check(lib.sb_map_gaussian_to_intersects_ordered(
            num_views, n, _abi.ptr(xys, F32, "xys"), xys_stride, _abi.ptr(depths, F32, "depths"),
            _abi.ptr(radii, I32, "radii"), _abi.ptr(order, I32, "order"), _abi.ptr(cum, I32, "cum_tiles_hit"),
            tiles_x, tiles_y, block_width, _abi.ptr(isect_ids, kdt, "isect_ids"),
            _abi.ptr(gaussian_ids, I32, "gaussian_ids"), _abi.stream()), "ordered key emission")
        tmp = _scratch(lib.sb_sort_temp_bytes(num_intersects), dev)
        _abi.check(lib.sb_sort_pairs_range(num_intersects, 32, 32 + tile_bits,
                                           _abi.ptr(isect_ids, I64, "isect_ids"),
                                           _abi.ptr(gaussian_ids, I32, "gaussian_ids"),
                                           _abi.ptr(keys_out, I64, "isect_ids_sorted"),
                                           _abi.ptr(vals_out, I32, "gaussian_ids_sorted"),
                                           _abi.ptr(tmp, torch.uint8, "temp"), tmp.numel(), _abi.stream()),
                   "sb_sort_pairs_range")
    del isect_ids, gaussian_ids
    tile_bins = get_tile_bin_edges(num_intersects, keys_out, tile_bounds, num_views)
    return num_intersects, keys_out, vals_out, tile_bins, view_ends
