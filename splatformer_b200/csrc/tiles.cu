// Tiling: intersection offsets, key emission, key sort, bin edges (upstream
// gsplat.utils.{compute_cumulative_intersects, map_gaussian_to_intersects, bin_and_sort_gaussians,
// get_tile_bin_edges}, reached from gsplat.rasterize_gaussians at
// /root/reference/utils/gs_utils.py:96-109).  All integer work, bit-exact by construction.
// The sort and the prefix sum themselves live in radix_sort.cu.
#include "sb_common.cuh"

namespace sb {

// One lane per (view, Gaussian).  A Gaussian covering a few tiles writes its own keys; one covering many
// (large screen radii: tens to hundreds of tiles) is handed to the whole warp, lane = tile, so that its
// keys and ids leave as coalesced 256-byte / 128-byte stores instead of one 8-byte store per lane and
// iteration, each to a different sector.
constexpr int kCoopTiles = 8;

// COMPACT: the key is the 32-bit tile index alone (depth-ordered emission: the depth is already encoded in
// the position), stored through the same pointer reinterpreted as uint32.
template <bool COMPACT>
__global__ void __launch_bounds__(256) map_isect_kernel(
    int num_views, int n, const float* __restrict__ xys, int xy_stride, const float* __restrict__ depths,
    const int* __restrict__ radii, const int* __restrict__ cum, int tiles_x, int tiles_y, int bw,
    int64_t* __restrict__ keys, int* __restrict__ gids, const int* __restrict__ order) {
  uint32_t* keys32 = reinterpret_cast<uint32_t*>(keys);
  // slot j of the emission order handles (view, Gaussian) index i = order[j] (identity without `order`);
  // cum is the prefix sum of the tile counts in emission order
  const size_t j = (size_t)blockIdx.x * 256 + threadIdx.x;
  const bool in_range = j < (size_t)num_views * n;
  const size_t i = (order && in_range) ? (size_t)__ldg(order + j) : j;
  const int lane = threadIdx.x & 31;
  int min_x = 0, min_y = 0, nx = 0, total = 0, g = 0;
  size_t k = 0;
  int64_t depth_bits = 0, tile0 = 0;
  if (in_range) {
    const int r = __ldg(radii + i);
    if (r > 0) {
      const int v = (int)(i / n);
      g = (int)(i - (size_t)v * n);
      const float2 c = __ldg(reinterpret_cast<const float2*>(xys + i * xy_stride));
      int max_x, max_y;
      tile_bbox(c.x, c.y, (float)r, tiles_x, tiles_y, bw, min_x, max_x, min_y, max_y);
      nx = max_x - min_x;
      total = nx * (max_y - min_y);
      k = j == 0 ? 0 : (size_t)__ldg(cum + j - 1);
      depth_bits = (int64_t)(uint32_t)__float_as_int(__ldg(depths + i));
      tile0 = (int64_t)v * tiles_x * tiles_y;
    }
  }
  if (total > 0 && total <= kCoopTiles) {
    for (int e = 0; e < total; ++e) {
      const int ky = e / nx, kx = e - ky * nx;
      const int64_t t = tile0 + (int64_t)(min_y + ky) * tiles_x + (min_x + kx);
      if (COMPACT) keys32[k + e] = (uint32_t)t; else keys[k + e] = (t << 32) | depth_bits;
      gids[k + e] = g;
    }
  }
  unsigned big = __ballot_sync(0xffffffffu, total > kCoopTiles);
  while (big) {
    const int src = __ffs(big) - 1;
    big &= big - 1;
    const int s_total = __shfl_sync(0xffffffffu, total, src), s_nx = __shfl_sync(0xffffffffu, nx, src);
    const int s_min_x = __shfl_sync(0xffffffffu, min_x, src), s_min_y = __shfl_sync(0xffffffffu, min_y, src);
    const int s_g = __shfl_sync(0xffffffffu, g, src);
    const size_t s_k = __shfl_sync(0xffffffffu, (unsigned long long)k, src);
    const int64_t s_depth = __shfl_sync(0xffffffffu, (long long)depth_bits, src);
    const int64_t s_tile0 = __shfl_sync(0xffffffffu, (long long)tile0, src);
    for (int e0 = 0; e0 < s_total; e0 += 32) {
      const int e = e0 + lane;
      const bool on = e < s_total;
      const int ky = e / s_nx, kx = e - ky * s_nx;
      const int64_t t = s_tile0 + (int64_t)(s_min_y + ky) * tiles_x + (s_min_x + kx);
      if (on) {
        if (COMPACT) keys32[s_k + e] = (uint32_t)t; else keys[s_k + e] = (t << 32) | s_depth;
        gids[s_k + e] = s_g;
      }
    }
  }
}

// Binning of already projected Gaussians into the fixed-capacity tile segments of tile_sort.cu (what the fused
// projection kernel does on the fly): one thread per (view, Gaussian) claims a slot in every tile of its bounding
// box and drops (depth bits << 32 | id) there.  Lets gsplat.rasterize_gaussians -- which receives xys / depths /
// radii from an earlier project_gaussians call -- use the per-tile sort and the rect-list rasterisers instead of
// cumsum + host read + emission + 64-bit global sort + bin edges.
__global__ void __launch_bounds__(256) bin_segments_kernel(int num_views, int n, const float* __restrict__ xys,
                                                           int xy_stride, const float* __restrict__ depths,
                                                           const int* __restrict__ radii, int tiles_x, int tiles_y,
                                                           int bw, int* __restrict__ tile_counts,
                                                           unsigned long long* __restrict__ tile_keys, int cap) {
  const size_t i = (size_t)blockIdx.x * 256 + threadIdx.x;
  if (i >= (size_t)num_views * n) return;
  const int r = __ldg(radii + i);
  if (r <= 0) return;
  const int v = (int)(i / n), g = (int)(i - (size_t)v * n);
  const float2 c = __ldg(reinterpret_cast<const float2*>(xys + i * xy_stride));
  int min_x, max_x, min_y, max_y;
  tile_bbox(c.x, c.y, (float)r, tiles_x, tiles_y, bw, min_x, max_x, min_y, max_y);
  const unsigned long long key = ((unsigned long long)(unsigned)__float_as_int(__ldg(depths + i)) << 32) | (unsigned)g;
  const size_t t0 = (size_t)v * tiles_x * tiles_y;
  for (int ty = min_y; ty < max_y; ++ty)
    for (int tx = min_x; tx < max_x; ++tx) {
      const size_t t = t0 + (size_t)ty * tiles_x + tx;
      const int pos = atomicAdd(tile_counts + t, 1);
      if (pos < cap) tile_keys[t * cap + pos] = key;
    }
}

// Depth pre-order of the fused path's global sort (tiles.cu: sb_depth_order_keys): key = view << 32 |
// float32 bits of the depth (culled Gaussians: all ones in the low word -- they emit nothing, their place
// does not matter), value = flat (view, Gaussian) index.
__global__ void __launch_bounds__(256) depth_keys_kernel(int num_views, int n, const float* __restrict__ depths,
                                                         const int* __restrict__ radii, int64_t* __restrict__ keys,
                                                         int* __restrict__ index) {
  const size_t i = (size_t)blockIdx.x * 256 + threadIdx.x;
  if (i >= (size_t)num_views * n) return;
  const int64_t v = (int64_t)(i / n);
  const uint32_t bits = __ldg(radii + i) > 0 ? (uint32_t)__float_as_int(__ldg(depths + i)) : 0xffffffffu;
  keys[i] = (v << 32) | (int64_t)bits;
  index[i] = (int)i;
}

__global__ void __launch_bounds__(256) gather_i32_kernel(int count, const int* __restrict__ src,
                                                         const int* __restrict__ index, int* __restrict__ dst) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i < count) dst[i] = __ldg(src + __ldg(index + i));
}

__global__ void __launch_bounds__(256) bin_edges_kernel(int num_isect, const int64_t* __restrict__ keys,
                                                        int2* __restrict__ bins) {
  const int idx = blockIdx.x * 256 + threadIdx.x;
  if (idx >= num_isect) return;
  const int cur = (int)(__ldg(keys + idx) >> 32);
  if (idx == 0) bins[cur].x = 0;
  else {
    const int prev = (int)(__ldg(keys + idx - 1) >> 32);
    if (prev != cur) { bins[prev].y = idx; bins[cur].x = idx; }
  }
  if (idx == num_isect - 1) bins[cur].y = num_isect;
}

__global__ void __launch_bounds__(256) bin_edges32_kernel(int count, const uint32_t* __restrict__ tiles,
                                                          int2* __restrict__ bins) {
  const int idx = blockIdx.x * 256 + threadIdx.x;
  if (idx >= count) return;
  const int cur = (int)__ldg(tiles + idx);
  if (idx == 0) bins[cur].x = 0;
  else {
    const int prev = (int)__ldg(tiles + idx - 1);
    if (prev != cur) { bins[prev].y = idx; bins[cur].x = idx; }
  }
  if (idx == count - 1) bins[cur].y = count;
}

}  // namespace sb

extern "C" int sb_map_gaussian_to_intersects(int num_views, int n, const float* xys, int xys_stride,
                                             const float* depths, const int* radii,
                                             const int* cum_tiles_hit, int tiles_x, int tiles_y,
                                             int block_width, int64_t* isect_ids, int* gaussian_ids,
                                             sb_stream_t stream) {
  if (n < 0 || num_views < 1 || block_width < 2 || block_width > 16 || xys_stride < 2 || (xys_stride & 1))
    return SB_ERR_BAD_ARG;
  const size_t total = (size_t)num_views * n;
  if (total == 0) return 0;
  sb::map_isect_kernel<false><<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      num_views, n, xys, xys_stride, depths, radii, cum_tiles_hit, tiles_x, tiles_y, block_width, isect_ids,
      gaussian_ids, nullptr);
  SB_RETURN_LAST();
}

extern "C" int sb_bin_to_tile_segments(int num_views, int n, const float* xys, int xys_stride, const float* depths,
                                       const int* radii, int tiles_x, int tiles_y, int block_width,
                                       int* tile_counts, uint64_t* tile_keys, int tile_capacity, sb_stream_t stream) {
  if (n < 0 || num_views < 1 || block_width < 2 || block_width > 16 || xys_stride < 2 || (xys_stride & 1) ||
      !tile_counts || !tile_keys || tile_capacity < 1)
    return SB_ERR_BAD_ARG;
  const size_t total = (size_t)num_views * n;
  if (total == 0) return 0;
  sb::bin_segments_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      num_views, n, xys, xys_stride, depths, radii, tiles_x, tiles_y, block_width, tile_counts,
      reinterpret_cast<unsigned long long*>(tile_keys), tile_capacity);
  SB_RETURN_LAST();
}

extern "C" int sb_map_gaussian_to_intersects_ordered(int num_views, int n, const float* xys, int xys_stride,
                                                     const float* depths, const int* radii, const int* order,
                                                     const int* cum_tiles_hit_ordered, int tiles_x, int tiles_y,
                                                     int block_width, int64_t* isect_ids, int* gaussian_ids,
                                                     sb_stream_t stream) {
  if (n < 0 || num_views < 1 || block_width < 2 || block_width > 16 || xys_stride < 2 || (xys_stride & 1) || !order)
    return SB_ERR_BAD_ARG;
  const size_t total = (size_t)num_views * n;
  if (total == 0) return 0;
  sb::map_isect_kernel<false><<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      num_views, n, xys, xys_stride, depths, radii, cum_tiles_hit_ordered, tiles_x, tiles_y, block_width, isect_ids,
      gaussian_ids, order);
  SB_RETURN_LAST();
}

extern "C" int sb_map_gaussian_to_tiles_ordered(int num_views, int n, const float* xys, int xys_stride,
                                                const float* depths, const int* radii, const int* order,
                                                const int* cum_tiles_hit_ordered, int tiles_x, int tiles_y,
                                                int block_width, uint32_t* tile_ids, int* gaussian_ids,
                                                sb_stream_t stream) {
  if (n < 0 || num_views < 1 || block_width < 2 || block_width > 16 || xys_stride < 2 || (xys_stride & 1) || !order)
    return SB_ERR_BAD_ARG;
  const size_t total = (size_t)num_views * n;
  if (total == 0) return 0;
  sb::map_isect_kernel<true><<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      num_views, n, xys, xys_stride, depths, radii, cum_tiles_hit_ordered, tiles_x, tiles_y, block_width,
      reinterpret_cast<int64_t*>(tile_ids), gaussian_ids, order);
  SB_RETURN_LAST();
}

extern "C" int sb_tile_bin_edges32(int count, const uint32_t* tile_ids_sorted, int num_tiles_total, int* tile_bins,
                                   sb_stream_t stream) {
  if (count < 0 || num_tiles_total < 0) return SB_ERR_BAD_ARG;
  if (count == 0) return 0;
  sb::bin_edges32_kernel<<<sb::ceil_div(count, 256), 256, 0, (cudaStream_t)stream>>>(
      count, tile_ids_sorted, reinterpret_cast<int2*>(tile_bins));
  SB_RETURN_LAST();
}

extern "C" int sb_depth_order_keys(int num_views, int n, const float* depths, const int* radii, int64_t* keys,
                                   int* index, sb_stream_t stream) {
  if (n < 0 || num_views < 1) return SB_ERR_BAD_ARG;
  const size_t total = (size_t)num_views * n;
  if (total == 0) return 0;
  sb::depth_keys_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(num_views, n, depths,
                                                                                          radii, keys, index);
  SB_RETURN_LAST();
}

extern "C" int sb_gather_i32(int count, const int* src, const int* index, int* dst, sb_stream_t stream) {
  if (count < 0) return SB_ERR_BAD_ARG;
  if (count == 0) return 0;
  sb::gather_i32_kernel<<<sb::ceil_div(count, 256), 256, 0, (cudaStream_t)stream>>>(count, src, index, dst);
  SB_RETURN_LAST();
}

extern "C" int sb_tile_bin_edges(int num_intersects, const int64_t* isect_ids_sorted,
                                 int num_tiles_total, int* tile_bins, sb_stream_t stream) {
  if (num_intersects < 0 || num_tiles_total < 0) return SB_ERR_BAD_ARG;
  if (num_intersects == 0) return 0;
  sb::bin_edges_kernel<<<sb::ceil_div(num_intersects, 256), 256, 0, (cudaStream_t)stream>>>(
      num_intersects, isect_ids_sorted, reinterpret_cast<int2*>(tile_bins));
  SB_RETURN_LAST();
}
