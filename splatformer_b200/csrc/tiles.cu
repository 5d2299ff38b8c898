// Tiling: intersection offsets, key emission, key sort, bin edges (upstream
// gsplat.utils.{compute_cumulative_intersects, map_gaussian_to_intersects, bin_and_sort_gaussians,
// get_tile_bin_edges}, reached from gsplat.rasterize_gaussians at
// /root/reference/utils/gs_utils.py:96-109).  All integer work, bit-exact by construction.
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include "sb_common.cuh"

namespace sb {

__global__ void __launch_bounds__(256) map_isect_kernel(
    int num_views, int n, const float* __restrict__ xys, int xy_stride, const float* __restrict__ depths,
    const int* __restrict__ radii, const int* __restrict__ cum, int tiles_x, int tiles_y, int bw,
    int64_t* __restrict__ keys, int* __restrict__ gids) {
  const size_t i = (size_t)blockIdx.x * 256 + threadIdx.x;
  if (i >= (size_t)num_views * n) return;
  const int r = __ldg(radii + i);
  if (r <= 0) return;
  const int v = (int)(i / n);
  const int g = (int)(i - (size_t)v * n);
  const float2 c = __ldg(reinterpret_cast<const float2*>(xys + i * xy_stride));
  int min_x, max_x, min_y, max_y;
  tile_bbox(c.x, c.y, (float)r, tiles_x, tiles_y, bw, min_x, max_x, min_y, max_y);
  size_t k = i == 0 ? 0 : (size_t)__ldg(cum + i - 1);
  const int64_t depth_bits = (int64_t)(uint32_t)__float_as_int(__ldg(depths + i));
  const int64_t tile0 = (int64_t)v * tiles_x * tiles_y;
  for (int ty = min_y; ty < max_y; ++ty)
    for (int tx = min_x; tx < max_x; ++tx) {
      keys[k] = ((tile0 + (int64_t)ty * tiles_x + tx) << 32) | depth_bits;
      gids[k] = g;
      ++k;
    }
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

}  // namespace sb

extern "C" size_t sb_cumsum_temp_bytes(int count) {
  size_t bytes = 0;
  cub::DeviceScan::InclusiveSum(nullptr, bytes, (const int*)nullptr, (int*)nullptr, count);
  return bytes;
}

extern "C" int sb_cumsum_i32(int count, const int* in, int* out, void* temp, size_t temp_bytes,
                             sb_stream_t stream) {
  if (count < 0) return SB_ERR_BAD_ARG;
  if (count == 0) return 0;
  size_t need = sb_cumsum_temp_bytes(count);
  if (temp_bytes < need) return SB_ERR_TEMP_TOO_SMALL;
  return (int)cub::DeviceScan::InclusiveSum(temp, temp_bytes, in, out, count, (cudaStream_t)stream);
}

extern "C" int sb_map_gaussian_to_intersects(int num_views, int n, const float* xys, int xys_stride,
                                             const float* depths, const int* radii,
                                             const int* cum_tiles_hit, int tiles_x, int tiles_y,
                                             int block_width, int64_t* isect_ids, int* gaussian_ids,
                                             sb_stream_t stream) {
  if (n < 0 || num_views < 1 || block_width < 2 || block_width > 16 || xys_stride < 2 || (xys_stride & 1))
    return SB_ERR_BAD_ARG;
  const size_t total = (size_t)num_views * n;
  if (total == 0) return 0;
  sb::map_isect_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      num_views, n, xys, xys_stride, depths, radii, cum_tiles_hit, tiles_x, tiles_y, block_width, isect_ids,
      gaussian_ids);
  SB_RETURN_LAST();
}

extern "C" size_t sb_sort_temp_bytes(int num_intersects) {
  size_t bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, bytes, (const int64_t*)nullptr, (int64_t*)nullptr,
                                  (const int*)nullptr, (int*)nullptr, num_intersects, 0, 64);
  return bytes;
}

extern "C" int sb_sort_pairs(int num_intersects, int end_bit, const int64_t* keys_in,
                             const int* vals_in, int64_t* keys_out, int* vals_out, void* temp,
                             size_t temp_bytes, sb_stream_t stream) {
  if (num_intersects < 0 || end_bit < 1 || end_bit > 64) return SB_ERR_BAD_ARG;
  if (num_intersects == 0) return 0;
  if (temp_bytes < sb_sort_temp_bytes(num_intersects)) return SB_ERR_TEMP_TOO_SMALL;
  return (int)cub::DeviceRadixSort::SortPairs(temp, temp_bytes, keys_in, keys_out, vals_in, vals_out,
                                              num_intersects, 0, end_bit, (cudaStream_t)stream);
}

extern "C" int sb_tile_bin_edges(int num_intersects, const int64_t* isect_ids_sorted,
                                 int num_tiles_total, int* tile_bins, sb_stream_t stream) {
  if (num_intersects < 0 || num_tiles_total < 0) return SB_ERR_BAD_ARG;
  if (num_intersects == 0) return 0;
  sb::bin_edges_kernel<<<sb::ceil_div(num_intersects, 256), 256, 0, (cudaStream_t)stream>>>(
      num_intersects, isect_ids_sorted, reinterpret_cast<int2*>(tile_bins));
  SB_RETURN_LAST();
}
