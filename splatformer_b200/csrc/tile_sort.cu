// Tile binning for the fused multi-view path: counting sort by tile + per-tile depth sort.
//
// Replaces, for scenes whose tiles hold at most kMaxSegment Gaussians, the chain
//   cumsum -> map_gaussian_to_intersects -> 64-bit radix sort -> get_tile_bin_edges
// (upstream gsplat.utils.bin_and_sort_gaussians reached from
// gsplat.rasterize_gaussians, /root/reference/utils/gs_utils.py:96-109) with
//   1. per-tile intersection counts, accumulated with global reductions by the projection kernel,
//   2. one exclusive scan over the V*T counts  -> tile_bins (start, end) directly,
//   3. a scatter of (depth bits << 32 | gaussian id) into the tile's segment (atomic cursor),
//   4. a bitonic sort of every segment in shared memory, keyed on (depth, id).
// The result is the same gaussian_ids_sorted / tile_bins the global sort produces: within a tile the
// stable radix sort orders by depth bits and, for equal depths, by emission order, which is
// ascending Gaussian id -- exactly the (depth, id) order used here.  Traffic: 12 bytes written + 8
// read + 4 written per intersection instead of six radix passes over 12-byte pairs.
#include "sb_common.cuh"

namespace sb {

constexpr int kMaxSegment = 8192;     // Gaussians per tile the shared-memory sort can take (96 KB)
constexpr int kScanBlock = 1024;
constexpr int kSortBlock = 256;

// Single-CTA exclusive scan of the tile counts.  totals[0] = number of intersections, totals[1] =
// largest tile count.
// view_end[v] = intersections of views 0..v (what the raster kernel uses to spot a view without any).
__global__ void __launch_bounds__(kScanBlock) tile_offsets_kernel(int num_tiles, int tiles_per_view,
                                                                  const int* __restrict__ counts,
                                                                  int2* __restrict__ bins, int* __restrict__ totals,
                                                                  int* __restrict__ view_end) {
  __shared__ int s_warp[kScanBlock / 32];
  __shared__ int s_carry;
  __shared__ int s_max[kScanBlock / 32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) s_carry = 0;
  int vmax = 0;
  __syncthreads();
  for (int base = 0; base < num_tiles; base += kScanBlock) {
    const int i = base + threadIdx.x;
    const int c = i < num_tiles ? __ldg(counts + i) : 0;
    vmax = max(vmax, c);
    int x = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= o) x += y;
    }
    if (lane == 31) s_warp[warp] = x;
    __syncthreads();
    if (warp == 0) {
      int w = s_warp[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int y = __shfl_up_sync(0xffffffffu, w, o);
        if (lane >= o) w += y;
      }
      s_warp[lane] = w;
    }
    __syncthreads();
    const int carry = s_carry;
    const int incl = carry + (warp ? s_warp[warp - 1] : 0) + x;
    // empty tiles keep (0, 0) like upstream's zero-initialised tile_bins
    if (i < num_tiles) {
      bins[i] = c > 0 ? make_int2(incl - c, incl) : make_int2(0, 0);
      if ((i + 1) % tiles_per_view == 0) view_end[i / tiles_per_view] = incl;
    }
    __syncthreads();
    if (threadIdx.x == kScanBlock - 1) s_carry = incl;
    __syncthreads();
  }
  vmax = __reduce_max_sync(0xffffffffu, vmax);
  if (lane == 0) s_max[warp] = vmax;
  __syncthreads();
  if (threadIdx.x == 0) {
    int m = 0;
    for (int w = 0; w < kScanBlock / 32; ++w) m = max(m, s_max[w]);
    totals[0] = s_carry;
    totals[1] = m;
  }
}

// One thread per (view, Gaussian): claims a slot in every tile of its bbox.  `cursors` holds the tile
// counts on entry and is counted down to zero (slot = start + old - 1), so no second zero-fill is needed.
// Consecutive blocks take consecutive VIEWS of the same Gaussian chunk: the resident blocks then hit
// the counters of all views at once (contention per counter drops by V) while loads stay coalesced.
__global__ void __launch_bounds__(256) scatter_isects_kernel(
    int num_views, int n, const float* __restrict__ xys, int xy_stride, const float* __restrict__ depths,
    const int* __restrict__ radii, int tiles_x, int tiles_y, int bw, const int2* __restrict__ bins,
    int* __restrict__ cursors, unsigned long long* __restrict__ keys) {
  const int v = blockIdx.x % num_views;
  const int g = (blockIdx.x / num_views) * 256 + threadIdx.x;
  if (g >= n) return;
  const size_t i = (size_t)v * n + g;
  const int r = __ldg(radii + i);
  if (r <= 0) return;
  const float2 c = __ldg(reinterpret_cast<const float2*>(xys + i * xy_stride));
  int min_x, max_x, min_y, max_y;
  tile_bbox(c.x, c.y, (float)r, tiles_x, tiles_y, bw, min_x, max_x, min_y, max_y);
  const unsigned long long key =
      ((unsigned long long)(unsigned)__float_as_int(__ldg(depths + i)) << 32) | (unsigned)g;
  const int tile0 = v * tiles_x * tiles_y;
  for (int ty = min_y; ty < max_y; ++ty)
    for (int tx = min_x; tx < max_x; ++tx) {
      const int t = tile0 + ty * tiles_x + tx;
      const int old = atomicSub(cursors + t, 1);
      keys[(size_t)__ldg(&bins[t].x) + (old - 1)] = key;
    }
}

// ------------------------------------------------------------------------------------------------
// Per-tile sort of the (depth bits << 32 | id) keys, one CTA per tile, in shared memory.
//
// Depths inside a tile spread over a narrow range, so a bucket sort on the depth bits does the bulk
// of the work in O(len): bucket = monotone map of the key's depth bits onto [0, B) with B >= len
// buckets, histogram + scan + scatter in shared memory, then every thread insertion-sorts the few
// keys of its buckets on the full 64-bit key.  If some bucket is too full for that (many equal
// depths) the CTA falls back to a bitonic sort of the whole segment -- same result, O(len log^2 len).
// ------------------------------------------------------------------------------------------------
constexpr int kBucketMax = 24;      // insertion-sort limit per bucket

__device__ __forceinline__ void bitonic_sort_smem(unsigned long long* s, int len_pow2) {
  for (int k = 2; k <= len_pow2; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int t = threadIdx.x; t < (len_pow2 >> 1); t += kSortBlock) {
        // element pair (a, a + j) of the j-stride stage; ascending blocks when (a & k) == 0
        const int a = ((t & ~(j - 1)) << 1) | (t & (j - 1));
        const int b = a | j;
        const unsigned long long x = s[a], y = s[b];
        const bool up = (a & k) == 0;
        if ((x > y) == up) { s[a] = y; s[b] = x; }
      }
      __syncthreads();
    }
  }
}

// dynamic shared memory: keys[cap] (8 B each) then ends[cap] (4 B each), cap = power of two >= max tile count
__global__ void __launch_bounds__(kSortBlock) sort_tiles_kernel(int cap, const int2* __restrict__ bins,
                                                                const unsigned long long* __restrict__ keys,
                                                                int* __restrict__ gids_sorted) {
  extern __shared__ unsigned long long s_keys[];
  int* s_end = reinterpret_cast<int*>(s_keys + cap);
  __shared__ unsigned s_min, s_max;
  __shared__ int s_warp[kSortBlock / 32], s_carry, s_overflow;
  const int2 range = __ldg(bins + blockIdx.x);
  const int len = range.y - range.x;
  if (len <= 0 || len > cap) return;      // (len > cap: the host never launches this path for such scenes)
  const unsigned long long* src = keys + range.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (len <= 32) {
    // one warp, one key per lane: rank by counting
    if (warp == 0) {
      const unsigned long long mine = lane < len ? __ldg(src + lane) : ~0ull;
      int rank = 0;
      for (int o = 0; o < len; ++o) rank += __shfl_sync(0xffffffffu, mine, o) < mine;
      if (lane < len) gids_sorted[range.x + rank] = (int)(unsigned)(mine & 0xffffffffull);
    }
    return;
  }
  int nb = 64;                             // buckets: power of two >= len
  while (nb < len) nb <<= 1;
  // pass 1: depth-bit range of the tile
  unsigned lo = 0xffffffffu, hi = 0u;
  for (int i = threadIdx.x; i < len; i += kSortBlock) {
    const unsigned d = (unsigned)(__ldg(src + i) >> 32);
    lo = min(lo, d); hi = max(hi, d);
  }
  lo = __reduce_min_sync(0xffffffffu, lo);
  hi = __reduce_max_sync(0xffffffffu, hi);
  if (threadIdx.x == 0) { s_min = 0xffffffffu; s_max = 0u; s_carry = 0; s_overflow = 0; }
  for (int i = threadIdx.x; i < nb; i += kSortBlock) s_end[i] = 0;
  __syncthreads();
  if (lane == 0) { atomicMin(&s_min, lo); atomicMax(&s_max, hi); }
  __syncthreads();
  const unsigned kmin = s_min;
  const float scale = (float)nb / ((float)(s_max - kmin) + 1.0f);
  auto bucket_of = [&](unsigned long long key) -> int {
    const int b = (int)((float)((unsigned)(key >> 32) - kmin) * scale);    // monotone in the depth bits
    return min(b, nb - 1);
  };
  // pass 2: histogram
  for (int i = threadIdx.x; i < len; i += kSortBlock) atomicAdd(&s_end[bucket_of(__ldg(src + i))], 1);
  __syncthreads();
  // exclusive scan of the histogram (s_end[b] becomes the start of bucket b and is then advanced to its end)
  for (int base = 0; base < nb; base += kSortBlock) {
    const bool in = base + (int)threadIdx.x < nb;
    const int c = in ? s_end[base + threadIdx.x] : 0;
    if (c > kBucketMax) s_overflow = 1;
    int x = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= o) x += y;
    }
    if (lane == 31) s_warp[warp] = x;
    __syncthreads();
    int wsum = 0;
    for (int w = 0; w < warp; ++w) wsum += s_warp[w];
    const int carry = s_carry;
    const int incl = carry + wsum + x;
    if (in) s_end[base + threadIdx.x] = incl - c;
    __syncthreads();
    if (threadIdx.x == kSortBlock - 1) s_carry = incl;
    __syncthreads();
  }
  // pass 3: scatter into buckets
  for (int i = threadIdx.x; i < len; i += kSortBlock) {
    const unsigned long long key = __ldg(src + i);
    s_keys[atomicAdd(&s_end[bucket_of(key)], 1)] = key;
  }
  __syncthreads();
  if (s_overflow) {
    int p2 = 64;
    while (p2 < len) p2 <<= 1;
    for (int i = len + threadIdx.x; i < p2; i += kSortBlock) s_keys[i] = ~0ull;
    __syncthreads();
    bitonic_sort_smem(s_keys, p2);
  } else {
    // every bucket [end[b-1], end[b]) holds at most kBucketMax keys: insertion sort on the full key
    for (int b = threadIdx.x; b < nb; b += kSortBlock) {
      const int e = s_end[b];
      const int st = b ? s_end[b - 1] : 0;
      for (int i = st + 1; i < e; ++i) {
        const unsigned long long key = s_keys[i];
        int j = i - 1;
        while (j >= st && s_keys[j] > key) { s_keys[j + 1] = s_keys[j]; --j; }
        s_keys[j + 1] = key;
      }
    }
    __syncthreads();
  }
  for (int i = threadIdx.x; i < len; i += kSortBlock) gids_sorted[range.x + i] = (int)(unsigned)(s_keys[i] & 0xffffffffull);
}

}  // namespace sb

extern "C" int sb_tile_sort_capacity(void) { return sb::kMaxSegment; }

extern "C" int sb_tile_offsets(int num_views, int tiles_per_view, const int* tile_counts, int* tile_bins,
                               int* totals, int* view_isect_end, sb_stream_t stream) {
  if (num_views < 1 || tiles_per_view < 1 || !tile_counts || !tile_bins || !totals || !view_isect_end)
    return SB_ERR_BAD_ARG;
  sb::tile_offsets_kernel<<<1, sb::kScanBlock, 0, (cudaStream_t)stream>>>(
      num_views * tiles_per_view, tiles_per_view, tile_counts, reinterpret_cast<int2*>(tile_bins), totals,
      view_isect_end);
  SB_RETURN_LAST();
}

extern "C" int sb_scatter_isects(int num_views, int n, const float* xys, int xys_stride, const float* depths,
                                 const int* radii, int tiles_x, int tiles_y, int block_width,
                                 const int* tile_bins, int* tile_counts, uint64_t* isect_keys,
                                 sb_stream_t stream) {
  if (n < 0 || num_views < 1 || block_width < 2 || block_width > 16 || xys_stride < 2 || (xys_stride & 1))
    return SB_ERR_BAD_ARG;
  if (n == 0) return 0;
  sb::scatter_isects_kernel<<<(unsigned)num_views * (unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      num_views, n, xys, xys_stride, depths, radii, tiles_x, tiles_y, block_width,
      reinterpret_cast<const int2*>(tile_bins), tile_counts, reinterpret_cast<unsigned long long*>(isect_keys));
  SB_RETURN_LAST();
}

extern "C" int sb_sort_tiles(int num_tiles_total, int max_tile_count, const int* tile_bins,
                             const uint64_t* isect_keys, int* gaussian_ids_sorted, sb_stream_t stream) {
  if (num_tiles_total < 1 || max_tile_count > sb::kMaxSegment) return SB_ERR_BAD_ARG;
  if (max_tile_count <= 0) return 0;
  int cap = 64;
  while (cap < max_tile_count) cap <<= 1;
  const size_t smem = (size_t)cap * 12;
  if (smem > 48 * 1024)
    cudaFuncSetAttribute(sb::sort_tiles_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  sb::sort_tiles_kernel<<<num_tiles_total, sb::kSortBlock, smem, (cudaStream_t)stream>>>(
      cap, reinterpret_cast<const int2*>(tile_bins), reinterpret_cast<const unsigned long long*>(isect_keys),
      gaussian_ids_sorted);
  SB_RETURN_LAST();
}
