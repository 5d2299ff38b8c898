// Tile binning for the fused multi-view path: fixed-capacity tile segments + per-tile depth sort.
//
// Replaces, for scenes whose tiles hold at most `cap` Gaussians, the chain
//   cumsum -> .item() -> map_gaussian_to_intersects -> 64-bit radix sort -> get_tile_bin_edges
// (upstream gsplat.utils.bin_and_sort_gaussians reached from
// gsplat.rasterize_gaussians, /root/reference/utils/gs_utils.py:96-109) with
//   1. the projection kernel dropping (depth bits << 32 | gaussian id) straight into the segment
//      [tile * cap, tile * cap + count) of every tile a Gaussian covers (one atomic slot claim each),
//   2. one kernel that sorts every segment in shared memory on (depth, id) and writes
//      gaussian_ids_sorted, tile_bins, the per-view "has intersections" flags and the totals.
// Within a tile the result is exactly what the stable global sort produces: ascending depth bits,
// ties in emission order = ascending Gaussian id.  Nothing on the host has to know the number of
// intersections, so the forward pass needs no synchronisation before the raster kernel.
// Traffic: 8 bytes written + 8 read + 4 written per intersection instead of a scan, a 12-byte
// emission and six radix passes over 12-byte pairs.
#include "raster_common.cuh"

namespace sb {

constexpr int kMaxSegment = 8192;     // largest tile capacity the shared-memory sort takes (96 KB)
constexpr int kSortBlock = 256;

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

// dynamic shared memory: keys[cap] (8 B each) then ends[cap] (4 B each); cap = segment capacity (power of two)
// totals: [0] += intersections, [1] = 1 if some tile overflowed its segment.
#ifndef SB_SORT_MINB
#define SB_SORT_MINB 6
#endif
__global__ void __launch_bounds__(kSortBlock, SB_SORT_MINB) sort_tiles_kernel(int cap, int tiles_per_view,
                                                                const int* __restrict__ counts,
                                                                const unsigned long long* __restrict__ keys,
                                                                int* __restrict__ gids_sorted, int2* __restrict__ bins,
                                                                int* __restrict__ view_has, int* __restrict__ totals,
                                                                int n, int tiles_x, const float4* __restrict__ packed,
                                                                int* __restrict__ rect_ids, int2* __restrict__ rect_meta,
                                                                int* __restrict__ pool_head, int pool_cap) {
  extern __shared__ unsigned long long s_keys[];
  int* s_end = reinterpret_cast<int*>(s_keys + cap);
  __shared__ unsigned s_min, s_max;
  __shared__ int s_warp[kSortBlock / 32], s_overflow;
  __shared__ int s_rcnt[8], s_roff[8], s_base;
  const int count = __ldg(counts + blockIdx.x);
  const int len = min(count, cap);
  const int2 range = make_int2(blockIdx.x * cap, blockIdx.x * cap + len);       // V*T*cap < 2^31 (host checks)
  if (threadIdx.x == 0) {
    // empty tiles keep (0, 0) like upstream's zero-initialised tile_bins
    bins[blockIdx.x] = len > 0 ? range : make_int2(0, 0);
    if (len > 0) { view_has[blockIdx.x / tiles_per_view] = 1; atomicAdd(totals, len); }
    if (count > cap) atomicOr(totals + 1, 1);
  }
  if (len <= 0) {
    // (start = -1: "the whole tile is empty" -- the forward rasteriser then writes the tile in one go)
    if (rect_meta && threadIdx.x < 8) rect_meta[(size_t)blockIdx.x * 8 + threadIdx.x] = make_int2(-1, 0);
    return;
  }
  const unsigned long long* src = keys + range.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (len <= 32) {
    // one warp, one key per lane: rank by counting
    if (warp == 0) {
      const unsigned long long mine = lane < len ? __ldg(src + lane) : ~0ull;
      int rank = 0;
      for (int o = 0; o < len; ++o) rank += __shfl_sync(0xffffffffu, mine, o) < mine;
      if (lane < len) s_keys[rank] = mine;
    }
    __syncthreads();
  } else {
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
  if (threadIdx.x == 0) { s_min = 0xffffffffu; s_max = 0u; s_overflow = 0; }
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
  // exclusive scan of the histogram (s_end[b] becomes the start of bucket b and is then advanced to its end):
  // every thread sums its own nb / 256 consecutive buckets, one block-wide scan of the 256 partial sums,
  // then the thread lays its buckets' starts down -- two barriers in all, whatever nb is
  {
    const int per = nb >= kSortBlock ? nb / kSortBlock : 1;     // nb is a power of two >= 64
    const int b0 = threadIdx.x * per;
    int sum = 0;
    if (b0 < nb) {
      for (int j = 0; j < per; ++j) {
        const int c = s_end[b0 + j];
        if (c > kBucketMax) s_overflow = 1;
        sum += c;
      }
    }
    int x = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= o) x += y;
    }
    if (lane == 31) s_warp[warp] = x;
    __syncthreads();
    int start = x - sum;
    for (int w = 0; w < warp; ++w) start += s_warp[w];
    if (b0 < nb) {
      for (int j = 0; j < per; ++j) {
        const int c = s_end[b0 + j];
        s_end[b0 + j] = start;
        start += c;
      }
    }
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
  }   // len > 32
  if (gids_sorted)
    for (int i = threadIdx.x; i < len; i += kSortBlock) gids_sorted[range.x + i] = (int)(unsigned)(s_keys[i] & 0xffffffffull);
  if (!rect_ids) return;

  // ---- rect lists: for each of the tile's eight 8x4 rectangles, the depth-ordered ids of the Gaussians
  // whose visible region reaches it (raster_common.cuh: rect_mask), compacted stably per rectangle into
  // rect_ids[(tile * 8 + r) * cap + ...]; the rect-list rasteriser gives every rectangle to one warp.
  // Phase A: every entry's 8-bit rectangle mask into shared memory (the bucket table s_end is free by now).
  // Phase B: stable compaction of the ids into the eight lists.
  unsigned char* s_mask = reinterpret_cast<unsigned char*>(s_end);
  const int v = blockIdx.x / tiles_per_view, tile = blockIdx.x - v * tiles_per_view;
  const int tile_y = tile / tiles_x, tile_x = tile - tile_y * tiles_x;
  const float4* vp = packed + (size_t)v * n * 3;
  if (threadIdx.x < 8) s_rcnt[threadIdx.x] = 0;
  __syncthreads();                                 // everyone is done with s_end as the bucket table
  {
    // the rectangles' list lengths are counted on the way: one vote per rectangle and 32 entries, lane r keeps
    // rectangle r's running count (the lists go into a POOL shared by all tiles -- no per-rectangle capacity, a
    // list takes exactly its length -- so the lengths must be known before anything is stored)
    int my_cnt = 0;
    for (int i0 = 0; i0 < len; i0 += kSortBlock) {
      const int i = i0 + threadIdx.x;
      unsigned m = 0u;
      if (i < len) {
        const float4* rec = vp + (size_t)(unsigned)(s_keys[i] & 0xffffffffull) * 3;
        m = rect_mask(__ldg(rec), __ldg(rec + 1), tile_x, tile_y);
        s_mask[i] = (unsigned char)m;
      }
#pragma unroll
      for (int r = 0; r < 8; ++r) {
        const int c = __popc(__ballot_sync(0xffffffffu, (m >> r) & 1u));
        if (lane == r) my_cnt += c;
      }
    }
    if (lane < 8 && my_cnt) atomicAdd(&s_rcnt[lane], my_cnt);
  }
  __syncthreads();
  {
    // the tile claims room for its eight lists with one atomic on the pool head; then warp r owns rectangle r and
    // walks the masks in list order, 32 entries per step: one ballot on bit r, the set lanes store their ids
    // behind the warp's running offset -- stable
    const int r = warp;                            // kSortBlock / 32 == 8 rectangles
    static_assert(kSortBlock == 8 * 32, "one warp per rectangle");
    if (threadIdx.x == 0) {
      int tot = 0;
#pragma unroll
      for (int k = 0; k < 8; ++k) { s_roff[k] = tot; tot += s_rcnt[k]; }
      int base = tot > 0 ? atomicAdd(pool_head, tot) : 0;        // (keeps counting past the end: the host learns the need)
      if (base + tot > pool_cap) { base = -1; atomicOr(totals + 1, 2); }
      s_base = base;
    }
    __syncthreads();
    const int cnt = s_rcnt[r];
    const int base = s_base;
    const int start = base < 0 ? 0 : base + s_roff[r];
    if (lane == 0) rect_meta[(size_t)blockIdx.x * 8 + r] = make_int2(start, base < 0 ? 0 : cnt);
    if (base < 0) return;
    int* out = rect_ids + start;
    const unsigned lt = (1u << lane) - 1u;
    int off = 0;
#pragma unroll 4
    for (int i0 = 0; i0 < len; i0 += 32) {
      const int i = i0 + lane;
      const bool hit = i < len && ((s_mask[i] >> r) & 1u);
      const unsigned bal = __ballot_sync(0xffffffffu, hit);
      if (hit) out[off + __popc(bal & lt)] = (int)(unsigned)(s_keys[i] & 0xffffffffull);
      off += __popc(bal);
    }
  }
}

}  // namespace sb

extern "C" int sb_tile_sort_capacity(void) { return sb::kMaxSegment; }

extern "C" int sb_sort_tiles(int num_views, int tiles_per_view, int tile_capacity, const int* tile_counts,
                             const uint64_t* tile_keys, int* gaussian_ids_sorted, int* tile_bins,
                             int* view_has_isects, int* totals, int n, int tiles_x, const float* packed,
                             int* rect_ids, int* rect_meta, int* pool_head, int pool_capacity,
                             sb_stream_t stream) {
  if (num_views < 1 || tiles_per_view < 1 || tile_capacity < 64 || tile_capacity > sb::kMaxSegment ||
      (tile_capacity & (tile_capacity - 1)) ||
      (long long)num_views * tiles_per_view * tile_capacity > 2147483647LL ||
      (rect_ids && ((!packed && n > 0) || !rect_meta || !pool_head || pool_capacity < 1 || tiles_x < 1 ||
                    tiles_per_view % tiles_x)))
    return SB_ERR_BAD_ARG;
  const size_t smem = (size_t)tile_capacity * 12;
  if (smem > 40 * 1024)      // the kernel also has a little static shared memory
    cudaFuncSetAttribute(sb::sort_tiles_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  sb::sort_tiles_kernel<<<num_views * tiles_per_view, sb::kSortBlock, smem, (cudaStream_t)stream>>>(
      tile_capacity, tiles_per_view, tile_counts, reinterpret_cast<const unsigned long long*>(tile_keys),
      gaussian_ids_sorted, reinterpret_cast<int2*>(tile_bins), view_has_isects, totals, n, tiles_x,
      reinterpret_cast<const float4*>(packed), rect_ids, reinterpret_cast<int2*>(rect_meta), pool_head, pool_capacity);
  SB_RETURN_LAST();
}
