// Hand-written device-wide LSD radix sort (one sweep per digit, decoupled look-back) and prefix sum for the
// tiling stage: the 64-bit (tile | depth) key sort that upstream gsplat
// does with torch.sort + torch.gather inside gsplat.rasterize_gaussians
// (/root/reference/utils/gs_utils.py:96-109), the uint32 tile-index sort of the fused path's global-sort
// branch, and compute_cumulative_intersects' cumsum.  No library calls: three kernels.
//
//   digit_histogram_kernel   one read of the keys: 256-bin histograms of EVERY pass's digit (every thread run-length
//                            encodes 16 consecutive keys, one shared atomic per run, one global atomic per bin and CTA).
//   onesweep_kernel          one pass = one read + one write of the pairs.  A CTA takes the next tile of
//                            512 x IPT consecutive pairs (dynamic tile index, so that every predecessor of a
//                            running tile is running or done), ranks them stably by the pass's 8-bit digit
//                            (per warp: every lane ORs its bit into the digit's word of a shared bitmap --
//                            ATOMS.OR, not MATCH.ANY, which costs ~60 cycles per warp on the SM's one ADU --
//                            and ranks itself by popc below its lane; the warp's digit counter grows by the
//                            word's popc; per CTA: a scan of the 16 warps' counters), publishes its 256
//                            digit counts, looks back over the preceding tiles' counts / inclusive prefixes
//                            (one thread per digit, flag in the top two bits of a 32-bit word), and scatters
//                            through shared memory so that the stores leave in digit-contiguous runs.
//                            The LAST pass of the tile-index sort keeps only the values (the sorted Gaussian
//                            ids -- the key is implied by the bin) and counts the pairs per tile on the fly,
//                            which replaces the bin-edge kernel's read of the sorted keys.
//   scan_kernel              inclusive int32 prefix sum, single pass, same look-back scheme.
//
// Stability: within a warp the 32 x IPT pairs are consecutive and visited in order, warps and tiles are
// ordered by position, so equal digits keep their input order -- the property the depth-pre-ordered
// emission (tiles.cu) and the tie rule "ascending Gaussian id" rest on.  Bit-identical to any stable sort.
#include <algorithm>

#include "sb_common.cuh"

namespace sb {

constexpr int kRadix = 256;
#ifndef SB_RS_THREADS
#define SB_RS_THREADS 512
#endif
#ifndef SB_RS_LOOKBACK
#define SB_RS_LOOKBACK 16
#endif
#ifndef SB_RS_IPT32
#define SB_RS_IPT32 8
#endif
#ifndef SB_RS_IPT64
#define SB_RS_IPT64 6
#endif
#ifndef SB_RS_MINB
#define SB_RS_MINB 2
#endif
constexpr int kRsThreads = SB_RS_THREADS;
constexpr int kRsWarps = kRsThreads / 32;
static_assert(kRsThreads >= 256 && kRsThreads % 32 == 0, "one thread per digit in the look-back");
constexpr unsigned kStAgg = 1u << 30, kStPre = 2u << 30, kStMask = (1u << 30) - 1u;
constexpr int kMaxPasses = 8;
constexpr int kLookBack = SB_RS_LOOKBACK;

// digit of a pass: key bits [shift, shift + 8), cut at the sort's end_bit by `mask` (the last pass of a
// sort whose bit count is not a multiple of 8 looks at fewer bits -- the ones above must not influence the order)
template <typename K>
__device__ __forceinline__ unsigned digit_of(K key, int shift, unsigned mask) {
  return (unsigned)(key >> shift) & mask;
}
__host__ __device__ inline unsigned pass_mask(int begin_bit, int end_bit, int p) {
  const int left = end_bit - (begin_bit + 8 * p);
  return left >= 8 ? 255u : ((1u << left) - 1u);
}

// ------------------------------------------------------------------------------------------------
// Histograms of all passes' digits in one read.  hist: [passes][256] (zero on entry).
// Every thread takes kHistBatch CONSECUTIVE keys (16-byte vector loads) and run-length encodes each pass's
// digit before touching shared memory: tile indices come in runs whose high bytes coincide (a Gaussian's
// rows), so the high passes cost one shared atomic per thread and batch instead of one per key.  No
// MATCH.ANY: it runs at ~60 cycles per warp on the SM's single ADU (scripts/micro/match_vs_vote.cu, measured
// on this B200), which is what bounds a library histogram / one-sweep pass on this chip.
// ------------------------------------------------------------------------------------------------
constexpr int kHistThreads = 256;
constexpr int kHistBatch = 16;

template <typename K>
__global__ void __launch_bounds__(kHistThreads) digit_histogram_kernel(const K* __restrict__ keys, size_t count,
                                                                      int begin_bit, int end_bit, int passes,
                                                                      unsigned* __restrict__ hist) {
  __shared__ unsigned s_hist[kMaxPasses][kRadix];
  for (int i = threadIdx.x; i < passes * kRadix; i += kHistThreads) (&s_hist[0][0])[i] = 0u;
  __syncthreads();
  constexpr size_t kChunk = (size_t)kHistThreads * kHistBatch;
  constexpr int kVec = 16 / sizeof(K);                         // keys per 16-byte load
  const size_t chunks = (count + kChunk - 1) / kChunk;
  const bool aligned = (reinterpret_cast<size_t>(keys) & 15) == 0;
  for (size_t c = blockIdx.x; c < chunks; c += gridDim.x) {
    const size_t base = c * kChunk + (size_t)threadIdx.x * kHistBatch;
    K key[kHistBatch];
    int have = kHistBatch;
    if (aligned && base + kHistBatch <= count) {
#pragma unroll
      for (int j = 0; j < kHistBatch / kVec; ++j) {
        const uint4 v = __ldg(reinterpret_cast<const uint4*>(keys + base) + j);
        if (sizeof(K) == 4) {
          key[4 * j] = (K)v.x; key[4 * j + 1] = (K)v.y; key[4 * j + 2] = (K)v.z; key[4 * j + 3] = (K)v.w;
        } else {
          key[2 * j] = (K)(((unsigned long long)v.y << 32) | v.x);
          key[2 * j + 1] = (K)(((unsigned long long)v.w << 32) | v.z);
        }
      }
    } else {
      have = base < count ? (int)min((size_t)kHistBatch, count - base) : 0;
#pragma unroll
      for (int j = 0; j < kHistBatch; ++j) key[j] = j < have ? keys[base + j] : (K)0;
    }
    if (have == 0) continue;
    for (int p = 0; p < passes; ++p) {
      const int shift = begin_bit + 8 * p;
      const unsigned mask = pass_mask(begin_bit, end_bit, p);
      unsigned cur = digit_of(key[0], shift, mask), run = 1;
#pragma unroll
      for (int j = 1; j < kHistBatch; ++j) {
        if (j < have) {
          const unsigned d = digit_of(key[j], shift, mask);
          if (d != cur) { atomicAdd(&s_hist[p][cur], run); cur = d; run = 0; }
          ++run;
        }
      }
      atomicAdd(&s_hist[p][cur], run);
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < passes * kRadix; i += kHistThreads) {
    const unsigned c = (&s_hist[0][0])[i];
    if (c) atomicAdd(hist + i, c);
  }
}

// ------------------------------------------------------------------------------------------------
// One sweep.  VALS_ONLY: the pass does not write the keys (last pass of the tile-index sort) and, if
// key_counts is given, adds the number of pairs per full key value to key_counts[key].
// status: [tiles][256] words, zero on entry; tile_counter: one word, zero on entry.
// hist: this pass's 256 digit counts (complete before the launch).
// ------------------------------------------------------------------------------------------------
template <typename K, int IPT>
struct SweepSmem {
  K keys[kRsThreads * IPT];
  int vals[kRsThreads * IPT];
  unsigned warp_hist[kRsWarps][kRadix];
  unsigned warp_bits[kRsWarps][kRadix];   // per warp and digit: the lanes holding that digit in the current round
  unsigned bin_start[kRadix];      // first position of the digit in the tile-sorted order
  unsigned global_base[kRadix];    // global index of tile-sorted position 0 of the digit's run, minus bin_start
  unsigned scan_tmp[2][kRadix / 32];
  unsigned tile;
};

template <typename K, int IPT, bool VALS_ONLY>
__global__ void __launch_bounds__(kRsThreads, SB_RS_MINB) onesweep_kernel(const K* __restrict__ keys_in,
                                                             const int* __restrict__ vals_in,
                                                             K* __restrict__ keys_out, int* __restrict__ vals_out,
                                                             unsigned count, int shift, unsigned dmask,
                                                             const unsigned* __restrict__ hist,
                                                             unsigned* __restrict__ status,
                                                             unsigned* __restrict__ tile_counter,
                                                             int* __restrict__ key_counts) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  SweepSmem<K, IPT>& S = *reinterpret_cast<SweepSmem<K, IPT>*>(smem_raw);
  constexpr int kTile = kRsThreads * IPT;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) S.tile = atomicAdd(tile_counter, 1u);
  for (int i = threadIdx.x; i < 2 * kRsWarps * kRadix; i += kRsThreads) (&S.warp_hist[0][0])[i] = 0u;   // + warp_bits
  __syncthreads();
  const unsigned tile = S.tile;
  const unsigned tile_base = tile * (unsigned)kTile;
  const unsigned tile_items = min((unsigned)kTile, count - tile_base);

  // ---- load: warp w owns the 32*IPT consecutive pairs from tile_base + w*32*IPT, round i = 32 of them ----
  K key[IPT];
  const unsigned wbase = tile_base + (unsigned)warp * 32u * IPT + (unsigned)lane;
#pragma unroll
  for (int i = 0; i < IPT; ++i) {
    const unsigned idx = wbase + 32u * i;
    key[i] = idx < count ? keys_in[idx] : (K)0;
  }
  // ---- stable rank inside the warp.  Round i looks at 32 consecutive pairs, one per lane.  The lanes that
  // hold the same digit find each other through a bitmap in shared memory (ATOMS.OR of the lane bit, then a
  // read) -- no MATCH.ANY, which costs ~60 cycles of the SM's one ADU per warp (scripts/micro/match_vs_vote.cu)
  // and is what bounds a library one-sweep pass on this chip; one vote first catches the round whose 32
  // digits all coincide (the high byte of tile indices).  The group's lowest lane advances the warp's counter.
  unsigned short rank[IPT];
  unsigned* wh = S.warp_hist[warp];
  unsigned* bm = S.warp_bits[warp];
  const unsigned lt = (1u << lane) - 1u;
#pragma unroll
  for (int i = 0; i < IPT; ++i) {
    const bool valid = wbase + 32u * i < count;
    const unsigned d = digit_of(key[i], shift, dmask);
    const unsigned d0 = __shfl_sync(0xffffffffu, d, 0);
    const bool same = __all_sync(0xffffffffu, valid && d == d0);
    unsigned m = 0xffffffffu;
    if (!same) {
      if (valid) atomicOr(bm + d, 1u << lane);
      __syncwarp();
      m = valid ? bm[d] : 0u;
    }
    const unsigned prev = valid ? wh[d] : 0u;
    __syncwarp();
    if (valid && lane == __ffs(m) - 1) {
      wh[d] = prev + (unsigned)__popc(m);
      if (!same) bm[d] = 0u;
    }
    rank[i] = (unsigned short)(prev + (unsigned)__popc(m & lt));
    __syncwarp();
  }
  __syncthreads();

  // ---- per digit: offsets of the warps, tile count; publish; scans of the tile counts and of the global histogram ----
  unsigned tile_cnt = 0, gh = 0;
  if (threadIdx.x < kRadix) {
    const int d = threadIdx.x;
    unsigned run = 0;
#pragma unroll
    for (int w = 0; w < kRsWarps; ++w) {
      const unsigned c = S.warp_hist[w][d];
      S.warp_hist[w][d] = run;
      run += c;
    }
    tile_cnt = run;
    volatile unsigned* st = status + (size_t)tile * kRadix + d;
    *st = (tile == 0 ? kStPre : kStAgg) | tile_cnt;
    gh = hist[d];
  }
  // block-wide exclusive scans over the 256 digits (8 warps x 32) of tile_cnt and gh
  unsigned x = tile_cnt, y = gh;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const unsigned xs = __shfl_up_sync(0xffffffffu, x, o), ys = __shfl_up_sync(0xffffffffu, y, o);
    if (lane >= o) { x += xs; y += ys; }
  }
  if (lane == 31 && warp < kRadix / 32) { S.scan_tmp[0][warp] = x; S.scan_tmp[1][warp] = y; }
  __syncthreads();
  if (threadIdx.x < kRadix) {
    unsigned xo = 0, yo = 0;
    for (int w = 0; w < warp; ++w) { xo += S.scan_tmp[0][w]; yo += S.scan_tmp[1][w]; }
    const unsigned bin_start = xo + x - tile_cnt;        // exclusive
    const unsigned g_excl = yo + y - gh;                 // exclusive over the digits of the whole input
    // ---- decoupled look-back: pairs with this digit in all preceding tiles.  The CTAs in flight reach this
    // point at about the same time, so the nearest tile that already knows its inclusive prefix is typically
    // (number of resident CTAs) tiles back: the words are fetched kLookBack at a time (independent loads, one
    // latency) and consumed nearest first; a word that is not there yet is polled again with the next batch ----
    unsigned before = 0;
    if (tile > 0) {
      int t = (int)tile - 1;
      bool done = false;
      while (!done) {
        unsigned w[kLookBack];
#pragma unroll
        for (int k = 0; k < kLookBack; ++k) {
          const int tt = t - k;
          w[k] = tt >= 0 ? *(volatile const unsigned*)(status + (size_t)tt * kRadix + threadIdx.x) : kStPre;
        }
        int used = 0;
#pragma unroll
        for (int k = 0; k < kLookBack; ++k) {
          if (!done && used == k) {
            const unsigned flag = w[k] >> 30;
            if (flag != 0u) {
              before += w[k] & kStMask;
              used = k + 1;
              done = flag == 2u;
            }
          }
        }
        t -= used;
      }
      volatile unsigned* mine = status + (size_t)tile * kRadix + threadIdx.x;
      *mine = kStPre | (before + tile_cnt);
    }
    S.bin_start[threadIdx.x] = bin_start;
    S.global_base[threadIdx.x] = g_excl + before - bin_start;
  }
  __syncthreads();

  // ---- scatter into the tile-sorted order in shared memory; the values are fetched only now (their registers
  // would otherwise be held across the ranking and the look-back) and follow the keys to the same positions ----
#pragma unroll
  for (int i = 0; i < IPT; ++i) {
    const unsigned idx = wbase + 32u * i;
    if (idx < count) {
      const unsigned d = digit_of(key[i], shift, dmask);
      const unsigned pos = S.bin_start[d] + S.warp_hist[warp][d] + rank[i];
      rank[i] = (unsigned short)pos;
      S.keys[pos] = key[i];
    }
  }
  {
    int val[IPT];
#pragma unroll
    for (int i = 0; i < IPT; ++i) {
      const unsigned idx = wbase + 32u * i;
      val[i] = idx < count ? vals_in[idx] : 0;
    }
#pragma unroll
    for (int i = 0; i < IPT; ++i)
      if (wbase + 32u * i < count) S.vals[rank[i]] = val[i];
  }
  __syncthreads();
  // ---- store: position j of the tile-sorted order goes to global_base[digit] + j ----
  for (unsigned j0 = 0; j0 < tile_items; j0 += kRsThreads) {
    const unsigned j = j0 + threadIdx.x;
    const bool valid = j < tile_items;
    K k = (K)0;
    if (valid) {
      k = S.keys[j];
      const unsigned g = S.global_base[digit_of(k, shift, dmask)] + j;
      if (!VALS_ONLY) keys_out[g] = k;
      vals_out[g] = S.vals[j];
    }
    if (VALS_ONLY && key_counts) {
      // pairs per key (= per tile): the tile-sorted order holds equal keys in runs; inside these 32 consecutive
      // positions a lane heads a run if its key differs from its left neighbour's, and the head adds the
      // distance to the next head (one vote, no MATCH)
      const unsigned left = __shfl_up_sync(0xffffffffu, (unsigned)k, 1);
      const bool head = valid && (lane == 0 || left != (unsigned)k);
      const unsigned heads = __ballot_sync(0xffffffffu, head);
      const unsigned live = __ballot_sync(0xffffffffu, valid);
      if (head) {
        const unsigned after = (lane == 31) ? 0u : (heads >> (lane + 1));
        const int end = after ? lane + __ffs(after) : __popc(live);       // live lanes are a prefix of the warp
        atomicAdd(key_counts + (unsigned)k, end - lane);
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------
// Inclusive int32 prefix sum, one pass.  status: [tiles] 64-bit words (flag << 32 | value), zero on entry.
// ------------------------------------------------------------------------------------------------
constexpr int kScanThreads = 256, kScanIpt = 16, kScanTile = kScanThreads * kScanIpt;

__global__ void __launch_bounds__(kScanThreads) scan_kernel(const int* __restrict__ in, int* __restrict__ out, int count,
                                                          unsigned long long* __restrict__ status,
                                                          unsigned* __restrict__ tile_counter) {
  __shared__ int s_warp[kScanThreads / 32];
  __shared__ int s_prefix;
  __shared__ unsigned s_tile;
  if (threadIdx.x == 0) s_tile = atomicAdd(tile_counter, 1u);
  __syncthreads();
  const unsigned tile = s_tile;
  const int base = (int)tile * kScanTile + threadIdx.x * kScanIpt;
  int v[kScanIpt];
  if (base + kScanIpt <= count && (reinterpret_cast<size_t>(in) & 15) == 0) {
#pragma unroll
    for (int q = 0; q < kScanIpt / 4; ++q) {
      const int4 t = __ldg(reinterpret_cast<const int4*>(in + base) + q);
      v[4 * q] = t.x; v[4 * q + 1] = t.y; v[4 * q + 2] = t.z; v[4 * q + 3] = t.w;
    }
  } else {
#pragma unroll
    for (int i = 0; i < kScanIpt; ++i) v[i] = base + i < count ? __ldg(in + base + i) : 0;
  }
  int sum = 0;
#pragma unroll
  for (int i = 0; i < kScanIpt; ++i) { sum += v[i]; v[i] = sum; }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int x = sum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int y = __shfl_up_sync(0xffffffffu, x, o);
    if (lane >= o) x += y;
  }
  if (lane == 31) s_warp[warp] = x;
  __syncthreads();
  int off = x - sum;
  for (int w = 0; w < warp; ++w) off += s_warp[w];
  if (threadIdx.x == kScanThreads - 1) {
    const int total = off + sum;                               // this tile's aggregate
    volatile unsigned long long* st = status + tile;
    int before = 0;
    if (tile == 0) {
      *st = (2ull << 32) | (unsigned)total;
    } else {
      *st = (1ull << 32) | (unsigned)total;
      int t = (int)tile - 1;
      while (true) {
        volatile unsigned long long* p = status + t;
        unsigned long long w;
        do { w = *p; } while ((w >> 32) == 0ull);
        before += (int)(unsigned)w;
        if ((w >> 32) == 2ull) break;
        --t;
      }
      *st = (2ull << 32) | (unsigned)(before + total);
    }
    s_prefix = before;
  }
  __syncthreads();
  off += s_prefix;
  if (base + kScanIpt <= count && (reinterpret_cast<size_t>(out) & 15) == 0) {
#pragma unroll
    for (int q = 0; q < kScanIpt / 4; ++q)
      reinterpret_cast<int4*>(out + base)[q] = make_int4(v[4 * q] + off, v[4 * q + 1] + off, v[4 * q + 2] + off, v[4 * q + 3] + off);
  } else {
#pragma unroll
    for (int i = 0; i < kScanIpt; ++i)
      if (base + i < count) out[base + i] = v[i] + off;
  }
}

// tile_bins (T,2) from the per-tile pair counts and their inclusive prefix sum: [start, end) in the sorted
// order, (0,0) for an empty tile (upstream's zero-initialised tile_bins).
__global__ void __launch_bounds__(256) bins_from_counts_kernel(int num_tiles, const int* __restrict__ counts,
                                                             const int* __restrict__ incl, int2* __restrict__ bins) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= num_tiles) return;
  const int c = __ldg(counts + i), e = __ldg(incl + i);
  bins[i] = c > 0 ? make_int2(e - c, e) : make_int2(0, 0);
}

// ------------------------------------------------------------------------------------------------
// Host side: temp layout and pass scheduling.
//   [ hist: passes*256 u32 | counters: 8 u32 | status: passes * tiles * 256 u32 | alt keys | alt vals ]
// ------------------------------------------------------------------------------------------------
template <typename K> struct SortTraits;
template <> struct SortTraits<uint32_t> { static constexpr int kIpt = SB_RS_IPT32; };
template <> struct SortTraits<uint64_t> { static constexpr int kIpt = SB_RS_IPT64; };

inline size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

template <typename K>
struct SortPlan {
  int passes, key_bufs;
  unsigned tiles;
  size_t off_hist, off_counters, off_status, off_keys[2], off_vals, total;
  // vals_only: the last pass writes no keys, so the keys never reach an output buffer and ping-pong between
  // alternate buffers of their own (one for two passes, two from three passes on)
  SortPlan(size_t count, int num_bits, bool vals_only) {
    passes = (num_bits + 7) / 8;
    key_bufs = passes <= 1 ? 0 : (vals_only && passes >= 3 ? 2 : 1);
    constexpr int kTile = kRsThreads * SortTraits<K>::kIpt;
    tiles = (unsigned)((count + kTile - 1) / kTile);
    size_t o = 0;
    off_hist = o; o = align256(o + (size_t)kMaxPasses * kRadix * 4);
    off_counters = o; o = align256(o + 64);
    off_status = o; o = align256(o + (size_t)passes * tiles * kRadix * 4);
    off_keys[0] = o; o = align256(o + (key_bufs >= 1 ? count * sizeof(K) : 0));
    off_keys[1] = o; o = align256(o + (key_bufs >= 2 ? count * sizeof(K) : 0));
    off_vals = o; o = align256(o + (passes > 1 ? count * sizeof(int) : 0));
    total = o;
  }
};

// Sorts `count` pairs on key bits [begin_bit, end_bit).  vals_only: the last pass drops the keys (keys_out is
// not used) and, if key_counts is given, counts the pairs per key.  hist_ext (may be NULL): the digit
// histograms [passes][256], already complete (the emission kernel counted them); otherwise one read of the keys.
template <typename K>
int radix_sort_pairs(size_t count, int begin_bit, int end_bit, const K* keys_in, const int* vals_in, K* keys_out,
                     int* vals_out, void* temp, size_t temp_bytes, bool vals_only, int* key_counts,
                     const unsigned* hist_ext, cudaStream_t st) {
  if (count == 0) return 0;
  if (count >= (1ull << 30)) return SB_ERR_BAD_ARG;            // look-back words carry 30-bit counts
  SortPlan<K> plan(count, end_bit - begin_bit, vals_only);
  if (plan.passes > kMaxPasses) return SB_ERR_BAD_ARG;
  if (temp_bytes < plan.total) return SB_ERR_TEMP_TOO_SMALL;
  unsigned char* base = reinterpret_cast<unsigned char*>(temp);
  unsigned* hist_own = reinterpret_cast<unsigned*>(base + plan.off_hist);
  unsigned* counters = reinterpret_cast<unsigned*>(base + plan.off_counters);
  unsigned* status = reinterpret_cast<unsigned*>(base + plan.off_status);
  K* alt_keys[2] = {reinterpret_cast<K*>(base + plan.off_keys[0]), reinterpret_cast<K*>(base + plan.off_keys[1])};
  int* alt_vals = reinterpret_cast<int*>(base + plan.off_vals);
  // zero: (own histograms,) tile counters, look-back words of every pass
  unsigned char* z0 = hist_ext ? reinterpret_cast<unsigned char*>(counters) : reinterpret_cast<unsigned char*>(hist_own);
  cudaError_t err = cudaMemsetAsync(z0, 0, (base + plan.off_keys[0]) - z0, st);
  if (err != cudaSuccess) return (int)err;
  const unsigned* hist = hist_ext;
  if (!hist_ext) {
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const size_t want = (count + (size_t)kHistThreads * kHistBatch - 1) / ((size_t)kHistThreads * kHistBatch);
    const unsigned grid = (unsigned)std::min<size_t>(want, (size_t)sms * 8);
    digit_histogram_kernel<K><<<grid, kHistThreads, 0, st>>>(keys_in, count, begin_bit, end_bit, plan.passes, hist_own);
    hist = hist_own;
  }
  constexpr int kIpt = SortTraits<K>::kIpt;
  const size_t smem = sizeof(SweepSmem<K, kIpt>);
  cudaFuncSetAttribute(onesweep_kernel<K, kIpt, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  cudaFuncSetAttribute(onesweep_kernel<K, kIpt, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  // buffer schedule: the last pass writes the outputs and the passes before it alternate between the
  // alternate buffers and the outputs, so that the input is never written: pass p (0-based, P passes) writes
  // the outputs when P - 1 - p is even.  Keys of a vals-only sort stay in the alternate key buffers.
  const K* kin = keys_in;
  const int* vin = vals_in;
  for (int p = 0; p < plan.passes; ++p) {
    const bool last = p == plan.passes - 1;
    const bool to_out = ((plan.passes - 1 - p) & 1) == 0;
    int* vout = to_out ? vals_out : alt_vals;
    K* kout = vals_only ? alt_keys[plan.key_bufs == 2 ? (p & 1) : 0] : (to_out ? keys_out : alt_keys[0]);
    unsigned* pst = status + (size_t)p * plan.tiles * kRadix;
    if (last && vals_only)
      onesweep_kernel<K, kIpt, true><<<plan.tiles, kRsThreads, smem, st>>>(
          kin, vin, nullptr, vout, (unsigned)count, begin_bit + 8 * p, pass_mask(begin_bit, end_bit, p),
          hist + p * kRadix, pst, counters + p, key_counts);
    else
      onesweep_kernel<K, kIpt, false><<<plan.tiles, kRsThreads, smem, st>>>(
          kin, vin, kout, vout, (unsigned)count, begin_bit + 8 * p, pass_mask(begin_bit, end_bit, p),
          hist + p * kRadix, pst, counters + p, nullptr);
    kin = kout; vin = vout;
  }
  return (int)cudaGetLastError();
}

}  // namespace sb

// ---- C ABI -------------------------------------------------------------------------------------------------
extern "C" size_t sb_cumsum_temp_bytes(int count) {
  const size_t tiles = ((size_t)std::max(count, 0) + sb::kScanTile - 1) / sb::kScanTile;
  return sb::align256(tiles * 8 + 64) + 256;
}

extern "C" int sb_cumsum_i32(int count, const int* in, int* out, void* temp, size_t temp_bytes, sb_stream_t stream) {
  if (count < 0) return SB_ERR_BAD_ARG;
  if (count == 0) return 0;
  if (temp_bytes < sb_cumsum_temp_bytes(count) || !temp) return SB_ERR_TEMP_TOO_SMALL;
  const unsigned tiles = (unsigned)(((size_t)count + sb::kScanTile - 1) / sb::kScanTile);
  // 256-byte aligned: [counter (64 B) | status words]
  unsigned char* base = reinterpret_cast<unsigned char*>((reinterpret_cast<size_t>(temp) + 255) & ~(size_t)255);
  cudaStream_t st = (cudaStream_t)stream;
  cudaError_t err = cudaMemsetAsync(base, 0, 64 + (size_t)tiles * 8, st);
  if (err != cudaSuccess) return (int)err;
  sb::scan_kernel<<<tiles, sb::kScanThreads, 0, st>>>(in, out, count, reinterpret_cast<unsigned long long*>(base + 64),
                                                      reinterpret_cast<unsigned*>(base));
  SB_RETURN_LAST();
}

extern "C" size_t sb_sort_temp_bytes(int num_intersects) {
  return sb::SortPlan<uint64_t>((size_t)std::max(num_intersects, 0), 64, false).total + 256;
}

extern "C" size_t sb_sort32_temp_bytes(int count) {
  // covers sb_sort_pairs32 and sb_sort_tiles_global for any number of key bits
  const size_t c = (size_t)std::max(count, 0);
  return std::max(sb::SortPlan<uint32_t>(c, 32, false).total, sb::SortPlan<uint32_t>(c, 32, true).total) + 256;
}

static void* align_temp(void* temp, size_t& bytes) {
  const size_t a = (reinterpret_cast<size_t>(temp) + 255) & ~(size_t)255;
  const size_t lost = a - reinterpret_cast<size_t>(temp);
  bytes = bytes > lost ? bytes - lost : 0;
  return reinterpret_cast<void*>(a);
}

extern "C" int sb_sort_pairs_range(int count, int begin_bit, int end_bit, const int64_t* keys_in, const int* vals_in,
                                   int64_t* keys_out, int* vals_out, void* temp, size_t temp_bytes,
                                   sb_stream_t stream) {
  if (count < 0 || begin_bit < 0 || end_bit <= begin_bit || end_bit > 64) return SB_ERR_BAD_ARG;
  if (count == 0) return 0;
  if (!temp) return SB_ERR_TEMP_TOO_SMALL;
  temp = align_temp(temp, temp_bytes);
  return sb::radix_sort_pairs<uint64_t>((size_t)count, begin_bit, end_bit, reinterpret_cast<const uint64_t*>(keys_in),
                                        vals_in, reinterpret_cast<uint64_t*>(keys_out), vals_out, temp, temp_bytes,
                                        false, nullptr, nullptr, (cudaStream_t)stream);
}

extern "C" int sb_sort_pairs(int num_intersects, int end_bit, const int64_t* keys_in, const int* vals_in,
                             int64_t* keys_out, int* vals_out, void* temp, size_t temp_bytes, sb_stream_t stream) {
  return sb_sort_pairs_range(num_intersects, 0, end_bit, keys_in, vals_in, keys_out, vals_out, temp, temp_bytes, stream);
}

extern "C" int sb_sort_pairs32(int count, int end_bit, const uint32_t* keys_in, const int* vals_in, uint32_t* keys_out,
                               int* vals_out, void* temp, size_t temp_bytes, sb_stream_t stream) {
  if (count < 0 || end_bit < 1 || end_bit > 32) return SB_ERR_BAD_ARG;
  if (count == 0) return 0;
  if (!temp) return SB_ERR_TEMP_TOO_SMALL;
  temp = align_temp(temp, temp_bytes);
  return sb::radix_sort_pairs<uint32_t>((size_t)count, 0, end_bit, keys_in, vals_in, keys_out, vals_out, temp,
                                        temp_bytes, false, nullptr, nullptr, (cudaStream_t)stream);
}

// Tile-index sort of the fused path's global-sort branch: stable sort of (tile index, Gaussian id) pairs on the
// tile index (num_tiles_total = V * T distinct values); only the sorted ids are produced -- the key is implied
// by the bin -- and tile_bins (num_tiles_total, 2) comes out of the same call: the last sweep counts the pairs
// per tile, a single-pass scan and one elementwise kernel turn the counts into [start, end).
// digit_hist (may be NULL): the 256-bin histograms of the key's bytes 0..passes-1, [passes][256] uint32, if the
// caller already has them; otherwise the sort reads the keys once more.  (Counting them in the emission kernel
// was tried and dropped: 2.2 ms of shared-memory atomics at 376 M pairs against a 0.5 ms histogram kernel.)
static size_t tiles_global_extra(int num_tiles_total) {
  // counts | inclusive sums | scan state
  return sb::align256((size_t)num_tiles_total * 4) * 2 + sb_cumsum_temp_bytes(num_tiles_total);
}

extern "C" size_t sb_sort_tiles_global_temp_bytes(int count, int num_tiles_total) {
  return sb_sort32_temp_bytes(count) + tiles_global_extra(std::max(num_tiles_total, 0)) + 512;
}

extern "C" int sb_sort_tiles_global(int count, int num_tiles_total, const uint32_t* tile_ids, const int* gaussian_ids,
                                    int* gaussian_ids_sorted, int* tile_bins, const uint32_t* digit_hist,
                                    void* temp, size_t temp_bytes, sb_stream_t stream) {
  if (count < 0 || num_tiles_total < 1) return SB_ERR_BAD_ARG;
  if (!temp || temp_bytes < sb_sort_tiles_global_temp_bytes(count, num_tiles_total)) return SB_ERR_TEMP_TOO_SMALL;
  cudaStream_t st = (cudaStream_t)stream;
  int tile_bits = 1;
  while (tile_bits < 32 && (1ll << tile_bits) < (long long)num_tiles_total) ++tile_bits;
  temp = align_temp(temp, temp_bytes);
  unsigned char* base = reinterpret_cast<unsigned char*>(temp);
  int* counts = reinterpret_cast<int*>(base);
  int* incl = reinterpret_cast<int*>(base + sb::align256((size_t)num_tiles_total * 4));
  unsigned char* scan_tmp = base + 2 * sb::align256((size_t)num_tiles_total * 4);
  const size_t scan_bytes = sb_cumsum_temp_bytes(num_tiles_total);
  unsigned char* sort_tmp = scan_tmp + sb::align256(scan_bytes);
  const size_t used = (size_t)(sort_tmp - base);
  cudaError_t err = cudaMemsetAsync(counts, 0, (size_t)num_tiles_total * 4, st);
  if (err != cudaSuccess) return (int)err;
  if (count > 0) {
    const int rc = sb::radix_sort_pairs<uint32_t>((size_t)count, 0, tile_bits, tile_ids, gaussian_ids, nullptr,
                                                  gaussian_ids_sorted, sort_tmp, temp_bytes - used, true, counts,
                                                  digit_hist, st);
    if (rc != 0) return rc;
  }
  const int rc = sb_cumsum_i32(num_tiles_total, counts, incl, scan_tmp, scan_bytes, stream);
  if (rc != 0) return rc;
  sb::bins_from_counts_kernel<<<sb::ceil_div(num_tiles_total, 256), 256, 0, st>>>(num_tiles_total, counts, incl,
                                                                                reinterpret_cast<int2*>(tile_bins));
  SB_RETURN_LAST();
}

// ---- the prefix sum for callers that may not allocate ---------------------------------------------------------
// sb_cumsum_i32_static: sb_cumsum_i32 with the look-back state in a file-scope device array instead of a buffer
// the caller allocates (code under CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).  No
// counterpart in the reference.  The array is part of the library's
// image: every GPU that loads the library holds its 32 832 bytes until the process exits.  It is shared by every
// call on a device and keeps the last call's contents, so (a) the tile counter and the status words the call is
// going to use are cleared on the call's stream before the kernel, (b) two calls that overlap on the GPU (different
// streams without an ordering between them) race on it -- callers serialise.  Nothing here calls cudaMalloc /
// cudaFree or a routine that does; cudaGetSymbolAddress runs once per device (the first call on that device, which
// should therefore be made outside a stream capture).
namespace sb {

constexpr int kStaticScanMaxCount = SB_CUMSUM_STATIC_MAX_COUNT;
static_assert(kStaticScanMaxCount % kScanTile == 0, "whole tiles");
constexpr int kStaticScanMaxTiles = kStaticScanMaxCount / kScanTile;
constexpr size_t kStaticScanBytes = 64 + (size_t)kStaticScanMaxTiles * 8;        // [counter (64 B) | status words]
static __device__ __align__(256) unsigned char g_static_scan_scratch[kStaticScanBytes];

// In host code `g_static_scan_scratch` is a host-side stand-in for the symbol, not a device address: the kernel
// gets the address cudaGetSymbolAddress reports for the CURRENT device, looked up once per device.
constexpr int kMaxDevices = 64;
static unsigned char* volatile g_static_scan_ptr[kMaxDevices];                    // zero-initialised; same value from any thread

static int static_scan_scratch(unsigned char** out) {
  int dev = -1;
  cudaError_t err = cudaGetDevice(&dev);
  if (err != cudaSuccess) return (int)err;
  if (dev < 0 || dev >= kMaxDevices) return SB_ERR_BAD_ARG;
  unsigned char* p = g_static_scan_ptr[dev];
  if (!p) {
    void* sym = nullptr;
    err = cudaGetSymbolAddress(&sym, g_static_scan_scratch);
    if (err != cudaSuccess) return (int)err;
    if (!sym) return (int)cudaErrorInvalidSymbol;
    p = reinterpret_cast<unsigned char*>(sym);
    g_static_scan_ptr[dev] = p;
  }
  *out = p;
  return 0;
}

}  // namespace sb

extern "C" int sb_cumsum_i32_static(int count, const int* in, int* out, sb_stream_t stream) {
  if (count < 0) return SB_ERR_BAD_ARG;
  if (count > sb::kStaticScanMaxCount) return SB_ERR_STATIC_SCRATCH_EXCEEDED;   // before any CUDA call
  if (count == 0) return 0;
  if (!in || !out) return SB_ERR_BAD_ARG;
  unsigned char* base = nullptr;
  const int rc = sb::static_scan_scratch(&base);
  if (rc != 0) return rc;
  const unsigned tiles = (unsigned)(((size_t)count + sb::kScanTile - 1) / sb::kScanTile);
  cudaStream_t st = (cudaStream_t)stream;
  cudaError_t err = cudaMemsetAsync(base, 0, 64 + (size_t)tiles * 8, st);        // what must start at zero
  if (err != cudaSuccess) return (int)err;
  sb::scan_kernel<<<tiles, sb::kScanThreads, 0, st>>>(in, out, count, reinterpret_cast<unsigned long long*>(base + 64),
                                                      reinterpret_cast<unsigned*>(base));
  SB_RETURN_LAST();
}
