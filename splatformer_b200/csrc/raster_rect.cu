// Rect-list rasteriser (forward and backward) for the fused multi-view path, block width 16.
// Same arithmetic per (pixel, Gaussian) pair as raster_fwd.cu / raster_bwd.cu (SURVEY.md A.5 / A.6;
// upstream _C.rasterize_forward / _backward behind gsplat.rasterize_gaussians,
// /root/reference/utils/gs_utils.py:96-109, train.py:289) -- what changes is who walks which list.
//
// sb_sort_tiles has already split every tile's depth-sorted list into eight lists, one per 8x4-pixel
// rectangle, keeping only the Gaussians whose visible region reaches the rectangle (exact test,
// raster_common.cuh: rect_mask).  Here ONE WARP owns one rectangle and its list:
//   * no block-level barrier and no shared staging between warps: a warp gathers the 32 next records
//     of ITS list lane-parallel (one 48-byte record per lane) into registers while it composites the
//     current 32 from its private 1.5 KB of shared memory -- the gather latency hides behind the walk;
//   * no per-warp culling pass and no ballot / find-first-set loop: every list entry is a hit, the walk
//     is a plain counted loop over consecutive shared-memory slots;
//   * early termination is per warp (all 32 pixels saturated), not per 256-pixel tile.
// The forward is persistent (warps pull rectangles from a counter, next rectangle's loads in flight); the
// backward splits the rectangle once more into its two 4x4 halves, one per half-warp (see below).
#include <algorithm>

#include "raster_common.cuh"

namespace sb {

#ifndef SB_RECT_FWD_WARPS
#define SB_RECT_FWD_WARPS 8
#endif
constexpr int kRectWarps = SB_RECT_FWD_WARPS;     // forward: warps are independent, the CTA is launch granularity only
constexpr int kRectWarpsB = 4;    // backward: two CTAs per tile (shared-memory budget of the slot buffers)

struct RectGeom {
  int px_i, py_i;      // this lane's pixel
  bool inside;
  float px, py;        // its centre
  float ox, oy;        // centre of the rectangle's corner pixel (lane 0)
};

__device__ __forceinline__ RectGeom rect_geom(int tile_x, int tile_y, int warp, int lane, int img_w, int img_h) {
  RectGeom rg;
  const int x0 = tile_x * 16 + (warp & 1) * 8, y0 = tile_y * 16 + (warp >> 1) * 4;
  rg.px_i = x0 + (lane & 7);
  rg.py_i = y0 + (lane >> 3);
  rg.inside = rg.px_i < img_w && rg.py_i < img_h;
  rg.px = (float)rg.px_i + 0.5f; rg.py = (float)rg.py_i + 0.5f;
  rg.ox = (float)x0 + 0.5f; rg.oy = (float)y0 + 0.5f;
  return rg;
}

// Shared-memory layout of a staged chunk of 32 list entries: 16 PAIRS of consecutive entries (2j, 2j+1),
// interleaved so that one LDS.128 feeds both halves of the packed-FP32 instructions:
//   q0[j] = { x0 x1 | y0 y1 }   q1[j] = { A0 A1 | B0 B1 }   q2[j] = { C0 C1 | lop0 lop1 }
//   q3[j] = { r0 g0 | r1 g1 }   q4[j] = { b0 b1 }        (r, g accumulate with one packed FMA per entry)
// A list position past the end of the list holds lop = -inf-like, i.e. alpha = 0: never composited.
struct PairSlots {
  float4 q0[16], q1[16], q2[16], q3[16];
  float4 q4[16];     // .xy used; float4 stride keeps one index register for all five arrays
};
constexpr float kNullLop = -1.0e30f;

// lane = list entry of the chunk: scatter its record into the pair layout
__device__ __forceinline__ void stage_pair_slots(PairSlots& sl, int lane, const float4& rA, const float4& rB,
                                                 const float4& rC) {
  const int j = lane >> 1, h = lane & 1;
  float* f0 = reinterpret_cast<float*>(sl.q0 + j) + h;
  float* f1 = reinterpret_cast<float*>(sl.q1 + j) + h;
  float* f2 = reinterpret_cast<float*>(sl.q2 + j) + h;
  f0[0] = rA.x; f0[2] = rA.y;
  f1[0] = rB.x; f1[2] = rB.y;
  f2[0] = rB.z; f2[2] = rB.w;
  *reinterpret_cast<float2*>(reinterpret_cast<float*>(sl.q3 + j) + 2 * h) = make_float2(rC.x, rC.y);
  reinterpret_cast<float*>(sl.q4 + j)[h] = rC.z;
}

// ------------------------------------------------------------------------------------------------
// Forward: persistent warps over a queue of rectangles.
// A rectangle's list is short (46 entries on average at the benchmark shape: 1.5 chunks), so a warp
// that handled ONE rectangle spent more time in its dependent-load prologue (count -> ids -> records)
// and in its store / exit epilogue than in the compositing loop.  Here the grid is sized to the machine
// (resident CTAs per SM x SMs), every warp pulls rectangle numbers from a global counter and keeps a
// two-deep lookahead, so that all of the next rectangle's loads are in flight while the current one
// composites:
//     top of rectangle k      : counter fetch for k+3;  count and first 32 ids of k+2
//     last chunk of rectangle k: the 32 first records of k+1 (ids arrived one rectangle ago)
// `work` is one int, zero on entry (the caller's per-step zeroed state block).
// ------------------------------------------------------------------------------------------------
#ifndef SB_RECT_FWD_MINB
#define SB_RECT_FWD_MINB 3
#endif
#ifndef SB_RECT_FWD_UNROLL
#define SB_RECT_FWD_UNROLL 4
#endif
#define SB_PRAGMA_(x) _Pragma(#x)
#define SB_UNROLL_(n) SB_PRAGMA_(unroll n)

__device__ __forceinline__ int grab_rect(int* work, int lane) {
  int r = 0;
  if (lane == 0) r = atomicAdd(work, 1);
  return __shfl_sync(0xffffffffu, r, 0);
}
// the same in two halves: the atomic is issued where the number is wanted EARLY, the broadcast (which waits for the
// atomic's round trip to L2) where it is needed
// `opaque_zero` is a lane's value that is 0 at run time but that ptxas cannot prove to be: with a provably uniform
// address ptxas wraps the atomic in its warp-aggregation pattern (VOTE / elect / ATOMG / SHFL), and that SHFL waits
// for the atomic on the spot (ncu source page: 5 % of the kernel's samples).
__device__ __forceinline__ int grab_rect_issue(int* work, int lane, int opaque_zero) {
  int r = 0;
  if (lane == 0) asm volatile("atom.global.add.u32 %0, [%1], 1;" : "=r"(r) : "l"(work + opaque_zero) : "memory");
  return r;
}

// One record of the forward walk: centre (rA.xy), conic + log2 opacity (rB), colour (rC.xyz).  rC.w (the opacity)
// is not needed here but is STORED with the rest (stage_pair_slots_fwd): a dead component of a vector load is a
// register ptxas hands to the next temporary, and the write to it then waits for the load (write-after-write on the
// load's destination) -- the whole gather latency, in front of the walk that was meant to hide it (ncu source page:
// 9 % of the kernel's samples on one LOP3).  Splitting the load does not help: ptxas fuses the pieces again.
__device__ __forceinline__ void load_record_fwd(const float4* __restrict__ p, float4& rA, float4& rB, float4& rC) {
  const float2 c = __ldg(reinterpret_cast<const float2*>(p));
  rA.x = c.x; rA.y = c.y;
  rB = __ldg(p + 1);
  rC = __ldg(p + 2);
}
__device__ __forceinline__ void stage_pair_slots_fwd(PairSlots& sl, int lane, const float4& rA, const float4& rB,
                                                     const float4& rC) {
  stage_pair_slots(sl, lane, rA, rB, rC);
  reinterpret_cast<float*>(sl.q4 + (lane >> 1))[2 + (lane & 1)] = rC.w;       // keeps rC.w live until here
}

// A tile without a single Gaussian (70 % of the tiles on the benchmark scene; sort_tiles marks its eight rectangles
// with start = -1): the warp that drew rectangle 0 writes all 256 pixels -- background, T = 1, no last index, the loss
// against the target -- and the tile's loss sums go to rectangle 0's entry (the caller adds the entries up); the
// seven other warps have nothing to do.  Eight generic rectangle passes cost ~1600 warp instructions per empty
// tile, this ~300.
__device__ __forceinline__ void fill_empty_tile(int v, int tile_x, int tile_y, size_t rect0, int lane, int img_w, int img_h,
                                                const float* __restrict__ background,
                                                const int* __restrict__ view_has_isects, float* __restrict__ out_img,
                                                float* __restrict__ final_ts, int* __restrict__ final_idx,
                                                float* __restrict__ out_alpha, int clamp_rgb, const void* __restrict__ gt,
                                                int gt_u8, float* __restrict__ loss_partial) {
  const bool empty_view = view_has_isects ? __ldg(view_has_isects + v) == 0 : false;
  float r = __ldg(background), g = __ldg(background + 1), b = __ldg(background + 2);
  if (clamp_rgb) { r = fminf(r, 1.0f); g = fminf(g, 1.0f); b = fminf(b, 1.0f); }
  const float T_out = empty_view ? 0.f : 1.f, a_out = empty_view ? 1.f : 0.f;
  float l1 = 0.f, l2 = 0.f;
#pragma unroll
  for (int half = 0; half < 2; ++half) {
    // four rectangles at a time: their target pixels are requested before the first is used
    unsigned raw[4][3];
    size_t pix[4];
    bool in[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const RectGeom rg = rect_geom(tile_x, tile_y, 4 * half + k, lane, img_w, img_h);
      in[k] = rg.inside;
      pix[k] = ((size_t)v * img_h + rg.py_i) * img_w + rg.px_i;
      raw[k][0] = raw[k][1] = raw[k][2] = 0u;
      if (gt && in[k]) load_gt_raw(gt, gt_u8, pix[k], raw[k][0], raw[k][1], raw[k][2]);
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      if (!in[k]) continue;
      final_ts[pix[k]] = T_out;
      final_idx[pix[k]] = -1;
      if (out_alpha) out_alpha[pix[k]] = a_out;
      out_img[3 * pix[k]] = r; out_img[3 * pix[k] + 1] = g; out_img[3 * pix[k] + 2] = b;
      if (gt) {
        float tr, tg_, tb;
        gt_from_raw(gt_u8, raw[k][0], raw[k][1], raw[k][2], tr, tg_, tb);
        const float dr = r - tr, dg = g - tg_, db = b - tb;
        l1 += fabsf(dr) + fabsf(dg) + fabsf(db);
        l2 += dr * dr + dg * dg + db * db;
      }
    }
  }
  if (loss_partial) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      l1 += __shfl_xor_sync(0xffffffffu, l1, o);
      l2 += __shfl_xor_sync(0xffffffffu, l2, o);
    }
    if (lane < 8) *reinterpret_cast<float2*>(loss_partial + (rect0 + lane) * 2) = lane == 0 ? make_float2(l1, l2) : make_float2(0.f, 0.f);
  }
}

__global__ void __launch_bounds__(kRectWarps * 32, SB_RECT_FWD_MINB) raster_rect_fwd_kernel(
    int n, const float4* __restrict__ packed, const int* __restrict__ rect_ids,
    const int2* __restrict__ rect_meta, const float* __restrict__ background, int tiles_x,
    int tiles_y, int img_w, int img_h, const int* __restrict__ view_has_isects,
    float* __restrict__ out_img, float* __restrict__ final_ts, int* __restrict__ final_idx,
    float* __restrict__ out_alpha, int clamp_rgb, const void* __restrict__ gt, int gt_u8,
    float* __restrict__ loss_partial, int total_rects, int* __restrict__ work, FastDiv div_tiles,
    FastDiv div_tiles_x) {
  __shared__ PairSlots sSlots[kRectWarps];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int tiles = tiles_x * tiles_y;
  PairSlots& sl = sSlots[warp];
  const float4 nullA = make_float4(0.f, 0.f, 0.f, 0.f), nullB = make_float4(0.f, 0.f, 0.f, kNullLop);
  const int2 none = make_int2(0, 0);

  // lookahead state: rectangle k (r0: list at meta0 = (start, count), first records in rA..rC), k+1 (r1: meta1 and
  // its first ids id1), k+2 (r2: meta2), k+3 (r3).  The lists live in a pool (tile_sort.cu), so a rectangle's ids
  // can only be addressed once its (start, count) pair has arrived: one stage more than fixed-capacity lists had.
  // Rectangles are handed out ONE AT A TIME: the eight rectangles of a tile then run on neighbouring warps at the
  // same moment and share their Gaussians' records in L1 -- a warp that takes a whole tile per grab (one atomic
  // and one (start, count) load per eight rectangles) measured 0.75 ms instead of 0.48 ms.
  int r0 = grab_rect(work, lane);
  if (r0 >= total_rects) return;
  int r1 = grab_rect(work, lane);
  int r2 = grab_rect(work, lane);
  int r3 = grab_rect(work, lane);
  int2 meta0 = __ldg(rect_meta + r0);
  int2 meta1 = r1 < total_rects ? __ldg(rect_meta + r1) : none;
  int2 meta2 = r2 < total_rects ? __ldg(rect_meta + r2) : none;
  int id1 = lane < meta1.y ? __ldg(rect_ids + meta1.x + lane) : 0;
  float4 rA = nullA, rB = nullB, rC = nullA;
  if (lane < meta0.y) {
    const int v0 = fast_div(r0 >> 3, div_tiles);
    load_record_fwd(packed + ((size_t)v0 * n + __ldg(rect_ids + meta0.x + lane)) * 3, rA, rB, rC);
  }

  while (true) {
    // ---- lookahead loads: ids of k+2, (start, count) of k+3, counter fetch for k+4 ----
    const int id2 = lane < meta2.y ? __ldg(rect_ids + meta2.x + lane) : 0;
    const int2 meta3 = r3 < total_rects ? __ldg(rect_meta + r3) : none;
    const int r4_lane0 = grab_rect_issue(work, lane, (int)((unsigned)(meta1.y | id1) >> 31));      // broadcast at the bottom of the rectangle
    const int count1 = meta1.y;
    // ids of this rectangle's SECOND chunk: the records of chunk c + 1 are requested at the top of chunk c, so
    // their ids must be in registers by then -- one chunk ahead of the records, two ahead of the walk
    int idn = 32 + lane < meta0.y ? __ldg(rect_ids + meta0.x + 32 + lane) : 0;

    const int tg = r0 >> 3, wt = r0 & 7;
    const int v = fast_div(tg, div_tiles), tile = tg - v * tiles;
    const int tile_y = fast_div(tile, div_tiles_x), tile_x = tile - tile_y * tiles_x;
    const float4* vp1 = packed + (size_t)fast_div(r1 >> 3, div_tiles) * n * 3;      // only used when r1 is valid
    if (meta0.x < 0) {
      // a rectangle of an EMPTY TILE: rectangle 0's warp writes the whole tile, the others only keep the queue going
      rB = nullB;
      if (lane < count1) load_record_fwd(vp1 + (size_t)id1 * 3, rA, rB, rC);
      if (wt == 0)
        fill_empty_tile(v, tile_x, tile_y, (size_t)r0, lane, img_w, img_h, background, view_has_isects, out_img, final_ts,
                        final_idx, out_alpha, clamp_rgb, gt, gt_u8, loss_partial);
    } else {
    const RectGeom rg = rect_geom(tile_x, tile_y, wt, lane, img_w, img_h);
    const int count = meta0.y;
    const int* ids = rect_ids + meta0.x;
    const float4* vp = packed + (size_t)v * n * 3;

    float T = 1.0f, pb = 0.f;
    unsigned long long prg = pk2(0.f, 0.f);        // (red, green) accumulators, one FFMA2 per entry
    int cur_idx = -1;
    bool done = !rg.inside;
    const unsigned long long px2 = pk2(rg.px, rg.px), py2 = pk2(rg.py, rg.py);
    bool fetched_next = false;

    // The loss epilogue's target pixel: its load sat exposed at the end of every rectangle (5 % of the samples).
    // A rectangle with a list asks for the line before the walk and loads it after; an EMPTY one (77 % of them on
    // the benchmark scene) has no walk to hide anything behind and no register pressure either: it loads at once.
    // Two separate branches, so that the early-loaded pixel is not live across the walk.
    const bool want_gt = gt && rg.inside;
    const size_t gt_pix = ((size_t)v * img_h + rg.py_i) * img_w + rg.px_i;
    unsigned gt0 = 0u, gt1 = 0u, gt2 = 0u;
    if (count == 0) {
      if (want_gt) load_gt_raw(gt, gt_u8, gt_pix, gt0, gt1, gt2);
      rB = nullB;
      if (lane < count1) load_record_fwd(vp1 + (size_t)id1 * 3, rA, rB, rC);
    } else {
    if (want_gt) prefetch_gt(gt, gt_u8, gt_pix);
    for (int c0 = 0; c0 < count; c0 += 32) {
      __syncwarp();
      stage_pair_slots_fwd(sl, lane, rA, rB, rC);
      __syncwarp();
      rB = nullB;
      if (c0 + 32 < count) {
        if (c0 + 32 + lane < count) load_record_fwd(vp + (size_t)idn * 3, rA, rB, rC);
        if (c0 + 64 + lane < count) idn = __ldg(ids + c0 + 64 + lane);
      } else {
        // last chunk of this rectangle: the registers now take the first chunk of the next one
        fetched_next = true;
        if (lane < count1) load_record_fwd(vp1 + (size_t)id1 * 3, rA, rB, rC);
      }
      const int npairs = (min(32, count - c0) + 1) >> 1;
      SB_UNROLL_(SB_RECT_FWD_UNROLL)
      for (int j = 0; j < npairs; ++j) {
        const float4 Q0 = sl.q0[j], Q1 = sl.q1[j], Q2 = sl.q2[j];
        // both entries of the pair at once; every half rounds exactly like the scalar expression
        //   pw = dx * fma(B, dy, A * dx) + (C * dy) * dy,   log2(alpha) = pw + lop
        const unsigned long long dx2 = sub2(pk2(Q0.x, Q0.y), px2), dy2 = sub2(pk2(Q0.z, Q0.w), py2);
        const unsigned long long t2 = fma2(pk2(Q1.z, Q1.w), dy2, mul2(pk2(Q1.x, Q1.y), dx2));
        const unsigned long long pw2 = fma2(dx2, t2, mul2(mul2(pk2(Q2.x, Q2.y), dy2), dy2));
        const unsigned long long e2 = add2(pw2, pk2(Q2.z, Q2.w));
        float pw0, pw1, e0, e1;
        upk2(pw2, pw0, pw1);
        upk2(e2, e0, e1);
        const float4 Q3 = sl.q3[j];
        const float2 Q4 = *reinterpret_cast<const float2*>(sl.q4 + j);
        {
          const float alpha = fminf(kAlphaMax, ex2_approx(e0));
          const bool valid = !done && (pw0 <= 0.f) && (alpha >= kAlphaMin);
          const float next_T = T * (1.0f - alpha);
          const bool stop = valid && (next_T <= kTMin);
          const bool upd = valid && !stop;
          done = done || stop;
          const float wgt = upd ? alpha * T : 0.f;
          prg = fma2(pk2(Q3.x, Q3.y), pk2(wgt, wgt), prg); pb = fmaf(Q4.x, wgt, pb);
          T = upd ? next_T : T;
          cur_idx = upd ? c0 + 2 * j : cur_idx;
        }
        {
          const float alpha = fminf(kAlphaMax, ex2_approx(e1));
          const bool valid = !done && (pw1 <= 0.f) && (alpha >= kAlphaMin);
          const float next_T = T * (1.0f - alpha);
          const bool stop = valid && (next_T <= kTMin);
          const bool upd = valid && !stop;
          done = done || stop;
          const float wgt = upd ? alpha * T : 0.f;
          prg = fma2(pk2(Q3.z, Q3.w), pk2(wgt, wgt), prg); pb = fmaf(Q4.y, wgt, pb);
          T = upd ? next_T : T;
          cur_idx = upd ? c0 + 2 * j + 1 : cur_idx;
        }
      }
      if (__all_sync(0xffffffffu, done)) break;
    }
    if (!fetched_next) {
      // every pixel saturated before the last chunk
      rB = nullB;
      if (lane < count1) load_record_fwd(vp1 + (size_t)id1 * 3, rA, rB, rC);
    }
    if (want_gt) load_gt_raw(gt, gt_u8, gt_pix, gt0, gt1, gt2);
    }   // count > 0

    float pr, pg;
    upk2(prg, pr, pg);
    float l1 = 0.f, l2 = 0.f;
    if (rg.inside) {
      const size_t pix = ((size_t)v * img_h + rg.py_i) * img_w + rg.px_i;
      // upstream quirk: a view without a single intersection reports final_T = 0
      // (alpha = 1) while its image is the plain background
      const bool empty_view = view_has_isects ? __ldg(view_has_isects + v) == 0 : false;
      float r = pr + T * __ldg(background), g = pg + T * __ldg(background + 1), b = pb + T * __ldg(background + 2);
      if (clamp_rgb) { r = fminf(r, 1.0f); g = fminf(g, 1.0f); b = fminf(b, 1.0f); }   // gs_utils.py:111
      final_ts[pix] = empty_view ? 0.f : T;
      final_idx[pix] = cur_idx;
      if (out_alpha) out_alpha[pix] = empty_view ? 1.f : 1.0f - T;
      out_img[3 * pix] = r; out_img[3 * pix + 1] = g; out_img[3 * pix + 2] = b;
      if (gt) {
        float gt_r, gt_g, gt_b;
        gt_from_raw(gt_u8, gt0, gt1, gt2, gt_r, gt_g, gt_b);
        const float dr = r - gt_r, dg = g - gt_g, db = b - gt_b;
        l1 = fabsf(dr) + fabsf(dg) + fabsf(db);
        l2 = dr * dr + dg * dg + db * db;
      }
    }
    if (loss_partial) {
      // fused L1 / squared-error partial sums (train.py:275,278): one deterministic (V, T, 8, 2) entry
      // per rectangle, summed by the caller
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        l1 += __shfl_xor_sync(0xffffffffu, l1, o);
        l2 += __shfl_xor_sync(0xffffffffu, l2, o);
      }
      if (lane == 0) *reinterpret_cast<float2*>(loss_partial + (size_t)r0 * 2) = make_float2(l1, l2);
    }
    }   // a tile with Gaussians

    const int r4 = __shfl_sync(0xffffffffu, r4_lane0, 0);
    if (r1 >= total_rects) break;
    r0 = r1; meta0 = meta1;
    r1 = r2; meta1 = meta2; id1 = id2;
    r2 = r3; meta2 = meta3;
    r3 = r4;
  }
}

// ------------------------------------------------------------------------------------------------
// Backward, split walk: the two half-warps own the two 4x4 halves of the rectangle (lanes 0-15
// the left one, lanes 16-31 the right one) and walk THEIR OWN sub-lists in lockstep.  Only 37 % of the
// (rectangle, Gaussian) entries reach both halves, so max(|L|, |R|) is 0.72 x the entries
// (scripts/half_rect_stats.py).  Per 32-entry chunk: every lane tests its entry against both halves
// (the exact test of rect_mask on a 4x4 rectangle), two ballots give the sub-list positions, the record is
// scattered into the half's pair layout back to front (slot 0 = deepest entry) together with its list
// index, unused slots up to the longer sub-list get null records.  A step then evaluates slot s of the left
// sub-list in lanes 0-15 and slot s of the right one in lanes 16-31; a parked slot holds two independent
// (Gaussian, half) sums of 16 pixels each, which flush_half_slots reduces without any cross-half shuffle.
// ------------------------------------------------------------------------------------------------
struct SplitLists { PairSlots half[2]; };

// conservative: does {alpha >= 1/255} reach the 4x4 pixel centres starting at (ox, oy) / (ox + 4, oy)?
__device__ __forceinline__ void half_hits(const float4& A, const float4& B, float ox, float oy, bool& hit_l,
                                          bool& hit_r) {
  hit_l = hit_r = false;
  if (A.z < 0.f) return;                                      // opacity < 1/255: never visible
  const float qa = -B.x, qb = -B.y, qc = -B.z;
  hit_l = hit_r = true;
  if (A.z > 1.0e37f || !(qa > 0.f) || !(qc > 0.f)) return;    // not an ellipse: no culling
  const float qmax = (B.w + 7.994353f) * 1.0001f + 1e-3f;
  const float rb = __fdividef(-0.5f * qb, qc), ra = __fdividef(-0.5f * qb, qa);
  const float ly = oy - A.y, hy = ly + 3.0f;
  const float dyc = fminf(fmaxf(0.f, ly), hy);
  const float tc = qc * dyc * dyc, rd = ra * dyc, tb = qb * dyc;
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    const float lx = ox + 4.0f * j - A.x, hx = lx + 3.0f;
    const float dxc = fminf(fmaxf(0.f, lx), hx);
    const float dy1 = fminf(fmaxf(rb * dxc, ly), hy);
    const float q1 = fmaf(dxc, fmaf(qb, dy1, qa * dxc), qc * dy1 * dy1);
    const float dx2 = fminf(fmaxf(rd, lx), hx);
    const float q2 = fmaf(dx2, fmaf(qa, dx2, tb), tc);
    const bool hit = !(fminf(q1, q2) > qmax);                 // (a NaN keeps the pair)
    if (j == 0) hit_l = hit; else hit_r = hit;
  }
}

__device__ __forceinline__ void stage_slot_at(PairSlots& sl, int s, const float4& rA, const float4& rB, float lop,
                                              float cr, float cg, float cb, int idx) {
  const int j = s >> 1, h = s & 1;
  float* f0 = reinterpret_cast<float*>(sl.q0 + j) + h;
  float* f1 = reinterpret_cast<float*>(sl.q1 + j) + h;
  float* f2 = reinterpret_cast<float*>(sl.q2 + j) + h;
  f0[0] = rA.x; f0[2] = rA.y;
  f1[0] = rB.x; f1[2] = rB.y;
  f2[0] = rB.z; f2[2] = lop;
  *reinterpret_cast<float2*>(reinterpret_cast<float*>(sl.q3 + j) + 2 * h) = make_float2(cr, cg);
  float* f4 = reinterpret_cast<float*>(sl.q4 + j);
  f4[h] = cb; f4[2 + h] = __int_as_float(idx);
}

// 16 parked steps x 2 halves: lane (sj, h) sums the 16 pixels of half h of step sj (one Gaussian)
__device__ __forceinline__ void flush_half_slots(const float2* __restrict__ myW, const float4* __restrict__ myVo,
                                                 int nslot, int lane, float own_x, float own_y, int own_g,
                                                 float ox, float oy, float* __restrict__ vgrad) {
  __syncwarp();
  const int sj = lane & 15, h = lane >> 4;
  const float2* row = myW + sj * kSlotStride + h * 16;
  const float4* vo = myVo + h * 16;
  float m0 = 0.f, sx = 0.f, sy = 0.f, sxx = 0.f, sxy = 0.f, syy = 0.f, cr = 0.f, cg = 0.f, cb = 0.f;
#pragma unroll
  for (int t = 0; t < 16; ++t) {
    const float2 wf = row[t];
    const float4 o = vo[t];
    const float ix = (float)(t & 3), iy = (float)(t >> 2);
    m0 += wf.x;
    if (t & 3) { sx = fmaf(wf.x, ix, sx); sxx = fmaf(wf.x, ix * ix, sxx); }
    if (t >> 2) { sy = fmaf(wf.x, iy, sy); syy = fmaf(wf.x, iy * iy, syy); }
    if ((t & 3) && (t >> 2)) sxy = fmaf(wf.x, ix * iy, sxy);
    cr = fmaf(wf.y, o.x, cr); cg = fmaf(wf.y, o.y, cg); cb = fmaf(wf.y, o.z, cb);
  }
  const float ax = own_x - (ox + (float)(4 * h)), ay = own_y - oy;
  const float m1x = ax * m0 - sx;
  const float m1y = ay * m0 - sy;
  const float m2xx = ax * (m1x - sx) + sxx;
  const float m2xy = ay * m1x - ax * sy + sxy;
  const float m2yy = ay * (m1y - sy) + syy;
  if (sj < nslot && own_g >= 0) {
    float* dst = vgrad + (size_t)own_g * 12;
    red_add_v4(dst, m1x, m1y, m2xx, m2xy);
    red_add_v4(dst + 4, m2yy, m0, cr, cg);
    if (cb != 0.f) atomicAdd(dst + 8, cb);
  }
  __syncwarp();
}

#ifndef SB_RECT_BWD_MINB
#define SB_RECT_BWD_MINB 6      // 80 registers without spills; 36 KB of shared memory per CTA also caps it at 6
#endif
__global__ void __launch_bounds__(kRectWarpsB * 32, SB_RECT_BWD_MINB) raster_rect_bwd_kernel(
    int n, const float4* __restrict__ packed, const int* __restrict__ rect_ids,
    const int2* __restrict__ rect_meta, const float* __restrict__ background, int tiles_x,
    int tiles_y, int img_w, int img_h, int clamped, const float* __restrict__ final_ts,
    const int* __restrict__ final_idx, const float* __restrict__ v_out_img,
    const float* __restrict__ v_out_alpha, const float* __restrict__ out_img,
    const void* __restrict__ gt, int gt_u8, const float* __restrict__ v_loss, float* __restrict__ v_packed,
    FastDiv div_tiles_x) {
  __shared__ SplitLists sLists[kRectWarpsB];
  __shared__ int sGid[kRectWarpsB][2][32];
  __shared__ float4 sVo[kRectWarpsB][32];
  __shared__ float2 sW[kRectWarpsB][kSlots * kSlotStride];
  __shared__ float4 sRaw[kRectWarpsB][32][3];      // next chunk's raw records, filled by cp.async
  constexpr int kParts = 8 / kRectWarpsB;
  const int tile = blockIdx.x / kParts, v = blockIdx.y;
  const int tile_y = fast_div(tile, div_tiles_x), tile_x = tile - tile_y * tiles_x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int wt = warp + (blockIdx.x % kParts) * kRectWarpsB;        // rectangle index within the tile
  const size_t rect = ((size_t)v * tiles_x * tiles_y + tile) * 8 + wt;
  const int2 meta = __ldg(rect_meta + rect);
  if (meta.y <= 0) return;
  // pixel of this lane: lanes 0-15 the left 4x4 half of the rectangle (row-major), lanes 16-31 the right one
  const int hh = lane >> 4;
  const int x0 = tile_x * 16 + (wt & 1) * 8, y0 = tile_y * 16 + (wt >> 1) * 4;
  const int px_i = x0 + 4 * hh + (lane & 3), py_i = y0 + ((lane >> 2) & 3);
  const bool inside = px_i < img_w && py_i < img_h;
  const float fpx = (float)px_i + 0.5f, fpy = (float)py_i + 0.5f;
  const float ox = (float)x0 + 0.5f, oy = (float)y0 + 0.5f;
  const int* ids = rect_ids + meta.x;
  const float4* vp = packed + (size_t)v * n * 3;
  float* vgrad = v_packed + (size_t)v * n * 12;
  PairSlots& listL = sLists[warp].half[0];
  PairSlots& listR = sLists[warp].half[1];
  const PairSlots& my = sLists[warp].half[hh];
  int* sGl = sGid[warp][0];
  int* sGr = sGid[warp][1];
  const int* sGm = sGid[warp][hh];

  float T_final = 1.f, vo_r = 0.f, vo_g = 0.f, vo_b = 0.f, vo_a = 0.f;
  int last = -1;
  if (inside) {
    const size_t pix = ((size_t)v * img_h + py_i) * img_w + px_i;
    T_final = __ldg(final_ts + pix);
    last = __ldg(final_idx + pix);
    if (v_out_img) {
      vo_r = __ldg(v_out_img + 3 * pix); vo_g = __ldg(v_out_img + 3 * pix + 1); vo_b = __ldg(v_out_img + 3 * pix + 2);
    }
    if (v_out_alpha) vo_a = __ldg(v_out_alpha + pix);
    if (gt) {
      // fused L1 loss (train.py:275): d|x - gt|/dx = sign(x - gt), scaled by the upstream gradient of
      // this view's loss term
      const float sc = __ldg(v_loss + v);
      float tr, tg, tb;
      load_gt(gt, gt_u8, pix, tr, tg, tb);
      const float dr = __ldg(out_img + 3 * pix) - tr;
      const float dg = __ldg(out_img + 3 * pix + 1) - tg;
      const float db = __ldg(out_img + 3 * pix + 2) - tb;
      vo_r += dr > 0.f ? sc : (dr < 0.f ? -sc : 0.f);
      vo_g += dg > 0.f ? sc : (dg < 0.f ? -sc : 0.f);
      vo_b += db > 0.f ? sc : (db < 0.f ? -sc : 0.f);
    }
    if (out_img && clamped) {
      // the forward stored min(rgb, 1) (gs_utils.py:111): no gradient through a clamped channel
      if (!(__ldg(out_img + 3 * pix) < 1.0f)) vo_r = 0.f;
      if (!(__ldg(out_img + 3 * pix + 1) < 1.0f)) vo_g = 0.f;
      if (!(__ldg(out_img + 3 * pix + 2) < 1.0f)) vo_b = 0.f;
    }
  }
  const int warp_last = __reduce_max_sync(0xffffffffu, last);
  if (warp_last < 0) return;                      // nothing was composited in this rectangle
  sVo[warp][lane] = make_float4(vo_r, vo_g, vo_b, 0.f);
  const float bg_dot = __ldg(background) * vo_r + __ldg(background + 1) * vo_g + __ldg(background + 2) * vo_b;
  // v_alpha = T * (rgb . v_out) - ra * acc,  acc = sum_{behind} fac_j (rgb_j . v_out) - T_final (v_out_alpha - bg . v_out)
  float T = T_final;
  float acc = -T_final * (vo_a - bg_dot);
  int nslot = 0;
  float own_x = 0.f, own_y = 0.f;
  int own_g = -1;

  const unsigned long long px2 = pk2(fpx, fpx), py2 = pk2(fpy, fpy);
  const float4 nullA = make_float4(0.f, 0.f, 0.f, 0.f), nullB = make_float4(0.f, 0.f, 0.f, kNullLop);
  // The records of the next chunk travel global -> shared by cp.async (no registers held across the
  // walk); their ids are read two chunks ahead so that the copy's address is known when it is issued.
  int c = warp_last >> 5;
  float4* raw = sRaw[warp][lane];
  auto fetch = [&](int g) {
    const float4* p = vp + (size_t)g * 3;
    cp_async16(raw, p); cp_async16(raw + 1, p + 1); cp_async16(raw + 2, p + 2);
  };
  int g_cur = 0, g_nxt = 0;
  if (32 * c + lane <= warp_last) { g_cur = __ldg(ids + 32 * c + lane); fetch(g_cur); }
  if (c > 0) g_nxt = __ldg(ids + 32 * (c - 1) + lane);
  const unsigned wbase = (unsigned)__cvta_generic_to_shared(&sW[warp][lane]);
  unsigned wcur = wbase;
  // one step, back to front: transmittance / accumulator update and parking of (w, fac)
  auto entry = [&](int idx, float pw, float e, float cr, float cg, float cb, float gx, float gy, int slot) {
    const float alpha_raw = ex2_approx(e);
    const float alpha = fminf(kAlphaMax, alpha_raw);
    const bool valid = (idx <= last) && (pw <= 0.f) && (alpha >= kAlphaMin);
    // a lane that skips the entry runs the same arithmetic with alpha = 0 (rcp(1) = 1: T, acc unchanged)
    const float a = valid ? alpha : 0.f;
    const float ra = rcp_approx(1.0f - a);
    T *= ra;
    const float fac = a * T;
    const float cdot = cr * vo_r + cg * vo_g + cb * vo_b;
    const float v_alpha = T * cdot - ra * acc;
    acc = fmaf(fac, cdot, acc);
    const float aw = (valid && alpha_raw <= kAlphaMax) ? alpha_raw : 0.f;
    const float w = -aw * v_alpha;
    asm volatile("st.shared.v2.f32 [%0], {%1, %2};" ::"r"(wcur), "f"(w), "f"(fac) : "memory");
    wcur += kSlotStride * (unsigned)sizeof(float2);
    if ((lane & 15) == nslot + (slot & 1)) { own_x = gx; own_y = gy; own_g = sGm[slot]; }
  };
  const unsigned lt = (1u << lane) - 1u;
  for (; c >= 0; --c) {
    const int top = min(31, warp_last - 32 * c);
    cp_async_wait_all();
    float4 rA = nullA, rB = nullB, rC = nullA;
    if (lane <= top) { rA = raw[0]; rB = raw[1]; rC = raw[2]; }
    const int rG = g_cur;
    bool hl = false, hr = false;
    if (lane <= top) half_hits(rA, rB, ox, oy, hl, hr);
    const unsigned ml = __ballot_sync(0xffffffffu, hl), mr = __ballot_sync(0xffffffffu, hr);
    const int nl = __popc(ml), nr = __popc(mr);
    const int npairs = (max(nl, nr) + 1) >> 1;
    __syncwarp();
    {
      // hit: slot counted from the back of the sub-list; no hit: one of the unused slots gets a null record
      const int sl_ = hl ? nl - 1 - __popc(ml & lt) : nl + __popc(~ml & lt);
      if (sl_ < 2 * npairs) {
        stage_slot_at(listL, sl_, rA, rB, hl ? rB.w : kNullLop, hl ? rC.x : 0.f, hl ? rC.y : 0.f, hl ? rC.z : 0.f,
                      32 * c + lane);
        sGl[sl_] = hl ? rG : -1;
      }
      const int sr_ = hr ? nr - 1 - __popc(mr & lt) : nr + __popc(~mr & lt);
      if (sr_ < 2 * npairs) {
        stage_slot_at(listR, sr_, rA, rB, hr ? rB.w : kNullLop, hr ? rC.x : 0.f, hr ? rC.y : 0.f, hr ? rC.z : 0.f,
                      32 * c + lane);
        sGr[sr_] = hr ? rG : -1;
      }
    }
    __syncwarp();
    g_cur = g_nxt;
    if (c > 0) fetch(g_cur);
    if (c > 1) g_nxt = __ldg(ids + 32 * (c - 2) + lane);
    for (int j = 0; j < npairs; ++j) {
      const float4 Q0 = my.q0[j], Q1 = my.q1[j], Q2 = my.q2[j];
      const unsigned long long dx2 = sub2(pk2(Q0.x, Q0.y), px2), dy2 = sub2(pk2(Q0.z, Q0.w), py2);
      const unsigned long long t2 = fma2(pk2(Q1.z, Q1.w), dy2, mul2(pk2(Q1.x, Q1.y), dx2));
      const unsigned long long pw2 = fma2(dx2, t2, mul2(mul2(pk2(Q2.x, Q2.y), dy2), dy2));
      const unsigned long long e2 = add2(pw2, pk2(Q2.z, Q2.w));
      float pw0, pw1, e0, e1;
      upk2(pw2, pw0, pw1);
      upk2(e2, e0, e1);
      const float4 Q3 = my.q3[j];
      const float4 Q4 = my.q4[j];
      entry(__float_as_int(Q4.z), pw0, e0, Q3.x, Q3.y, Q4.x, Q0.x, Q0.z, 2 * j);
      entry(__float_as_int(Q4.w), pw1, e1, Q3.z, Q3.w, Q4.y, Q0.y, Q0.w, 2 * j + 1);
      // two steps per iteration: the slot count stays even, one flush check per pair
      nslot += 2;
      if (nslot == kSlots) {
        flush_half_slots(sW[warp], sVo[warp], kSlots, lane, own_x, own_y, own_g, ox, oy, vgrad);
        nslot = 0;
        wcur = wbase;
      }
    }
  }
  if (nslot) flush_half_slots(sW[warp], sVo[warp], nslot, lane, own_x, own_y, own_g, ox, oy, vgrad);
}

}  // namespace sb

namespace {
// resident CTAs of a persistent kernel on the current device (cached per device)
template <typename K>
int persistent_grid(K kernel, int threads, int* cache) {
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) dev = 0;
  if (cache[dev] == 0) {
    int per_sm = 0, sms = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, threads, 0);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cache[dev] = per_sm > 0 && sms > 0 ? per_sm * sms : 1;
  }
  return cache[dev];
}
}  // namespace

extern "C" int sb_rasterize_rect_fwd(int num_views, int n, const float* packed, const int* rect_ids,
                                     const int* rect_meta, const float* background,
                                     int img_h, int img_w, const int* view_has_isects, float* out_img,
                                     float* final_ts, int* final_idx, float* out_alpha,
                                     int clamp_rgb_max1, const void* gt, int gt_is_u8,
                                     float* loss_partial, int* work_counter, sb_stream_t stream) {
  if (num_views < 1 || n < 0 || img_h < 1 || img_w < 1 || !rect_meta || (loss_partial && !gt) ||
      !work_counter)
    return SB_ERR_BAD_ARG;
  const int tiles_x = sb::ceil_div(img_w, 16), tiles_y = sb::ceil_div(img_h, 16);
  const long long total = (long long)num_views * tiles_x * tiles_y * 8;
  if (total > 0x7fffffffLL - (1 << 24)) return SB_ERR_BAD_ARG;
  static int grid_cache[64];
  const int resident = persistent_grid(sb::raster_rect_fwd_kernel, sb::kRectWarps * 32, grid_cache);
  const int grid = (int)std::min<long long>(resident, (total + sb::kRectWarps - 1) / sb::kRectWarps);
  sb::raster_rect_fwd_kernel<<<grid, sb::kRectWarps * 32, 0, (cudaStream_t)stream>>>(
      n, reinterpret_cast<const float4*>(packed), rect_ids, reinterpret_cast<const int2*>(rect_meta), background, tiles_x,
      tiles_y, img_w, img_h, view_has_isects, out_img, final_ts, final_idx, out_alpha, clamp_rgb_max1, gt,
      gt_is_u8, loss_partial, (int)total, work_counter, sb::make_fast_div(tiles_x * tiles_y), sb::make_fast_div(tiles_x));
  SB_RETURN_LAST();
}

extern "C" int sb_rasterize_rect_bwd(int num_views, int n, const float* packed, const int* rect_ids,
                                     const int* rect_meta, const float* background,
                                     int img_h, int img_w, const float* final_ts, const int* final_idx,
                                     const float* v_out_img, const float* v_out_alpha,
                                     const float* out_img, int clamped_rgb_max1, const void* gt,
                                     int gt_is_u8, const float* v_loss, float* v_packed,
                                     sb_stream_t stream) {
  if (num_views < 1 || n < 0 || img_h < 1 || img_w < 1 || !rect_meta ||
      (!v_out_img && !gt && !v_out_alpha) || (gt && (!out_img || !v_loss)) || (clamped_rgb_max1 && !out_img))
    return SB_ERR_BAD_ARG;
  const int tiles_x = sb::ceil_div(img_w, 16), tiles_y = sb::ceil_div(img_h, 16);
  dim3 grid(tiles_x * tiles_y * (8 / sb::kRectWarpsB), num_views);
  sb::raster_rect_bwd_kernel<<<grid, sb::kRectWarpsB * 32, 0, (cudaStream_t)stream>>>(
      n, reinterpret_cast<const float4*>(packed), rect_ids, reinterpret_cast<const int2*>(rect_meta), background, tiles_x,
      tiles_y, img_w, img_h, clamped_rgb_max1, final_ts, final_idx, v_out_img, v_out_alpha, out_img, gt,
      gt_is_u8, v_loss, v_packed, sb::make_fast_div(tiles_x));
  SB_RETURN_LAST();
}
