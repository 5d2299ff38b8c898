// Rasterise backward: reverse compositing (upstream _C.rasterize_backward behind
// gsplat.rasterize_gaussians' autograd, /root/reference/utils/gs_utils.py:96-109, train.py:289).
//
// Same CTA/warp decomposition as the forward kernel, walking each tile's list back to front.
// Per (pixel, Gaussian) pair the walk produces two numbers: w = dL/dsigma and fac = alpha * T (the
// weight of the colour gradient).  What has to leave the warp are nine sums per (warp, Gaussian):
// the moments {w, w dx, w dy, w dx^2, w dx dy, w dy^2} and fac * v_out[rgb] (raster_common.cuh
// explains how the consumers turn moments into v_xy / v_conic / v_opacity).
//
// TRANSPOSED REDUCTION.  A butterfly reduction per hit costs ~55 issue slots for 9 values.  Instead
// every lane parks its (w, fac) pair of a hit in a per-warp shared-memory slot; once 16 hits are
// parked the warp transposes roles: lane j takes slot (j & 15) and sums that hit over the 16 pixels
// of half (j >> 4) of the warp's 8x4 rectangle, with the pixel offsets as compile-time immediates
// (raw moments about the rectangle corner, shifted to the Gaussian centre afterwards).  One
// shuffle step joins the halves and the 16 owners issue two red.global.add.v4.f32 + one scalar red
// each.  Cost: ~14 issue slots per hit.
#include "raster_common.cuh"

namespace sb {

// NW warps per CTA: 8 = one CTA per tile; 4 = two CTAs per 16x16 tile (upper / lower half), which halves
// the number of warps that wait for each other at the staging barrier.  Each CTA stages NW*32 Gaussians
// per batch.
constexpr size_t bwd_smem(int nw) {
  return sizeof(float4) * (3 * nw * 32 + nw * 32) + sizeof(float2) * nw * kSlots * kSlotStride + sizeof(int) * nw * 32;
}


template <bool RECT, int NW>
__global__ void __launch_bounds__(NW * 32) raster_bwd_kernel(
    int n, const float4* __restrict__ packed, const int* __restrict__ ids_sorted,
    const int2* __restrict__ tile_bins, const float* __restrict__ background, int tiles_x,
    int tiles_y, int img_w, int img_h, int bw, int clamped, const float* __restrict__ final_ts,
    const int* __restrict__ final_idx, const float* __restrict__ v_out_img,
    const float* __restrict__ v_out_alpha, const float* __restrict__ out_img_clamped,
    const void* __restrict__ gt, int gt_u8, const float* __restrict__ v_loss, float* __restrict__ v_packed) {
  extern __shared__ float4 smem4[];
  constexpr int kB = NW * 32;      // Gaussians staged per batch
  constexpr int kParts = 8 / NW;   // CTAs per tile
  float4* sA = smem4;
  float4* sB = sA + kB;
  float4* sC = sB + kB;
  float4* sVo = sC + kB;                                       // [NW][32]  v_out rgb per pixel
  float2* sW = reinterpret_cast<float2*>(sVo + NW * 32);       // [NW][kSlots][kSlotStride]
  int* sG = reinterpret_cast<int*>(sW + NW * kSlots * kSlotStride);
  const int tile = blockIdx.x / kParts, part = blockIdx.x % kParts, v = blockIdx.y;
  const int tile_y = tile / tiles_x, tile_x = tile - tile_y * tiles_x;
  const PixelMap pm = make_pixel_map(tile_x, tile_y, bw, img_w, img_h, part * NW);
  const float px = (float)pm.px_i + 0.5f, py = (float)pm.py_i + 0.5f;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int warp_t = warp + part * NW;                          // warp index within the tile
  const int2 range = __ldg(tile_bins + (size_t)v * tiles_x * tiles_y + tile);
  const int count = range.y - range.x;
  if (count <= 0) return;
  const float4* vp = packed + (size_t)v * n * 3;
  float* vgrad = v_packed + (size_t)v * n * 12;

  float T_final = 1.f, vo_r = 0.f, vo_g = 0.f, vo_b = 0.f, vo_a = 0.f;
  int bin_final = -1;
  if (pm.inside) {
    const size_t pix = ((size_t)v * img_h + pm.py_i) * img_w + pm.px_i;
    T_final = __ldg(final_ts + pix);
    bin_final = __ldg(final_idx + pix);
    if (v_out_img) {
      vo_r = __ldg(v_out_img + 3 * pix); vo_g = __ldg(v_out_img + 3 * pix + 1); vo_b = __ldg(v_out_img + 3 * pix + 2);
    }
    if (v_out_alpha) vo_a = __ldg(v_out_alpha + pix);
    if (gt) {
      // fused L1 loss (train.py:275): the forward stored the image, d|x - gt|/dx = sign(x - gt),
      // scaled by the upstream gradient of this view's loss term
      const float sc = __ldg(v_loss + v);
      float tr, tg, tb;
      load_gt(gt, gt_u8, pix, tr, tg, tb);
      const float dr = __ldg(out_img_clamped + 3 * pix) - tr;
      const float dg = __ldg(out_img_clamped + 3 * pix + 1) - tg;
      const float db = __ldg(out_img_clamped + 3 * pix + 2) - tb;
      vo_r += dr > 0.f ? sc : (dr < 0.f ? -sc : 0.f);
      vo_g += dg > 0.f ? sc : (dg < 0.f ? -sc : 0.f);
      vo_b += db > 0.f ? sc : (db < 0.f ? -sc : 0.f);
    }
    if (out_img_clamped && clamped) {
      // the forward stored min(rgb, 1) (gs_utils.py:111): no gradient through a clamped channel
      if (!(__ldg(out_img_clamped + 3 * pix) < 1.0f)) vo_r = 0.f;
      if (!(__ldg(out_img_clamped + 3 * pix + 1) < 1.0f)) vo_g = 0.f;
      if (!(__ldg(out_img_clamped + 3 * pix + 2) < 1.0f)) vo_b = 0.f;
    }
  }
  float4* myVo = sVo + warp * 32;
  float2* myW = sW + warp * kSlots * kSlotStride;
  myVo[lane] = make_float4(vo_r, vo_g, vo_b, 0.f);
  __shared__ float2 sP[RECT ? 1 : NW * 32];          // per-lane pixel centres (block widths other than 16)
  const float2* myP = sP;
  if (!RECT) { sP[warp * 32 + lane] = make_float2(px, py); myP = sP + warp * 32; }
  const float ox = (float)(tile_x * bw + (warp_t & 1) * 8) + 0.5f;
  const float oy = (float)(tile_y * bw + (warp_t >> 1) * 4) + 0.5f;
  const float bg_dot = __ldg(background) * vo_r + __ldg(background + 1) * vo_g + __ldg(background + 2) * vo_b;
  const int warp_final = __reduce_max_sync(0xffffffffu, bin_final);
  // The walk starts at the last Gaussian any pixel of this CTA composited, not at the end of the tile's
  // list: with early termination (deep tiles, opaque scenes) almost all of the list lies behind it and
  // must not even be staged.
  __shared__ int sFinal;
  if (threadIdx.x == 0) sFinal = -1;
  __syncthreads();
  if (lane == 0) atomicMax(&sFinal, warp_final);
  __syncthreads();
  const int list_end = min(range.y, sFinal + 1);
  const int walk = list_end - range.x;
  if (walk <= 0) return;
  // v_alpha = T * (rgb . v_out) - ra * acc,  acc = sum_{behind} fac_j (rgb_j . v_out) - T_final (v_out_alpha - bg . v_out)
  float T = T_final;
  float acc = -T_final * (vo_a - bg_dot);
  int nslot = 0;
  float own_x = 0.f, own_y = 0.f;
  int own_g = 0;

  for (int b0 = 0; b0 < walk; b0 += kB) {
    __syncthreads();
    const int bsize = min(kB, walk - b0);
    if (threadIdx.x < bsize) {
      const int g = __ldg(ids_sorted + list_end - 1 - b0 - threadIdx.x);   // slot s <-> idx = list_end-1-b0-s
      const float4* p = vp + (size_t)g * 3;
      sA[threadIdx.x] = __ldg(p); sB[threadIdx.x] = __ldg(p + 1); sC[threadIdx.x] = __ldg(p + 2);
      sG[threadIdx.x] = g;
    }
    __syncthreads();
    for (int c0 = 0; c0 < bsize; c0 += 32) {
      // everything at list positions above the warp's last composited Gaussian is untouched
      const int hi_idx = list_end - 1 - b0 - c0;      // list index of slot c0
      if (hi_idx - 31 > warp_final) continue;
      const int j = c0 + lane;
      bool hit = false;
      if (j < bsize) hit = (hi_idx - lane <= warp_final) && touches(sA[j], pm);
      unsigned mask = __ballot_sync(0xffffffffu, hit);
      while (mask) {
        const int k = c0 + __ffs(mask) - 1;
        mask &= mask - 1;
        const int idx = list_end - 1 - b0 - k;
        const float4 A = sA[k], B = sB[k];
        const float dx = A.x - px, dy = A.y - py;
        const float pw = dx * fmaf(B.y, dy, B.x * dx) + B.z * dy * dy;
        const float alpha_raw = ex2_approx(pw + B.w);
        const float alpha = fminf(kAlphaMax, alpha_raw);
        const bool valid = (idx <= bin_final) && (pw <= 0.f) && (alpha >= kAlphaMin);
        if (!__any_sync(0xffffffffu, valid)) continue;
        const float4 C = sC[k];
        const float ra = rcp_approx(1.0f - alpha);
        T = valid ? T * ra : T;
        const float fac = valid ? alpha * T : 0.f;
        const float cdot = C.x * vo_r + C.y * vo_g + C.z * vo_b;
        const float v_alpha = T * cdot - ra * acc;
        acc = fmaf(fac, cdot, acc);
        // d alpha / d sigma = -alpha_raw where the 0.999 clamp is inactive, 0 where it is active
        const float w = (valid && alpha_raw <= kAlphaMax) ? -alpha_raw * v_alpha : 0.f;
        myW[nslot * kSlotStride + lane] = make_float2(w, fac);
        if ((lane & 15) == nslot) { own_x = A.x; own_y = A.y; own_g = sG[k]; }
        if (++nslot == kSlots) {
          flush_slots<RECT>(myW, myVo, myP, kSlots, lane, own_x, own_y, own_g, ox, oy, vgrad);
          nslot = 0;
        }
      }
    }
  }
  if (nslot) flush_slots<RECT>(myW, myVo, myP, nslot, lane, own_x, own_y, own_g, ox, oy, vgrad);
}

}  // namespace sb

extern "C" int sb_rasterize_bwd(int num_views, int n, const float* packed,
                                const int* gaussian_ids_sorted, const int* tile_bins,
                                const float* background, int img_h, int img_w, int block_width,
                                const float* final_ts, const int* final_idx, const float* v_out_img,
                                const float* v_out_alpha, const float* out_img,
                                int clamped_rgb_max1, const void* gt, int gt_is_u8,
                                const float* v_loss,
                                float* v_packed, sb_stream_t stream) {
  if (num_views < 1 || n < 0 || img_h < 1 || img_w < 1 || block_width < 2 || block_width > 16 ||
      (!v_out_img && !gt && !v_out_alpha) || (gt && (!out_img || !v_loss)) || (clamped_rgb_max1 && !out_img))
    return SB_ERR_BAD_ARG;
  const int tiles_x = sb::ceil_div(img_w, block_width), tiles_y = sb::ceil_div(img_h, block_width);
  // opting in to > 48 KB of dynamic shared memory is idempotent per-function state, not library state
  cudaFuncSetAttribute(sb::raster_bwd_kernel<false, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                       (int)sb::bwd_smem(8));
#define SB_LAUNCH(RECT, NW)                                                                               \
  sb::raster_bwd_kernel<RECT, NW><<<dim3(tiles_x * tiles_y * (8 / NW), num_views), NW * 32, sb::bwd_smem(NW), \
                                    (cudaStream_t)stream>>>(                                              \
      n, reinterpret_cast<const float4*>(packed), gaussian_ids_sorted,                                    \
      reinterpret_cast<const int2*>(tile_bins), background, tiles_x, tiles_y, img_w, img_h, block_width,  \
      clamped_rgb_max1, final_ts, final_idx, v_out_img, v_out_alpha, out_img, gt, gt_is_u8, v_loss, v_packed)
  if (block_width == 16) SB_LAUNCH(true, 4); else SB_LAUNCH(false, 8);
#undef SB_LAUNCH
  SB_RETURN_LAST();
}
