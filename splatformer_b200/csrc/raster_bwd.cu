// Rasterise backward: reverse compositing (upstream _C.rasterize_backward behind
// gsplat.rasterize_gaussians' autograd, /root/reference/utils/gs_utils.py:96-109, train.py:289).
//
// Same CTA/warp decomposition as the forward kernel, walking each tile's list back to front.
// Per pair the kernel produces w = dL/dsigma and three colour gradients; what leaves the warp are
// nine sums per (warp, Gaussian): the moments {w, w dx, w dy, w dx^2, w dx dy, w dy^2} and the rgb
// gradient (raster_common.cuh explains how the consumers turn moments into v_xy / v_conic /
// v_opacity).  Eight of them are summed with a halving butterfly (4+2+1 exchange steps + 2 plain
// steps = 9 shuffles for 8 values), the ninth with a plain 5-step butterfly, and nine lanes issue
// one red.global.add.f32 each.
#include "raster_common.cuh"

namespace sb {

#define SB_XCHG(a, b, up, delta)                                        \
  {                                                                     \
    const float send_ = (up) ? (a) : (b);                               \
    const float keep_ = (up) ? (b) : (a);                               \
    (a) = keep_ + __shfl_xor_sync(0xffffffffu, send_, (delta));         \
  }

// On return lane L holds in v[0] the warp-wide sum of value (L >> 2) of v[0..7], and the return
// value is the warp-wide sum of v8 (in every lane).
__device__ __forceinline__ float warp_sum_8p1(float (&v)[8], float v8, int lane) {
  const bool u16 = lane & 16, u8 = lane & 8, u4 = lane & 4;
  SB_XCHG(v[0], v[4], u16, 16) SB_XCHG(v[1], v[5], u16, 16) SB_XCHG(v[2], v[6], u16, 16) SB_XCHG(v[3], v[7], u16, 16)
  SB_XCHG(v[0], v[2], u8, 8) SB_XCHG(v[1], v[3], u8, 8)
  SB_XCHG(v[0], v[1], u4, 4)
  v[0] += __shfl_xor_sync(0xffffffffu, v[0], 2);
  v[0] += __shfl_xor_sync(0xffffffffu, v[0], 1);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v8 += __shfl_xor_sync(0xffffffffu, v8, o);
  return v8;
}
#undef SB_XCHG

__global__ void __launch_bounds__(kRasterBlock) raster_bwd_kernel(
    int n, const float4* __restrict__ packed, const int* __restrict__ ids_sorted,
    const int2* __restrict__ tile_bins, const float* __restrict__ background, int tiles_x,
    int tiles_y, int img_w, int img_h, int bw, const float* __restrict__ final_ts,
    const int* __restrict__ final_idx, const float* __restrict__ v_out_img,
    const float* __restrict__ v_out_alpha, const float* __restrict__ out_img_clamped,
    float* __restrict__ v_packed) {
  __shared__ float4 sA[kBatch], sB[kBatch], sC[kBatch];
  __shared__ int sG[kBatch];
  const int tile = blockIdx.x, v = blockIdx.y;
  const int tile_y = tile / tiles_x, tile_x = tile - tile_y * tiles_x;
  const PixelMap pm = make_pixel_map(tile_x, tile_y, bw, img_w, img_h);
  const float px = (float)pm.px_i + 0.5f, py = (float)pm.py_i + 0.5f;
  const int lane = threadIdx.x & 31;
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
    vo_r = __ldg(v_out_img + 3 * pix); vo_g = __ldg(v_out_img + 3 * pix + 1); vo_b = __ldg(v_out_img + 3 * pix + 2);
    if (v_out_alpha) vo_a = __ldg(v_out_alpha + pix);
    if (out_img_clamped) {
      // the forward stored min(rgb, 1) (gs_utils.py:111): no gradient through a clamped channel
      if (!(__ldg(out_img_clamped + 3 * pix) < 1.0f)) vo_r = 0.f;
      if (!(__ldg(out_img_clamped + 3 * pix + 1) < 1.0f)) vo_g = 0.f;
      if (!(__ldg(out_img_clamped + 3 * pix + 2) < 1.0f)) vo_b = 0.f;
    }
  }
  const float bg_dot = __ldg(background) * vo_r + __ldg(background + 1) * vo_g + __ldg(background + 2) * vo_b;
  const int warp_final = __reduce_max_sync(0xffffffffu, bin_final);
  // v_alpha = T * (rgb . v_out) - ra * acc,  acc = sum_{behind} fac_j (rgb_j . v_out) - T_final (v_out_alpha - bg . v_out)
  float T = T_final;
  float acc = -T_final * (vo_a - bg_dot);
  // lane -> which reduced value it owns, and whether it issues the global reduction
  const int my_slot = (lane == 1) ? 8 : (lane >> 2);
  const bool writer = ((lane & 3) == 0) || (lane == 1);

  for (int b0 = 0; b0 < count; b0 += kBatch) {
    // everything at list positions above the warp's last composited Gaussian is untouched; once the
    // whole CTA is past its positions nothing is left to do
    __syncthreads();
    const int bsize = min(kBatch, count - b0);
    if (threadIdx.x < bsize) {
      const int g = __ldg(ids_sorted + range.y - 1 - b0 - threadIdx.x);   // slot s <-> idx = end-1-b0-s
      const float4* p = vp + (size_t)g * 3;
      sA[threadIdx.x] = __ldg(p); sB[threadIdx.x] = __ldg(p + 1); sC[threadIdx.x] = __ldg(p + 2);
      sG[threadIdx.x] = g;
    }
    __syncthreads();
    for (int c0 = 0; c0 < bsize; c0 += 32) {
      const int hi_idx = range.y - 1 - b0 - c0;       // list index of slot c0
      if (hi_idx - 31 > warp_final) continue;
      const int j = c0 + lane;
      bool hit = false;
      if (j < bsize) hit = (hi_idx - lane <= warp_final) && touches(sA[j], pm);
      unsigned mask = __ballot_sync(0xffffffffu, hit);
      while (mask) {
        const int k = c0 + __ffs(mask) - 1;
        mask &= mask - 1;
        const int idx = range.y - 1 - b0 - k;
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
        float vals[8];
        const float wdx = w * dx, wdy = w * dy;
        vals[0] = wdx; vals[1] = wdy; vals[2] = wdx * dx; vals[3] = wdx * dy;
        vals[4] = wdy * dy; vals[5] = w; vals[6] = fac * vo_r; vals[7] = fac * vo_g;
        const float s8 = warp_sum_8p1(vals, fac * vo_b, lane);
        const float total = (lane == 1) ? s8 : vals[0];
        if (writer && total != 0.f) atomicAdd(vgrad + (size_t)sG[k] * 12 + my_slot, total);
      }
    }
  }
}

}  // namespace sb

extern "C" int sb_rasterize_bwd(int num_views, int n, const float* packed,
                                const int* gaussian_ids_sorted, const int* tile_bins,
                                const float* background, int img_h, int img_w, int block_width,
                                const float* final_ts, const int* final_idx, const float* v_out_img,
                                const float* v_out_alpha, const float* out_img_clamped,
                                float* v_packed, sb_stream_t stream) {
  if (num_views < 1 || n < 0 || img_h < 1 || img_w < 1 || block_width < 2 || block_width > 16)
    return SB_ERR_BAD_ARG;
  const int tiles_x = sb::ceil_div(img_w, block_width), tiles_y = sb::ceil_div(img_h, block_width);
  dim3 grid(tiles_x * tiles_y, num_views);
  sb::raster_bwd_kernel<<<grid, sb::kRasterBlock, 0, (cudaStream_t)stream>>>(
      n, reinterpret_cast<const float4*>(packed), gaussian_ids_sorted,
      reinterpret_cast<const int2*>(tile_bins), background, tiles_x, tiles_y, img_w, img_h, block_width,
      final_ts, final_idx, v_out_img, v_out_alpha, out_img_clamped, v_packed);
  SB_RETURN_LAST();
}
