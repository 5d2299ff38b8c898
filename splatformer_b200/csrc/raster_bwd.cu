// Rasterise backward: reverse compositing (upstream _C.rasterize_backward behind
// gsplat.rasterize_gaussians' autograd, /root/reference/utils/gs_utils.py:96-109, train.py:289).
//
// Same CTA/warp decomposition as the forward kernel, walking each tile's list back to front.
// For every (warp, Gaussian) hit the nine per-Gaussian gradient terms are summed across the
// warp with a packed butterfly (9 values cost 16 shuffles instead of 45) and leave the SM as
// three vector reductions (red.global.add.v4.f32 / .f32) into the packed gradient record.
#include "raster_common.cuh"

namespace sb {

// Butterfly that sums 16 per-lane values across the warp; on return lane L holds in `out` the
// warp-wide sum of value (L & 15)  (lanes L and L^16 hold the same value).
__device__ __forceinline__ float warp_sum16(float (&v)[16], int lane) {
  // step 1: 16 -> 8 values, partner lane ^ 16 ... wait: we keep all lanes busy by halving the
  // value set at every step while doubling the number of lanes already summed.
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const bool up = lane & 8;
    const float send = up ? v[i] : v[i + 8];
    const float keep = up ? v[i + 8] : v[i];
    v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const bool up = lane & 4;
    const float send = up ? v[i] : v[i + 4];
    const float keep = up ? v[i + 4] : v[i];
    v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
  }
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const bool up = lane & 2;
    const float send = up ? v[i] : v[i + 2];
    const float keep = up ? v[i + 2] : v[i];
    v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 2);
  }
  {
    const bool up = lane & 1;
    const float send = up ? v[0] : v[1];
    const float keep = up ? v[1] : v[0];
    v[0] = keep + __shfl_xor_sync(0xffffffffu, send, 1);
  }
  // now lane L holds the sum over the 16 lanes sharing L's bit 4 of value index
  // ((L&8)?8:0) + ((L&4)?4:0) + ((L&2)?2:0) + ((L&1)?1:0) = L & 15
  return v[0] + __shfl_xor_sync(0xffffffffu, v[0], 16);
}

__global__ void __launch_bounds__(kRasterBlock) raster_bwd_kernel(
    int n, const float4* __restrict__ packed, const int* __restrict__ ids_sorted,
    const int2* __restrict__ tile_bins, const float* __restrict__ background, int tiles_x,
    int tiles_y, int img_w, int img_h, int bw, const float* __restrict__ final_ts,
    const int* __restrict__ final_idx, const float* __restrict__ v_out_img,
    const float* __restrict__ v_out_alpha, float* __restrict__ v_packed) {
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
  }
  const float bg_dot = __ldg(background) * vo_r + __ldg(background + 1) * vo_g + __ldg(background + 2) * vo_b;
  const int warp_final = __reduce_max_sync(0xffffffffu, bin_final);
  float T = T_final, buf_r = 0.f, buf_g = 0.f, buf_b = 0.f;

  for (int b0 = 0; b0 < count; b0 += kBatch) {
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
      // lowest list index covered by this chunk; everything above warp_final is untouched
      const int hi_idx = range.y - 1 - b0 - c0;
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
        const float sigma = 0.5f * (B.x * dx * dx + B.z * dy * dy) + B.y * dx * dy;
        const float vis = __expf(-sigma);
        const float alpha_raw = B.w * vis;
        const float alpha = fminf(kAlphaMax, alpha_raw);
        const bool valid = (idx <= bin_final) && (sigma >= 0.f) && (alpha >= kAlphaMin);
        if (!__any_sync(0xffffffffu, valid)) continue;
        float vals[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) vals[i] = 0.f;
        if (valid) {
          const float4 C = sC[k];
          const float ra = 1.0f / (1.0f - alpha);
          T *= ra;
          const float fac = alpha * T;
          float v_alpha = (C.x * T - buf_r * ra) * vo_r + (C.y * T - buf_g * ra) * vo_g + (C.z * T - buf_b * ra) * vo_b;
          v_alpha += T_final * ra * (vo_a - bg_dot);
          buf_r += C.x * fac; buf_g += C.y * fac; buf_b += C.z * fac;
          vals[6] = fac * vo_r; vals[7] = fac * vo_g; vals[8] = fac * vo_b;
          if (alpha_raw <= kAlphaMax) {
            const float v_sigma = -alpha_raw * v_alpha;
            vals[0] = v_sigma * (B.x * dx + B.y * dy);
            vals[1] = v_sigma * (B.y * dx + B.z * dy);
            vals[2] = 0.5f * v_sigma * dx * dx;
            vals[3] = v_sigma * dx * dy;
            vals[4] = 0.5f * v_sigma * dy * dy;
            vals[5] = vis * v_alpha;
          }
        }
        const float total = warp_sum16(vals, lane);
        // lane L (< 16) holds the sum of value L: [vx vy vca vcb | vcc vop vr vg | vb ...]
        float* dst = vgrad + (size_t)sG[k] * 12;
        if (lane < 9 && total != 0.f) atomicAdd(dst + lane, total);
      }
    }
  }
}

}  // namespace sb

extern "C" int sb_rasterize_bwd(int num_views, int n, const float* packed,
                                const int* gaussian_ids_sorted, const int* tile_bins,
                                const float* background, int img_h, int img_w, int block_width,
                                const float* final_ts, const int* final_idx, const float* v_out_img,
                                const float* v_out_alpha, float* v_packed, sb_stream_t stream) {
  if (num_views < 1 || n < 0 || img_h < 1 || img_w < 1 || block_width < 2 || block_width > 16)
    return SB_ERR_BAD_ARG;
  const int tiles_x = sb::ceil_div(img_w, block_width), tiles_y = sb::ceil_div(img_h, block_width);
  dim3 grid(tiles_x * tiles_y, num_views);
  sb::raster_bwd_kernel<<<grid, sb::kRasterBlock, 0, (cudaStream_t)stream>>>(
      n, reinterpret_cast<const float4*>(packed), gaussian_ids_sorted,
      reinterpret_cast<const int2*>(tile_bins), background, tiles_x, tiles_y, img_w, img_h, block_width,
      final_ts, final_idx, v_out_img, v_out_alpha, v_packed);
  SB_RETURN_LAST();
}
