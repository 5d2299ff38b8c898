// Rasterise forward: per-16x16-tile front-to-back alpha compositing (upstream
// _C.rasterize_forward behind gsplat.rasterize_gaussians, /root/reference/utils/gs_utils.py:96-109).
//
// One CTA per (tile, view).  The tile's depth-sorted Gaussians are staged through shared memory
// in batches of 256 packed 48-byte records (raster_common.cuh), double-buffered with cp.async so
// that the gather of batch b+1 overlaps the compositing of batch b (one barrier per batch).  Each warp owns an 8x4 pixel
// rectangle: its lanes first test 32 staged Gaussians against the rectangle in parallel (one
// ballot), then the warp walks only the hits in depth order.  The skip is conservative
// (sb::visible_extent), so the composited result is the one the plain per-pixel loop produces.
// The walk is branch-free per lane (predicated selects), the kernel being issue-bound: per
// (warp, hit) it costs 3 LDS + ~25 FP32/ALU operations + one MUFU.EX2.
#include "raster_common.cuh"

namespace sb {

__global__ void __launch_bounds__(kRasterBlock) raster_fwd_kernel(
    int n, const float4* __restrict__ packed, const int* __restrict__ ids_sorted,
    const int2* __restrict__ tile_bins, const float* __restrict__ background, int tiles_x,
    int tiles_y, int img_w, int img_h, int bw, const int* __restrict__ view_has_isects,
    float* __restrict__ out_img, float* __restrict__ final_ts, int* __restrict__ final_idx,
    float* __restrict__ out_alpha, int clamp_rgb, const void* __restrict__ gt, int gt_u8,
    float* __restrict__ loss_partial) {
  __shared__ float4 sAll[2][3][kBatch];          // double-buffered staging: [buffer][record word][slot]
  const int tile = blockIdx.x, v = blockIdx.y;
  const int tile_y = tile / tiles_x, tile_x = tile - tile_y * tiles_x;
  const PixelMap pm = make_pixel_map(tile_x, tile_y, bw, img_w, img_h);
  const float px = (float)pm.px_i + 0.5f, py = (float)pm.py_i + 0.5f;
  const int lane = threadIdx.x & 31;
  const int2 range = __ldg(tile_bins + (size_t)v * tiles_x * tiles_y + tile);
  const int count = range.y - range.x;
  const float4* vp = packed + (size_t)v * n * 3;

  float T = 1.0f, pr = 0.f, pg = 0.f, pb = 0.f;
  int cur_idx = 0;
  bool done = !pm.inside;

  // software pipeline: ids are fetched two batches ahead (register), records one batch ahead (cp.async)
  auto fetch_id = [&](int b0) -> int {
    const int i = b0 + (int)threadIdx.x;
    return i < count ? __ldg(ids_sorted + range.x + i) : -1;
  };
  auto stage = [&](int buf, int g) {
    if (g >= 0) {
      const float4* p = vp + (size_t)g * 3;
      cp_async16(&sAll[buf][0][threadIdx.x], p);
      cp_async16(&sAll[buf][1][threadIdx.x], p + 1);
      cp_async16(&sAll[buf][2][threadIdx.x], p + 2);
    }
    cp_async_commit();
  };
  int g_next = fetch_id(0);
  stage(0, g_next);
  g_next = fetch_id(kBatch);

  for (int b0 = 0, buf = 0; b0 < count; b0 += kBatch, buf ^= 1) {
    cp_async_wait_all();
    if (__syncthreads_count(done) == kRasterBlock) break;
    if (b0 + kBatch < count) {
      stage(buf ^ 1, g_next);
      g_next = fetch_id(b0 + 2 * kBatch);
    }
    const int bsize = min(kBatch, count - b0);
    const float4* sA = sAll[buf][0];
    const float4* sB = sAll[buf][1];
    const float4* sC = sAll[buf][2];
    if (__all_sync(0xffffffffu, done)) continue;
    const int idx0 = range.x + b0;
    for (int c0 = 0; c0 < bsize; c0 += 32) {
      const int j = c0 + lane;
      bool hit = false;
      if (j < bsize) hit = touches(sA[j], pm);
      unsigned mask = __ballot_sync(0xffffffffu, hit);
      while (mask) {
        const int k = c0 + __ffs(mask) - 1;
        mask &= mask - 1;
        const float4 A = sA[k], B = sB[k], C = sC[k];
        const float dx = A.x - px, dy = A.y - py;
        const float pw = dx * fmaf(B.y, dy, B.x * dx) + B.z * dy * dy;      // -sigma * log2(e)
        const float alpha = fminf(kAlphaMax, ex2_approx(pw + B.w));
        const bool valid = !done && (pw <= 0.f) && (alpha >= kAlphaMin);
        const float next_T = T * (1.0f - alpha);
        const bool stop = valid && (next_T <= kTMin);
        const bool upd = valid && !stop;
        done = done || stop;
        const float wgt = upd ? alpha * T : 0.f;
        pr = fmaf(C.x, wgt, pr); pg = fmaf(C.y, wgt, pg); pb = fmaf(C.z, wgt, pb);
        T = upd ? next_T : T;
        cur_idx = upd ? idx0 + k : cur_idx;
      }
      if (__all_sync(0xffffffffu, done)) break;
    }
  }
  cp_async_wait_all();
  float l1 = 0.f, l2 = 0.f;
  if (pm.inside) {
    const size_t pix = ((size_t)v * img_h + pm.py_i) * img_w + pm.px_i;
    bool empty_view = false;
    // upstream quirk: a view without a single intersection reports final_T = 0
    // (alpha = 1) while its image is the plain background
    if (view_has_isects) empty_view = __ldg(view_has_isects + v) == 0;
    float r = pr + T * __ldg(background), g = pg + T * __ldg(background + 1), b = pb + T * __ldg(background + 2);
    if (clamp_rgb) { r = fminf(r, 1.0f); g = fminf(g, 1.0f); b = fminf(b, 1.0f); }   // gs_utils.py:111
    final_ts[pix] = empty_view ? 0.f : T;
    final_idx[pix] = cur_idx;
    if (out_alpha) out_alpha[pix] = empty_view ? 1.f : 1.0f - T;
    out_img[3 * pix] = r; out_img[3 * pix + 1] = g; out_img[3 * pix + 2] = b;
    if (gt) {
      float tr, tg, tb;
      load_gt(gt, gt_u8, pix, tr, tg, tb);
      const float dr = r - tr, dg = g - tg, db = b - tb;
      l1 = fabsf(dr) + fabsf(dg) + fabsf(db);
      l2 = dr * dr + dg * dg + db * db;
    }
  }
  if (loss_partial) {
    // fused L1 / squared-error partial sums of the tile (train.py:275,278): one deterministic
    // (V, T, 2) entry per CTA, summed by the caller
    __shared__ float sRed[2][kRasterBlock / 32];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      l1 += __shfl_xor_sync(0xffffffffu, l1, o);
      l2 += __shfl_xor_sync(0xffffffffu, l2, o);
    }
    if (lane == 0) { sRed[0][threadIdx.x >> 5] = l1; sRed[1][threadIdx.x >> 5] = l2; }
    __syncthreads();
    if (threadIdx.x < 2) {
      float t = 0.f;
#pragma unroll
      for (int w = 0; w < kRasterBlock / 32; ++w) t += sRed[threadIdx.x][w];
      loss_partial[((size_t)v * tiles_x * tiles_y + tile) * 2 + threadIdx.x] = t;
    }
  }
}

}  // namespace sb

extern "C" int sb_rasterize_fwd(int num_views, int n, const float* packed,
                                const int* gaussian_ids_sorted, const int* tile_bins,
                                const float* background, int img_h, int img_w, int block_width,
                                const int* view_has_isects, float* out_img, float* final_ts,
                                int* final_idx, float* out_alpha, int clamp_rgb_max1,
                                const void* gt, int gt_is_u8, float* loss_partial,
                                sb_stream_t stream) {
  if (num_views < 1 || n < 0 || img_h < 1 || img_w < 1 || block_width < 2 || block_width > 16 ||
      (loss_partial && !gt))
    return SB_ERR_BAD_ARG;
  const int tiles_x = sb::ceil_div(img_w, block_width), tiles_y = sb::ceil_div(img_h, block_width);
  dim3 grid(tiles_x * tiles_y, num_views);
  sb::raster_fwd_kernel<<<grid, sb::kRasterBlock, 0, (cudaStream_t)stream>>>(
      n, reinterpret_cast<const float4*>(packed), gaussian_ids_sorted,
      reinterpret_cast<const int2*>(tile_bins), background, tiles_x, tiles_y, img_w, img_h, block_width,
      view_has_isects, out_img, final_ts, final_idx, out_alpha, clamp_rgb_max1, gt, gt_is_u8, loss_partial);
  SB_RETURN_LAST();
}
