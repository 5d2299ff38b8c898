// Rasterise forward: per-16x16-tile front-to-back alpha compositing (upstream
// _C.rasterize_forward behind gsplat.rasterize_gaussians, /root/reference/utils/gs_utils.py:96-109).
//
// One CTA per (tile, view).  The tile's depth-sorted Gaussians are staged through shared memory
// in batches of 256 packed 48-byte records.  Each warp owns an 8x4 pixel rectangle: its lanes
// first test 32 staged Gaussians against the rectangle in parallel (one ballot), then the warp
// walks only the hits in depth order.  The skip is conservative (sb::visible_extent), so the
// composited result is the one the plain per-pixel loop produces.
#include "raster_common.cuh"

namespace sb {

__global__ void __launch_bounds__(kRasterBlock) raster_fwd_kernel(
    int n, const float4* __restrict__ packed, const int* __restrict__ ids_sorted,
    const int2* __restrict__ tile_bins, const float* __restrict__ background, int tiles_x,
    int tiles_y, int img_w, int img_h, int bw, const int* __restrict__ cum_tiles_hit,
    float* __restrict__ out_img, float* __restrict__ final_ts, int* __restrict__ final_idx) {
  __shared__ float4 sA[kBatch], sB[kBatch], sC[kBatch];
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

  for (int b0 = 0; b0 < count; b0 += kBatch) {
    if (__syncthreads_count(done) == kRasterBlock) break;
    const int bsize = min(kBatch, count - b0);
    if (threadIdx.x < bsize) {
      const int g = __ldg(ids_sorted + range.x + b0 + threadIdx.x);
      const float4* p = vp + (size_t)g * 3;
      sA[threadIdx.x] = __ldg(p); sB[threadIdx.x] = __ldg(p + 1); sC[threadIdx.x] = __ldg(p + 2);
    }
    __syncthreads();
    if (__all_sync(0xffffffffu, done)) continue;
    for (int c0 = 0; c0 < bsize; c0 += 32) {
      const int j = c0 + lane;
      bool hit = false;
      if (j < bsize) hit = touches(sA[j], pm);
      unsigned mask = __ballot_sync(0xffffffffu, hit);
      while (mask) {
        const int k = c0 + __ffs(mask) - 1;
        mask &= mask - 1;
        if (done) continue;
        const float4 A = sA[k], B = sB[k];
        const float dx = A.x - px, dy = A.y - py;
        const float sigma = 0.5f * (B.x * dx * dx + B.z * dy * dy) + B.y * dx * dy;
        if (sigma < 0.f) continue;
        const float alpha = fminf(kAlphaMax, B.w * __expf(-sigma));
        if (alpha < kAlphaMin) continue;
        const float next_T = T * (1.0f - alpha);
        if (next_T <= kTMin) { done = true; continue; }
        const float4 C = sC[k];
        const float wgt = alpha * T;
        pr += C.x * wgt; pg += C.y * wgt; pb += C.z * wgt;
        T = next_T;
        cur_idx = range.x + b0 + k;
      }
      if (__all_sync(0xffffffffu, done)) break;
    }
  }
  if (pm.inside) {
    const size_t pix = ((size_t)v * img_h + pm.py_i) * img_w + pm.px_i;
    if (cum_tiles_hit) {
      // upstream quirk: a view without a single intersection reports final_T = 0
      // (alpha = 1) while its image is the plain background
      const int hi = __ldg(cum_tiles_hit + (size_t)(v + 1) * n - 1);
      const int lo = v ? __ldg(cum_tiles_hit + (size_t)v * n - 1) : 0;
      if (hi - lo < 1) {
        final_ts[pix] = 0.f; final_idx[pix] = 0;
        out_img[3 * pix] = __ldg(background); out_img[3 * pix + 1] = __ldg(background + 1);
        out_img[3 * pix + 2] = __ldg(background + 2);
        return;
      }
    }
    final_ts[pix] = T;
    final_idx[pix] = cur_idx;
    out_img[3 * pix] = pr + T * __ldg(background);
    out_img[3 * pix + 1] = pg + T * __ldg(background + 1);
    out_img[3 * pix + 2] = pb + T * __ldg(background + 2);
  }
}

}  // namespace sb

extern "C" int sb_rasterize_fwd(int num_views, int n, const float* packed,
                                const int* gaussian_ids_sorted, const int* tile_bins,
                                const float* background, int img_h, int img_w, int block_width,
                                const int* cum_tiles_hit, float* out_img, float* final_ts,
                                int* final_idx, sb_stream_t stream) {
  if (num_views < 1 || n < 0 || img_h < 1 || img_w < 1 || block_width < 2 || block_width > 16)
    return SB_ERR_BAD_ARG;
  const int tiles_x = sb::ceil_div(img_w, block_width), tiles_y = sb::ceil_div(img_h, block_width);
  dim3 grid(tiles_x * tiles_y, num_views);
  sb::raster_fwd_kernel<<<grid, sb::kRasterBlock, 0, (cudaStream_t)stream>>>(
      n, reinterpret_cast<const float4*>(packed), gaussian_ids_sorted,
      reinterpret_cast<const int2*>(tile_bins), background, tiles_x, tiles_y, img_w, img_h, block_width,
      cum_tiles_hit, out_img, final_ts, final_idx);
  SB_RETURN_LAST();
}
