// Projection forward (reference call site /root/reference/utils/gs_utils.py:82-95).
//
// THIS TRANSLATION UNIT IS COMPILED WITH --fmad=false: the arithmetic contract in sb_math.cuh /
// oracle/ref_torch.py requires every float operation of the projection chain to be rounded
// separately, so that radii, tile counts, depth keys and pixel centres are bit-identical to the
// oracle.  The kernel is a pure stream (100 B per Gaussian per view), so losing FMA costs nothing.
#include "sb_common.cuh"

namespace sb {

constexpr int kProjBlock = 256;

__global__ void __launch_bounds__(kProjBlock) project_fwd_kernel(
    int n, const float* __restrict__ means, const float* __restrict__ scales, float glob,
    const float* __restrict__ quats, CamSrc cams, int img_h, int img_w, int bw, float clip,
    float* __restrict__ xys, float* __restrict__ depths, int* __restrict__ radii,
    float* __restrict__ conics, float* __restrict__ comp, int* __restrict__ tiles,
    float* __restrict__ cov3d) {
  const int g = blockIdx.x * kProjBlock + threadIdx.x;
  const int v = blockIdx.y;
  if (g >= n) return;
  const Cam cam = load_cam(cams, v);
  const float m[3] = {__ldg(means + 3 * g), __ldg(means + 3 * g + 1), __ldg(means + 3 * g + 2)};
  const float s[3] = {__ldg(scales + 3 * g), __ldg(scales + 3 * g + 1), __ldg(scales + 3 * g + 2)};
  const float4 q4 = __ldg(reinterpret_cast<const float4*>(quats) + g);
  const float q[4] = {q4.x, q4.y, q4.z, q4.w};
  ProjFwd o;
  project_fwd_one(m, s, glob, q, cam, img_w, img_h, bw, clip, o);
  const size_t i = (size_t)v * n + g;
  reinterpret_cast<float2*>(xys)[i] = make_float2(o.px, o.py);
  depths[i] = o.depth;
  radii[i] = o.radius;
  conics[3 * i] = o.ca; conics[3 * i + 1] = o.cb; conics[3 * i + 2] = o.cc;
  if (comp) comp[i] = o.comp;
  tiles[i] = o.tiles;
  if (cov3d) {
#pragma unroll
    for (int k = 0; k < 6; ++k) cov3d[6 * i + k] = o.cov3d[k];
  }
}

}  // namespace sb

extern "C" int sb_project_fwd(int num_views, int n, const float* means, const float* scales,
                              float glob_scale, const float* quats, const float* viewmats,
                              int viewmat_stride, const float* intrins, float fx, float fy, float cx,
                              float cy, int img_h, int img_w, int block_width, float clip_thresh,
                              float* xys, float* depths, int* radii, float* conics,
                              float* compensation, int* num_tiles_hit, float* cov3d,
                              sb_stream_t stream) {
  if (n < 0 || num_views < 1 || block_width < 2 || block_width > 16 || img_h < 1 || img_w < 1)
    return SB_ERR_BAD_ARG;
  if (n == 0) return 0;
  sb::CamSrc cams{viewmats, viewmat_stride, intrins, fx, fy, cx, cy};
  dim3 grid(sb::ceil_div(n, sb::kProjBlock), num_views);
  sb::project_fwd_kernel<<<grid, sb::kProjBlock, 0, (cudaStream_t)stream>>>(
      n, means, scales, glob_scale, quats, cams, img_h, img_w, block_width, clip_thresh, xys, depths,
      radii, conics, compensation, num_tiles_hit, cov3d);
  SB_RETURN_LAST();
}
