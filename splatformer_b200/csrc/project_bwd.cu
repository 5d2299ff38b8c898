// Projection backward (autograd counterpart of gsplat.project_gaussians,
// reference call site /root/reference/utils/gs_utils.py:82-95).  One thread per Gaussian walks
// the views, recomputes the forward record (cheaper than storing ~50 floats per view) and sums
// the per-view VJPs in registers, so v_means / v_scales / v_quats are written once.
#include "sb_common.cuh"

namespace sb {

constexpr int kProjBwdBlock = 128;

template <bool POSE_GRAD>
__global__ void __launch_bounds__(kProjBwdBlock) project_bwd_kernel(
    int num_views, int n, const float* __restrict__ means, const float* __restrict__ scales,
    float glob, const float* __restrict__ quats, CamSrc cams, int img_h, int img_w, int bw,
    float clip, const int* __restrict__ radii, const float* __restrict__ v_xys,
    const float* __restrict__ v_depths, const float* __restrict__ v_conics,
    const float* __restrict__ v_comp, const float* __restrict__ v_cov3d,
    float* __restrict__ v_means, float* __restrict__ v_scales, float* __restrict__ v_quats,
    float* __restrict__ v_viewmats) {
  const int g = blockIdx.x * kProjBwdBlock + threadIdx.x;
  const bool active = g < n;
  float m[3] = {0, 0, 0}, s[3] = {1, 1, 1}, q[4] = {1, 0, 0, 0};
  if (active) {
    m[0] = __ldg(means + 3 * g); m[1] = __ldg(means + 3 * g + 1); m[2] = __ldg(means + 3 * g + 2);
    s[0] = __ldg(scales + 3 * g); s[1] = __ldg(scales + 3 * g + 1); s[2] = __ldg(scales + 3 * g + 2);
    const float4 q4 = __ldg(reinterpret_cast<const float4*>(quats) + g);
    q[0] = q4.x; q[1] = q4.y; q[2] = q4.z; q[3] = q4.w;
  }
  float am[3] = {0, 0, 0}, as[3] = {0, 0, 0}, aq[4] = {0, 0, 0, 0};
  __shared__ float s_vm[12];
  for (int v = 0; v < num_views; ++v) {
    const Cam cam = load_cam(cams, v);
    float vvm[12];
    if (POSE_GRAD) {
#pragma unroll
      for (int k = 0; k < 12; ++k) vvm[k] = 0.f;
      if (threadIdx.x < 12) s_vm[threadIdx.x] = 0.f;
      __syncthreads();
    }
    if (active) {
      const size_t i = (size_t)v * n + g;
      const bool drawn = __ldg(radii + i) > 0;
      const float* vc3 = v_cov3d ? v_cov3d + 6 * i : nullptr;
      if (drawn || vc3) {
        ProjFwd f;
        project_fwd_one(m, s, glob, q, cam, img_w, img_h, bw, clip, f);
        float vm[3], vs[3], vq[4];
        float c3[6] = {0, 0, 0, 0, 0, 0};
        if (vc3) {
#pragma unroll
          for (int k = 0; k < 6; ++k) c3[k] = __ldg(vc3 + k);
        }
        if (drawn && f.valid) {
          const float2 vxy = __ldg(reinterpret_cast<const float2*>(v_xys) + i);
          project_bwd_one(m, s, glob, cam, f, vxy.x, vxy.y, v_depths ? __ldg(v_depths + i) : 0.f,
                          __ldg(v_conics + 3 * i), __ldg(v_conics + 3 * i + 1), __ldg(v_conics + 3 * i + 2),
                          v_comp ? __ldg(v_comp + i) : 0.f, vc3 ? c3 : nullptr, vm, vs, vq,
                          POSE_GRAD ? vvm : nullptr);
#pragma unroll
          for (int k = 0; k < 3; ++k) { am[k] += vm[k]; as[k] += vs[k]; }
#pragma unroll
          for (int k = 0; k < 4; ++k) aq[k] += vq[k];
        } else if (f.front && vc3) {
          cov3d_bwd_one(s, glob, f, c3, vs, vq);
#pragma unroll
          for (int k = 0; k < 3; ++k) as[k] += vs[k];
#pragma unroll
          for (int k = 0; k < 4; ++k) aq[k] += vq[k];
        }
      }
    }
    if (POSE_GRAD) {
      // block reduction of the 12 pose-gradient terms, one atomic per block and entry
#pragma unroll
      for (int k = 0; k < 12; ++k) {
        float x = vvm[k];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
        if ((threadIdx.x & 31) == 0 && x != 0.f) atomicAdd(&s_vm[k], x);
      }
      __syncthreads();
      if (threadIdx.x < 12 && s_vm[threadIdx.x] != 0.f)
        atomicAdd(v_viewmats + 12 * v + threadIdx.x, s_vm[threadIdx.x]);
      __syncthreads();
    }
  }
  if (active) {
#pragma unroll
    for (int k = 0; k < 3; ++k) { v_means[3 * g + k] = am[k]; v_scales[3 * g + k] = as[k]; }
    reinterpret_cast<float4*>(v_quats)[g] = make_float4(aq[0], aq[1], aq[2], aq[3]);
  }
}

}  // namespace sb

extern "C" int sb_project_bwd(int num_views, int n, const float* means, const float* scales,
                              float glob_scale, const float* quats, const float* viewmats,
                              int viewmat_stride, const float* intrins, float fx, float fy, float cx,
                              float cy, int img_h, int img_w, int block_width, float clip_thresh,
                              const int* radii, const float* v_xys, const float* v_depths,
                              const float* v_conics, const float* v_compensation,
                              const float* v_cov3d, float* v_means, float* v_scales, float* v_quats,
                              float* v_viewmats, sb_stream_t stream) {
  if (n < 0 || num_views < 1 || block_width < 2 || block_width > 16) return SB_ERR_BAD_ARG;
  if (n == 0) return 0;
  sb::CamSrc cams{viewmats, viewmat_stride, intrins, fx, fy, cx, cy};
  dim3 grid(sb::ceil_div(n, sb::kProjBwdBlock));
  cudaStream_t st = (cudaStream_t)stream;
  if (v_viewmats)
    sb::project_bwd_kernel<true><<<grid, sb::kProjBwdBlock, 0, st>>>(
        num_views, n, means, scales, glob_scale, quats, cams, img_h, img_w, block_width, clip_thresh,
        radii, v_xys, v_depths, v_conics, v_compensation, v_cov3d, v_means, v_scales, v_quats, v_viewmats);
  else
    sb::project_bwd_kernel<false><<<grid, sb::kProjBwdBlock, 0, st>>>(
        num_views, n, means, scales, glob_scale, quats, cams, img_h, img_w, block_width, clip_thresh,
        radii, v_xys, v_depths, v_conics, v_compensation, v_cov3d, v_means, v_scales, v_quats, nullptr);
  SB_RETURN_LAST();
}
