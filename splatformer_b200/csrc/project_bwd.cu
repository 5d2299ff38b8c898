// Projection backward (autograd counterpart of gsplat.project_gaussians,
// reference call site /root/reference/utils/gs_utils.py:82-95).  One thread per Gaussian walks
// the views, recomputes the forward record (cheaper than storing ~50 floats per view) and sums
// the per-view VJPs in registers, so v_means / v_scales / v_quats are written once.
#include "sh_stage.cuh"
#include "raster_common.cuh"

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

// ------------------------------------------------------------------------------------------------
// Fused multi-view backward: counterpart of views_project_fwd_kernel.  Reads the packed gradient
// record [v_x v_y v_ca v_cb | v_cc v_op v_r v_g | v_b ...] the raster backward accumulated for
// every (view, Gaussian), recomputes the forward record (this TU is built with --fmad=false like
// the forward, so `valid` is reproduced exactly), and sums over the views in registers (four threads
// per Gaussian take every fourth view each and are joined with two shuffle steps):
// v_means, v_scales, v_quats, v_opacities and v_coeffs (or v_rgb) are each written once.
// ------------------------------------------------------------------------------------------------
constexpr int kViewsBwdBlock = 128;
constexpr int kViewLanesB = 4;                               // threads sharing one Gaussian (every 4th view each)
constexpr int kViewsBwdRows = kViewsBwdBlock / kViewLanesB;  // Gaussians per block

__device__ __forceinline__ float sum_view_lanes(float x) {
  x += __shfl_xor_sync(0xffffffffu, x, 1);
  x += __shfl_xor_sync(0xffffffffu, x, 2);
  return x;
}

template <int DEG>
#ifndef SB_PROJ_BWD_MINB
#define SB_PROJ_BWD_MINB 4
#endif
__global__ void __launch_bounds__(kViewsBwdBlock, SB_PROJ_BWD_MINB) views_project_bwd_kernel(
    int num_views, int n, const float* __restrict__ means, const float* __restrict__ scales, float glob,
    const float* __restrict__ quats, int flags, CamSrc cams, const float* __restrict__ campos, int img_h,
    int img_w, int bw, float clip, int k3_store, const float4* __restrict__ packed,
    const float4* __restrict__ v_packed, float* __restrict__ v_means, float* __restrict__ v_scales,
    float* __restrict__ v_quats, float* __restrict__ v_opacities, float* __restrict__ v_color_src,
    float* __restrict__ v_color_rest) {
  extern __shared__ float smem[];
  constexpr int KU = DEG < 0 ? 1 : (DEG + 1) * (DEG + 1);
  const int g0 = blockIdx.x * kViewsBwdRows;
  const int nb = min(kViewsBwdRows, n - g0);
  const int rowi = threadIdx.x / kViewLanesB, sub = threadIdx.x % kViewLanesB;
  const int g = g0 + rowi;
  const bool active = rowi < nb;
  float acc[KU * 3];
#pragma unroll
  for (int i = 0; i < KU * 3; ++i) acc[i] = 0.f;
  {
    // inactive rows (tail block) run on Gaussian n-1 so that the lane-group shuffles stay convergent
    const int gl = active ? g : n - 1;
    const float m[3] = {__ldg(means + 3 * gl), __ldg(means + 3 * gl + 1), __ldg(means + 3 * gl + 2)};
    float s[3] = {__ldg(scales + 3 * gl), __ldg(scales + 3 * gl + 1), __ldg(scales + 3 * gl + 2)};
    if (flags & kFlagRawScales) { s[0] = expf(s[0]); s[1] = expf(s[1]); s[2] = expf(s[2]); }
    const float4 q4 = __ldg(reinterpret_cast<const float4*>(quats) + gl);
    const float q[4] = {q4.x, q4.y, q4.z, q4.w};
    float am[3] = {0, 0, 0}, as[3] = {0, 0, 0}, aq[4] = {0, 0, 0, 0}, aop = 0.f;
    float op = 0.f, col0 = 0.f, col1 = 0.f, col2 = 0.f;
    for (int v = sub; v < num_views; v += kViewLanesB) {
      const Cam cam = load_cam(cams, v);
      ProjFwd f;
      project_fwd_one(m, s, glob, q, cam, img_w, img_h, bw, clip, f);
      if (!f.valid) continue;
      const size_t i = (size_t)v * n + gl;
      const float4 ga = __ldg(v_packed + 3 * i), gb = __ldg(v_packed + 3 * i + 1), gc = __ldg(v_packed + 3 * i + 2);
      float vm[3], vs[3], vq[4];
      const float4 col = __ldg(packed + 3 * i + 2);       // r, g, b (after the clamp), opacity
      op = col.w; col0 = col.x; col1 = col.y; col2 = col.z;
      float vx, vy, vca, vcb, vcc, vop;
      moments_to_grads(ga, gb, f.ca, f.cb, f.cc, op, vx, vy, vca, vcb, vcc, vop);
      project_bwd_one(m, s, glob, cam, f, vx, vy, 0.f, vca, vcb, vcc, 0.f, nullptr, vm, vs, vq, nullptr);
#pragma unroll
      for (int k = 0; k < 3; ++k) { am[k] += vm[k]; as[k] += vs[k]; }
#pragma unroll
      for (int k = 0; k < 4; ++k) aq[k] += vq[k];
      aop += vop;
      float vr = gb.z, vg = gb.w, vb = gc.x;
      if (DEG >= 0) {
        // clamp(x + 0.5, min 0): gradient passes where > 0
        vr = col.x > 0.f ? vr : 0.f; vg = col.y > 0.f ? vg : 0.f; vb = col.z > 0.f ? vb : 0.f;
        float Y[KU], cp[3];
        load_campos(cams, campos, v, cp);
        sh_basis(DEG, m[0] - cp[0], m[1] - cp[1], m[2] - cp[2], Y);
#pragma unroll
        for (int k = 0; k < KU; ++k) { acc[3 * k] += Y[k] * vr; acc[3 * k + 1] += Y[k] * vg; acc[3 * k + 2] += Y[k] * vb; }
      } else {
        acc[0] += vr; acc[1] += vg; acc[2] += vb;
      }
    }
    // join the four view lanes of the Gaussian (op / colour are view-independent: take any lane that saw them)
#pragma unroll
    for (int k = 0; k < 3; ++k) { am[k] = sum_view_lanes(am[k]); as[k] = sum_view_lanes(as[k]); }
#pragma unroll
    for (int k = 0; k < 4; ++k) aq[k] = sum_view_lanes(aq[k]);
    aop = sum_view_lanes(aop);
#pragma unroll
    for (int i = 0; i < KU * 3; ++i) acc[i] = sum_view_lanes(acc[i]);
    op = fmaxf(fmaxf(op, __shfl_xor_sync(0xffffffffu, op, 1)), fmaxf(__shfl_xor_sync(0xffffffffu, op, 2), __shfl_xor_sync(0xffffffffu, op, 3)));
    if (DEG < 0 && (flags & kFlagColorSigmoid)) {
      // colours are in (0, 1) where a view drew the Gaussian and 0 in lanes that saw none
      col0 = fmaxf(fmaxf(col0, __shfl_xor_sync(0xffffffffu, col0, 1)), fmaxf(__shfl_xor_sync(0xffffffffu, col0, 2), __shfl_xor_sync(0xffffffffu, col0, 3)));
      col1 = fmaxf(fmaxf(col1, __shfl_xor_sync(0xffffffffu, col1, 1)), fmaxf(__shfl_xor_sync(0xffffffffu, col1, 2), __shfl_xor_sync(0xffffffffu, col1, 3)));
      col2 = fmaxf(fmaxf(col2, __shfl_xor_sync(0xffffffffu, col2, 1)), fmaxf(__shfl_xor_sync(0xffffffffu, col2, 2), __shfl_xor_sync(0xffffffffu, col2, 3)));
    }
    // chain rule through the activations the kernel applied itself (gs_utils.py:45,54,65)
    if (flags & kFlagRawScales) { as[0] *= s[0]; as[1] *= s[1]; as[2] *= s[2]; }
    if (flags & kFlagRawOpacities) aop *= op * (1.0f - op);
    if (DEG < 0 && (flags & kFlagColorSigmoid)) {
      acc[0] *= col0 * (1.0f - col0); acc[1] *= col1 * (1.0f - col1); acc[2] *= col2 * (1.0f - col2);
    }
    if (active && sub == 0) {
#pragma unroll
      for (int k = 0; k < 3; ++k) { v_means[3 * g + k] = am[k]; v_scales[3 * g + k] = as[k]; }
      reinterpret_cast<float4*>(v_quats)[g] = make_float4(aq[0], aq[1], aq[2], aq[3]);
      v_opacities[g] = aop;
      if (DEG < 0) { v_color_src[3 * g] = acc[0]; v_color_src[3 * g + 1] = acc[1]; v_color_src[3 * g + 2] = acc[2]; }
    }
  }
  if (DEG >= 0) {
    const int row = padded_row(k3_store);
    if (active && sub == 0) {
#pragma unroll
      for (int i = 0; i < KU * 3; ++i) smem[rowi * row + i] = acc[i];
    }
    __syncthreads();
    if (v_color_rest) {
      stage_rows_out(smem, v_color_src + (size_t)g0 * 3, nb, 3, kViewsBwdBlock, row, 0);
      if (k3_store > 3) stage_rows_out(smem, v_color_rest + (size_t)g0 * (k3_store - 3), nb, k3_store - 3, kViewsBwdBlock, row, 3);
    } else {
      stage_rows_out(smem, v_color_src + (size_t)g0 * k3_store, nb, k3_store, kViewsBwdBlock, row, 0);
    }
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
  sb::CamSrc cams{viewmats, viewmat_stride, intrins, fx, fy, cx, cy, 0};
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

extern "C" int sb_views_project_bwd(int num_views, int n, const float* means, const float* scales,
                                    float glob_scale, const float* quats, int sh_degree, int flags,
                                    const float* viewmats, int viewmat_stride, const float* intrins,
                                    float fx, float fy, float cx, float cy, const float* campos,
                                    int img_h, int img_w, int block_width, float clip_thresh,
                                    const float* packed, const float* v_packed, float* v_means,
                                    float* v_scales, float* v_quats, float* v_opacities,
                                    float* v_color_src, float* v_color_rest, sb_stream_t stream) {
  const bool c2w = flags & sb::kFlagC2wOpenGL;
  if (n < 0 || num_views < 1 || block_width < 2 || block_width > 16 || sh_degree < -1 || sh_degree > 4 ||
      (sh_degree >= 0 && !campos && !c2w) || (sh_degree < 0 && v_color_rest))
    return SB_ERR_BAD_ARG;
  if (n == 0) return 0;
  sb::CamSrc cams{viewmats, viewmat_stride, intrins, fx, fy, cx, cy, c2w ? 1 : 0,
                  (flags & sb::kFlagIntrinsShared) ? 1 : 0};
  const int k3 = sh_degree < 0 ? 3 : (sh_degree + 1) * (sh_degree + 1) * 3;
  const size_t smem = sh_degree < 0 ? 0 : (size_t)sb::kViewsBwdRows * (k3 | 1) * sizeof(float);
  dim3 grid(sb::ceil_div(n, sb::kViewsBwdRows));
  cudaStream_t st = (cudaStream_t)stream;
  const float4* pk = reinterpret_cast<const float4*>(packed);
  const float4* vpk = reinterpret_cast<const float4*>(v_packed);
#define SB_CASE(D)                                                                                   \
  case D:                                                                                            \
    sb::views_project_bwd_kernel<D><<<grid, sb::kViewsBwdBlock, smem, st>>>(                         \
        num_views, n, means, scales, glob_scale, quats, flags, cams, campos, img_h, img_w,           \
        block_width, clip_thresh, k3, pk, vpk, v_means, v_scales, v_quats, v_opacities, v_color_src, \
        v_color_rest);                                                                               \
    break;
  switch (sh_degree) { SB_CASE(-1) SB_CASE(0) SB_CASE(1) SB_CASE(2) SB_CASE(3) SB_CASE(4) }
#undef SB_CASE
  SB_RETURN_LAST();
}
