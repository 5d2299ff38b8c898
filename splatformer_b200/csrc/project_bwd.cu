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
// record [M1x M1y M2xx M2xy | M2yy M0 v_r v_g | v_b ...] the raster backward accumulated for every
// (view, Gaussian), recomputes the forward record (this TU is built with --fmad=false like the
// forward, so `valid` is reproduced exactly) and sums the VJPs over the views.
//
// Most (view, Gaussian) records are all-zero: a Gaussian behind the opaque front of the object never
// reaches a pixel before that pixel's transmittance runs out, and a culled one is never touched.  A
// thread that owned fixed (Gaussian, view) pairs would still run the ~1000-instruction recompute +
// VJP whenever ANY lane of its warp had work.  So every warp first scans the records of its 8
// Gaussians x V views (coalesced: 8 consecutive 48-byte records per view), compacts the non-zero
// pairs into a list in shared memory with one ballot per 32 pairs, and then the lanes take pairs from
// that list -- all 32 lanes busy until the list runs out.  The per-Gaussian sums live in shared memory
// (red.shared.add.f32; the SH block is the very buffer that is staged out to v_color_src / v_color_rest
// afterwards), so no thread carries 48 accumulators in registers across the loop.
// ------------------------------------------------------------------------------------------------
constexpr int kViewsBwdBlock = 128;
constexpr int kBwdWarps = kViewsBwdBlock / 32;
constexpr int kBwdRowsPerWarp = 8;                            // Gaussians whose pairs one warp compacts together
constexpr int kViewsBwdRows = kBwdWarps * kBwdRowsPerWarp;    // Gaussians per block
constexpr int kBwdViewChunk = 64;                             // views scanned per compaction round (list capacity)
constexpr int kBwdAcc = 20;    // per-Gaussian sums: v_mean[3] v_scale[3] v_quat[4] v_op | op | v_rgb[3] pad | rgb[3] pad
constexpr int kBwdPar = 12;    // per-Gaussian inputs: mean[3] scale[3] quat[4] |q_raw| pad

__host__ __device__ inline size_t views_bwd_smem_bytes(int k3_store, bool sh) {
  return ((sh ? (size_t)kViewsBwdRows * (k3_store | 1) : 0) + (size_t)kViewsBwdRows * (kBwdAcc + kBwdPar)) * sizeof(float) +
         (size_t)kBwdWarps * kBwdRowsPerWarp * kBwdViewChunk * sizeof(unsigned short);
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
    float* __restrict__ v_color_rest, int* __restrict__ stats) {
  extern __shared__ float smem[];
  constexpr int KU = DEG < 0 ? 1 : (DEG + 1) * (DEG + 1);
  const int row_stride = DEG >= 0 ? padded_row(k3_store) : 0;
  float* s_sh = smem;                                              // [rows][row_stride]  SH gradient sums
  float* s_acc = s_sh + kViewsBwdRows * row_stride;                // [rows][kBwdAcc]
  float* s_par = s_acc + kViewsBwdRows * kBwdAcc;                  // [rows][kBwdPar]
  unsigned short* s_list = reinterpret_cast<unsigned short*>(s_par + kViewsBwdRows * kBwdPar);
  const int g0 = blockIdx.x * kViewsBwdRows;
  const int nb = min(kViewsBwdRows, n - g0);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;

  for (int i = threadIdx.x; i < kViewsBwdRows * (row_stride + kBwdAcc); i += kViewsBwdBlock) smem[i] = 0.f;
  if (threadIdx.x < nb) {
    // view-independent inputs of the block's Gaussians, activations applied once (gs_utils.py:45-51)
    const int g = g0 + threadIdx.x;
    float* par = s_par + threadIdx.x * kBwdPar;
    par[0] = __ldg(means + 3 * g); par[1] = __ldg(means + 3 * g + 1); par[2] = __ldg(means + 3 * g + 2);
    float s0 = __ldg(scales + 3 * g), s1 = __ldg(scales + 3 * g + 1), s2 = __ldg(scales + 3 * g + 2);
    if (flags & kFlagRawScales) { s0 = expf(s0); s1 = expf(s1); s2 = expf(s2); }
    par[3] = s0; par[4] = s1; par[5] = s2;
    const float4 q4 = __ldg(reinterpret_cast<const float4*>(quats) + g);
    float q[4] = {q4.x, q4.y, q4.z, q4.w};
    float qraw = 1.0f;
    if (flags & kFlagRawQuats) qraw = wrapper_normalise_quat(q);
    par[6] = q[0]; par[7] = q[1]; par[8] = q[2]; par[9] = q[3]; par[10] = qraw;
  }
  __syncthreads();

  const int rows_w = max(0, min(kBwdRowsPerWarp, nb - warp * kBwdRowsPerWarp));     // this warp's Gaussians
  const int row0 = warp * kBwdRowsPerWarp;
  unsigned short* list = s_list + warp * (kBwdRowsPerWarp * kBwdViewChunk);
  int n_pairs = 0, n_live = 0;
  for (int vc0 = 0; vc0 < num_views && rows_w > 0; vc0 += kBwdViewChunk) {
    const int vc = min(kBwdViewChunk, num_views - vc0);
    // ---- scan: pair p = (view vc0 + p / 8, row p % 8); keep the ones with a non-zero gradient record ----
    int cnt = 0;
    const int total = vc * kBwdRowsPerWarp;
    for (int p0 = 0; p0 < total; p0 += 32) {
      const int p = p0 + lane;
      const int r = p & (kBwdRowsPerWarp - 1), vv = p >> 3;
      bool nz = false;
      if (p < total && r < rows_w) {
        const float4* rec = v_packed + 3 * ((size_t)(vc0 + vv) * n + g0 + row0 + r);
        const float4 a = __ldg(rec), b = __ldg(rec + 1);
        const float c = __ldg(reinterpret_cast<const float*>(rec + 2));
        nz = a.x != 0.f || a.y != 0.f || a.z != 0.f || a.w != 0.f || b.x != 0.f || b.y != 0.f || b.z != 0.f ||
             b.w != 0.f || c != 0.f;
      }
      const unsigned bal = __ballot_sync(0xffffffffu, nz);
      if (nz) list[cnt + __popc(bal & ((1u << lane) - 1u))] = (unsigned short)p;
      cnt += __popc(bal);
    }
    __syncwarp();
    n_pairs += total; n_live += cnt;
    // ---- work: lanes take the live pairs ----
    for (int k = lane; k < cnt; k += 32) {
      const int p = list[k];
      const int r = row0 + (p & (kBwdRowsPerWarp - 1)), v = vc0 + (p >> 3);
      const float* par = s_par + r * kBwdPar;
      const float m[3] = {par[0], par[1], par[2]};
      const float s[3] = {par[3], par[4], par[5]};
      const float q[4] = {par[6], par[7], par[8], par[9]};
      const Cam cam = load_cam(cams, v);
      ProjFwd f;
      project_fwd_one(m, s, glob, q, cam, img_w, img_h, bw, clip, f);
      if (!f.valid) continue;                          // (a culled pair's record is never written: not reached)
      const size_t i = (size_t)v * n + g0 + r;
      const float4 ga = __ldg(v_packed + 3 * i), gb = __ldg(v_packed + 3 * i + 1);
      const float gcx = __ldg(reinterpret_cast<const float*>(v_packed + 3 * i + 2));
      const float4 col = __ldg(packed + 3 * i + 2);       // r, g, b (after the clamp), opacity
      float vm[3], vs[3], vq[4];
      float vx, vy, vca, vcb, vcc, vop;
      moments_to_grads(ga, gb, f.ca, f.cb, f.cc, col.w, vx, vy, vca, vcb, vcc, vop);
      project_bwd_one(m, s, glob, cam, f, vx, vy, 0.f, vca, vcb, vcc, 0.f, nullptr, vm, vs, vq, nullptr);
      float* acc = s_acc + r * kBwdAcc;
#pragma unroll
      for (int j = 0; j < 3; ++j) { atomicAdd(acc + j, vm[j]); atomicAdd(acc + 3 + j, vs[j]); }
#pragma unroll
      for (int j = 0; j < 4; ++j) atomicAdd(acc + 6 + j, vq[j]);
      atomicAdd(acc + 10, vop);
      acc[11] = col.w;                                   // view-independent: every writer stores the same value
      float vr = gb.z, vg = gb.w, vb = gcx;
      if (DEG >= 0) {
        // clamp(x + 0.5, min 0): gradient passes where > 0
        vr = col.x > 0.f ? vr : 0.f; vg = col.y > 0.f ? vg : 0.f; vb = col.z > 0.f ? vb : 0.f;
        float Y[KU], cp[3];
        load_campos(cams, campos, v, cp);
        sh_basis(DEG, m[0] - cp[0], m[1] - cp[1], m[2] - cp[2], Y);
        float* sh = s_sh + r * row_stride;
#pragma unroll
        for (int j = 0; j < KU; ++j) {
          atomicAdd(sh + 3 * j, Y[j] * vr); atomicAdd(sh + 3 * j + 1, Y[j] * vg); atomicAdd(sh + 3 * j + 2, Y[j] * vb);
        }
      } else {
        atomicAdd(acc + 12, vr); atomicAdd(acc + 13, vg); atomicAdd(acc + 14, vb);
        acc[16] = col.x; acc[17] = col.y; acc[18] = col.z;     // activated colours (view-independent)
      }
    }
    __syncwarp();
  }
  if (stats && lane == 0 && n_pairs > 0) { atomicAdd(stats, n_pairs); atomicAdd(stats + 1, n_live); }
  __syncthreads();

  // ---- epilogue: chain rule through the activations the kernels applied themselves (gs_utils.py:45-51,54,65) ----
  if (threadIdx.x < nb) {
    const int g = g0 + threadIdx.x;
    const float* acc = s_acc + threadIdx.x * kBwdAcc;
    const float* par = s_par + threadIdx.x * kBwdPar;
    float as0 = acc[3], as1 = acc[4], as2 = acc[5];
    if (flags & kFlagRawScales) { as0 *= par[3]; as1 *= par[4]; as2 *= par[5]; }
    v_means[3 * g] = acc[0]; v_means[3 * g + 1] = acc[1]; v_means[3 * g + 2] = acc[2];
    v_scales[3 * g] = as0; v_scales[3 * g + 1] = as1; v_scales[3 * g + 2] = as2;
    float aq0 = acc[6], aq1 = acc[7], aq2 = acc[8], aq3 = acc[9];
    if (flags & kFlagRawQuats) {
      // through q / |q| of the wrapper; a substituted quaternion (|q| reported as 0) gets no gradient
      const float qn = par[10];
      if (qn > 0.f) {
        const float d = par[6] * aq0 + par[7] * aq1 + par[8] * aq2 + par[9] * aq3;
        const float inv = 1.0f / qn;
        aq0 = (aq0 - par[6] * d) * inv; aq1 = (aq1 - par[7] * d) * inv;
        aq2 = (aq2 - par[8] * d) * inv; aq3 = (aq3 - par[9] * d) * inv;
      } else { aq0 = aq1 = aq2 = aq3 = 0.f; }
    }
    reinterpret_cast<float4*>(v_quats)[g] = make_float4(aq0, aq1, aq2, aq3);
    float aop = acc[10];
    if (flags & kFlagRawOpacities) { const float op = acc[11]; aop *= op * (1.0f - op); }
    v_opacities[g] = aop;
    if (DEG < 0) {
      float c0 = acc[12], c1 = acc[13], c2 = acc[14];
      if (flags & kFlagColorSigmoid) {
        c0 *= acc[16] * (1.0f - acc[16]); c1 *= acc[17] * (1.0f - acc[17]); c2 *= acc[18] * (1.0f - acc[18]);
      }
      v_color_src[3 * g] = c0; v_color_src[3 * g + 1] = c1; v_color_src[3 * g + 2] = c2;
    }
  }
  if (DEG >= 0) {
    if (v_color_rest) {
      stage_rows_out(s_sh, v_color_src + (size_t)g0 * 3, nb, 3, kViewsBwdBlock, row_stride, 0);
      if (k3_store > 3) stage_rows_out(s_sh, v_color_rest + (size_t)g0 * (k3_store - 3), nb, k3_store - 3, kViewsBwdBlock, row_stride, 3);
    } else {
      stage_rows_out(s_sh, v_color_src + (size_t)g0 * k3_store, nb, k3_store, kViewsBwdBlock, row_stride, 0);
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
                                    float* v_color_src, float* v_color_rest, int* stats,
                                    sb_stream_t stream) {
  const bool c2w = flags & sb::kFlagC2wOpenGL;
  if (n < 0 || num_views < 1 || block_width < 2 || block_width > 16 || sh_degree < -1 || sh_degree > 4 ||
      (sh_degree >= 0 && !campos && !c2w) || (sh_degree < 0 && v_color_rest))
    return SB_ERR_BAD_ARG;
  if (n == 0) return 0;
  sb::CamSrc cams{viewmats, viewmat_stride, intrins, fx, fy, cx, cy, c2w ? 1 : 0,
                  (flags & sb::kFlagIntrinsShared) ? 1 : 0};
  const int k3 = sh_degree < 0 ? 3 : (sh_degree + 1) * (sh_degree + 1) * 3;
  const size_t smem = sb::views_bwd_smem_bytes(k3, sh_degree >= 0);
  dim3 grid(sb::ceil_div(n, sb::kViewsBwdRows));
  cudaStream_t st = (cudaStream_t)stream;
  const float4* pk = reinterpret_cast<const float4*>(packed);
  const float4* vpk = reinterpret_cast<const float4*>(v_packed);
#define SB_CASE(D)                                                                                   \
  case D:                                                                                            \
    sb::views_project_bwd_kernel<D><<<grid, sb::kViewsBwdBlock, smem, st>>>(                         \
        num_views, n, means, scales, glob_scale, quats, flags, cams, campos, img_h, img_w,           \
        block_width, clip_thresh, k3, pk, vpk, v_means, v_scales, v_quats, v_opacities, v_color_src, \
        v_color_rest, stats);                                                                        \
    break;
  switch (sh_degree) { SB_CASE(-1) SB_CASE(0) SB_CASE(1) SB_CASE(2) SB_CASE(3) SB_CASE(4) }
#undef SB_CASE
  SB_RETURN_LAST();
}
