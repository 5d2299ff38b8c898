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
// (view, Gaussian), recomputes the forward record and sums the VJPs over the views in registers: four lanes
// share a Gaussian and are joined with two shuffle steps, so every gradient is written once.
// This unit is built WITH FMA contraction (the forward, project_fwd.cu, is not: its outputs are under the bit-exact
// tiling contract, gradients are not): 20 % fewer instructions.  The recomputed record may therefore differ from
// the forward's in the last bit; which pairs were drawn is decided by the forward alone (a pair the rasteriser never
// touched has an all-zero gradient record and is not walked), the recompute's own `valid` only guards the arithmetic.
//
// Most (view, Gaussian) records are all-zero: a Gaussian behind the opaque front of the object never
// reaches a pixel before that pixel's transmittance runs out, and a culled one is never touched.  With
// fixed (lane, view) pairs the ~1000-instruction recompute + VJP still ran whenever ANY lane of the warp
// had work.  So every warp first scans the records of its 8 Gaussians x V views (coalesced: 8 consecutive
// 48-byte records per view) and compacts, per Gaussian, the views with a non-zero record into a list in
// shared memory (one ballot per 32 records); then the Gaussian's four lanes walk ITS list, k = lane, lane
// + 4, ...: the trip count is the longest list / 4 instead of V / 4.
// ------------------------------------------------------------------------------------------------
constexpr int kViewsBwdBlock = 128;
constexpr int kBwdWarps = kViewsBwdBlock / 32;
constexpr int kViewLanesB = 4;                                // lanes sharing one Gaussian
constexpr int kBwdRowsPerWarp = 32 / kViewLanesB;             // Gaussians per warp (8)
constexpr int kViewsBwdRows = kBwdWarps * kBwdRowsPerWarp;    // Gaussians per block
constexpr int kBwdViewChunk = 64;                             // views scanned per compaction round (list capacity)

__device__ __forceinline__ float sum_view_lanes(float x) {
  x += __shfl_xor_sync(0xffffffffu, x, 1);
  x += __shfl_xor_sync(0xffffffffu, x, 2);
  return x;
}
__device__ __forceinline__ float max_view_lanes(float x) {
  x = fmaxf(x, __shfl_xor_sync(0xffffffffu, x, 1));
  x = fmaxf(x, __shfl_xor_sync(0xffffffffu, x, 2));
  return x;
}

// shared memory: [rows][padded row] staging of the SH gradients | [3K][block] per-thread SH sums | counts | lists
__host__ __device__ inline size_t views_bwd_smem_bytes(int k3_store, bool sh, int num_views) {
  const size_t stage = sh ? (size_t)kViewsBwdRows * (k3_store | 1) * sizeof(float) : 0;
  const size_t priv = sh ? (size_t)k3_store * kViewsBwdBlock * sizeof(float) : 0;
  const size_t lists = (size_t)kViewsBwdRows * kBwdViewChunk + (size_t)kViewsBwdRows * sizeof(int);
  const size_t cams = num_views <= kMaxStagedCams ? (size_t)num_views * kCamFloats * sizeof(float) : 0;
  return stage + priv + ((lists + 15) & ~(size_t)15) + cams;
}

template <int DEG>
#ifndef SB_PROJ_BWD_MINB
#define SB_PROJ_BWD_MINB 5        // 96 registers, no spills: 0.170 ms (4: 128 regs 0.175, 3: 0.191, 6: 80 regs with spills)
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
  // the SH sums of a thread live in shared memory (column threadIdx.x of a [3K][block] array: conflict-free,
  // private, no atomics) -- 48 accumulators less in registers than the projection chain can spare
  float* s_priv = smem + kViewsBwdRows * row_stride + threadIdx.x;
  const int k3u = DEG >= 0 ? KU * 3 : 0;
  int* s_cnt = reinterpret_cast<int*>(smem + kViewsBwdRows * row_stride + (DEG >= 0 ? k3_store * kViewsBwdBlock : 0));   // [rows]
  unsigned char* s_list = reinterpret_cast<unsigned char*>(s_cnt + kViewsBwdRows);         // [rows][kBwdViewChunk]
  // cameras: built once per block when there are few enough views
  const bool cams_staged = num_views <= kMaxStagedCams;
  float* s_cams = reinterpret_cast<float*>(
      reinterpret_cast<unsigned char*>(s_cnt) +
      (((size_t)kViewsBwdRows * kBwdViewChunk + kViewsBwdRows * sizeof(int) + 15) & ~(size_t)15));
  if (cams_staged) {
    stage_cams(cams, campos, num_views, img_w, img_h, s_cams);
    __syncthreads();
  }
  const int g0 = blockIdx.x * kViewsBwdRows;
  const int nb = min(kViewsBwdRows, n - g0);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int rowi = threadIdx.x / kViewLanesB, sub = threadIdx.x % kViewLanesB;          // work phase: my Gaussian
  const int g = g0 + rowi;
  const bool active = rowi < nb;
  float acc[3] = {0.f, 0.f, 0.f};                      // DEG < 0: the rgb gradient sums
#pragma unroll
  for (int i = 0; i < k3u; ++i) s_priv[i * kViewsBwdBlock] = 0.f;
  {
    // inactive rows (tail block) read Gaussian n-1 and never find work
    const int gl = active ? g : n - 1;
    const float m[3] = {__ldg(means + 3 * gl), __ldg(means + 3 * gl + 1), __ldg(means + 3 * gl + 2)};
    float s[3] = {__ldg(scales + 3 * gl), __ldg(scales + 3 * gl + 1), __ldg(scales + 3 * gl + 2)};
    if (flags & kFlagRawScales) { s[0] = expf(s[0]); s[1] = expf(s[1]); s[2] = expf(s[2]); }
    const float4 q4 = __ldg(reinterpret_cast<const float4*>(quats) + gl);
    float q[4] = {q4.x, q4.y, q4.z, q4.w};
    float qraw = 1.0f;
    if (flags & kFlagRawQuats) qraw = wrapper_normalise_quat(q);      // gs_utils.py:46-51
    // Sigma is view-independent; so is the chain Sigma -> (scale, quat), which is linear in dL/dSigma: the
    // views' dL/dSigma are summed (six entries) and the chain runs once per Gaussian after the loop
    float sg[6];
    {
      GaussCov gc;
      gaussian_cov3d(s, glob, q, gc);
#pragma unroll
      for (int i = 0; i < 6; ++i) sg[i] = gc.sg[i];
    }
    float am[3] = {0, 0, 0}, ag[6] = {0, 0, 0, 0, 0, 0}, aop = 0.f;
    float op = 0.f, col0 = 0.f, col1 = 0.f, col2 = 0.f;

    const int row0 = warp * kBwdRowsPerWarp;                       // scan phase: the warp's first Gaussian
    const int rows_w = max(0, min(kBwdRowsPerWarp, nb - row0));
    const int scan_r = lane & (kBwdRowsPerWarp - 1);               // scan phase: lane looks at row scan_r, views lane / 8 + 4 j
    const unsigned row_mask = 0x01010101u << scan_r;
    const unsigned lt = (1u << lane) - 1u;
    int n_scanned = 0, n_live = 0;
    for (int vc0 = 0; vc0 < num_views; vc0 += kBwdViewChunk) {
      const int vc = min(kBwdViewChunk, num_views - vc0);
      // ---- scan: record (view vc0 + p / 8, row p % 8); per row, the views whose record is not all-zero ----
      int cnt = 0;
      const int total = vc * kBwdRowsPerWarp;
      // two steps' records are requested before the first is looked at: a step's ballot depends on its loads, so a
      // plain loop exposed one memory round trip per 32 records (ncu source page: 16 % of the kernel's samples on
      // the first compare).  "Not all-zero" is taken on the bits (a -0.0 counts as set: such a pair is walked
      // and adds zeros).
      for (int p0 = 0; p0 < total; p0 += 64) {
        unsigned bits[2] = {0u, 0u};
#pragma unroll
        for (int u = 0; u < 2; ++u) {
          const int p = p0 + 32 * u + lane;
          if (p < total && scan_r < rows_w) {
            const uint4* rec = reinterpret_cast<const uint4*>(v_packed + 3 * ((size_t)(vc0 + (p >> 3)) * n + g0 + row0 + scan_r));
            const uint4 a = __ldg(rec), b = __ldg(rec + 1);
            const unsigned c = __ldg(reinterpret_cast<const unsigned*>(rec + 2));
            bits[u] = a.x | a.y | a.z | a.w | b.x | b.y | b.z | b.w | c;
          }
        }
#pragma unroll
        for (int u = 0; u < 2; ++u) {
          const bool nz = (bits[u] << 1) != 0u;            // the sign bit alone is -0.0
          const unsigned bal = __ballot_sync(0xffffffffu, nz) & row_mask;
          if (nz) s_list[(row0 + scan_r) * kBwdViewChunk + cnt + __popc(bal & lt)] = (unsigned char)((p0 + 32 * u + lane) >> 3);
          cnt += __popc(bal);
        }
      }
      if (lane < kBwdRowsPerWarp) s_cnt[row0 + lane] = cnt;
      n_scanned += (lane < rows_w) ? vc : 0;
      n_live += (lane < kBwdRowsPerWarp) ? cnt : 0;
      __syncwarp();
      // ---- work: the Gaussian's four lanes walk its list ----
      const int my_cnt = s_cnt[rowi];
      const unsigned char* my_list = s_list + rowi * kBwdViewChunk;
      for (int k = sub; k < my_cnt; k += kViewLanesB) {
        const int v = vc0 + my_list[k];
        Cam cam;
        float lim_x, lim_y;
        if (cams_staged) {
          cam = cam_from_smem(s_cams + v * kCamFloats, lim_x, lim_y);
        } else {
          cam = load_cam(cams, v);
          cam_limits(cam, img_w, img_h, lim_x, lim_y);
        }
        // the pair's gradient record and colour row are requested BEFORE the projection is recomputed (its ~150
        // dependent instructions are what hides them; behind it the first use waited a whole round trip: 15 % of
        // the kernel's samples)
        const size_t i = (size_t)v * n + gl;
        const float4 ga = __ldg(v_packed + 3 * i), gb = __ldg(v_packed + 3 * i + 1);
        const float gcx = __ldg(reinterpret_cast<const float*>(v_packed + 3 * i + 2));
        const float4 col = __ldg(packed + 3 * i + 2);       // r, g, b (after the clamp), opacity
        ProjFwd f;
        project_fwd_view(m, sg, cam, lim_x, lim_y, img_w, img_h, bw, clip, f);
        if (!f.valid) continue;                        // (a culled pair's record is never written: not reached)
        op = col.w; col0 = col.x; col1 = col.y; col2 = col.z;
        float vm[3], G[9];
        float vx, vy, vca, vcb, vcc, vop;
        moments_to_grads(ga, gb, f.ca, f.cb, f.cc, op, vx, vy, vca, vcb, vcc, vop);
        project_bwd_view(m, sg, cam, f, vx, vy, 0.f, vca, vcb, vcc, 0.f, vm, G, nullptr);
#pragma unroll
        for (int j = 0; j < 3; ++j) am[j] += vm[j];
        ag[0] += G[0]; ag[1] += G[1]; ag[2] += G[2]; ag[3] += G[4]; ag[4] += G[5]; ag[5] += G[8];
        aop += vop;
        float vr = gb.z, vg = gb.w, vb = gcx;
        if (DEG >= 0) {
          // clamp(x + 0.5, min 0): gradient passes where > 0
          vr = col.x > 0.f ? vr : 0.f; vg = col.y > 0.f ? vg : 0.f; vb = col.z > 0.f ? vb : 0.f;
          float Y[KU], cp[3];
          if (cams_staged) { const float* d = s_cams + v * kCamFloats; cp[0] = d[18]; cp[1] = d[19]; cp[2] = d[20]; }
          else load_campos(cams, campos, v, cp);
          sh_basis(DEG, m[0] - cp[0], m[1] - cp[1], m[2] - cp[2], Y);
#pragma unroll
          for (int j = 0; j < KU; ++j) {
            s_priv[(3 * j) * kViewsBwdBlock] = fmaf(Y[j], vr, s_priv[(3 * j) * kViewsBwdBlock]);
            s_priv[(3 * j + 1) * kViewsBwdBlock] = fmaf(Y[j], vg, s_priv[(3 * j + 1) * kViewsBwdBlock]);
            s_priv[(3 * j + 2) * kViewsBwdBlock] = fmaf(Y[j], vb, s_priv[(3 * j + 2) * kViewsBwdBlock]);
          }
        } else {
          acc[0] += vr; acc[1] += vg; acc[2] += vb;
        }
      }
      __syncwarp();
    }
    if (stats) {
      n_scanned = __reduce_add_sync(0xffffffffu, n_scanned);
      n_live = __reduce_add_sync(0xffffffffu, n_live);
      if (lane == 0 && n_scanned > 0) { atomicAdd(stats, n_scanned); atomicAdd(stats + 1, n_live); }
    }
    // join the four view lanes of the Gaussian (op / colour are view-independent: take any lane that saw them)
#pragma unroll
    for (int k = 0; k < 3; ++k) am[k] = sum_view_lanes(am[k]);
#pragma unroll
    for (int k = 0; k < 6; ++k) ag[k] = sum_view_lanes(ag[k]);
    aop = sum_view_lanes(aop);
    // Sigma -> (scale, quat), once, from the summed dL/dSigma
    float as[3], aq[4];
    {
      GaussCov gc;
      gaussian_cov3d(s, glob, q, gc);
      const float G[9] = {ag[0], ag[1], ag[2], ag[1], ag[3], ag[4], ag[2], ag[4], ag[5]};
      cov3d_chain_bwd(s, glob, gc.M, gc.R, gc.qn, gc.qnorm, G, as, aq);
    }
    if (DEG < 0) { acc[0] = sum_view_lanes(acc[0]); acc[1] = sum_view_lanes(acc[1]); acc[2] = sum_view_lanes(acc[2]); }
    op = max_view_lanes(op);
    if (DEG < 0 && (flags & kFlagColorSigmoid)) {
      // colours are in (0, 1) where a view drew the Gaussian and 0 in lanes that saw none
      col0 = max_view_lanes(col0); col1 = max_view_lanes(col1); col2 = max_view_lanes(col2);
    }
    // chain rule through the activations the kernel applied itself (gs_utils.py:45-51,54,65)
    if (flags & kFlagRawScales) { as[0] *= s[0]; as[1] *= s[1]; as[2] *= s[2]; }
    if (flags & kFlagRawQuats) {
      // through the wrapper's q / |q|; a substituted quaternion (|q| reported as 0) gets no gradient
      if (qraw > 0.f) {
        const float d = q[0] * aq[0] + q[1] * aq[1] + q[2] * aq[2] + q[3] * aq[3];
        const float inv = 1.0f / qraw;
#pragma unroll
        for (int k = 0; k < 4; ++k) aq[k] = (aq[k] - q[k] * d) * inv;
      } else {
        aq[0] = aq[1] = aq[2] = aq[3] = 0.f;
      }
    }
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
    // join the four lanes' SH sums into the staging rows (lane `sub` takes every fourth coefficient)
    __syncwarp();
    if (active) {
      const float* col = smem + kViewsBwdRows * row_stride + rowi * kViewLanesB;
      for (int i = sub; i < k3u; i += kViewLanesB) {
        const float* c = col + i * kViewsBwdBlock;
        smem[rowi * row_stride + i] = (c[0] + c[1]) + (c[2] + c[3]);
      }
      for (int i = k3u + sub; i < k3_store; i += kViewLanesB) smem[rowi * row_stride + i] = 0.f;   // (bases not used)
    }
    __syncthreads();
    if (v_color_rest) {
      stage_rows_out(smem, v_color_src + (size_t)g0 * 3, nb, 3, kViewsBwdBlock, row_stride, 0);
      if (k3_store > 3) stage_rows_out(smem, v_color_rest + (size_t)g0 * (k3_store - 3), nb, k3_store - 3, kViewsBwdBlock, row_stride, 3);
    } else {
      stage_rows_out(smem, v_color_src + (size_t)g0 * k3_store, nb, k3_store, kViewsBwdBlock, row_stride, 0);
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
  const size_t smem = sb::views_bwd_smem_bytes(k3, sh_degree >= 0, num_views);
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
