// Launch-side helpers shared by the .cu translation units of libsplat_b200.so.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/splat_b200.h"
#include "sb_math.cuh"

#define SB_RETURN_LAST() return (int)cudaGetLastError()

namespace sb {

__host__ __device__ inline int ceil_div(int a, int b) { return (a + b - 1) / b; }

// Division of a non-negative int (< 2^31) by a fixed positive divisor with one multiply-high, one add and one shift
// (Granlund & Montgomery, "Division by invariant integers using multiplication"): the rasterisers split a
// rectangle number into (view, tile row, tile column) once per rectangle, three hardware-less ~20-instruction
// divisions where the list is often empty and the rest of the rectangle's work is a few dozen instructions.
struct FastDiv {
  unsigned mul, shr;
  int d;
};
inline FastDiv make_fast_div(int d) {
  FastDiv f;
  f.d = d;
  unsigned l = 0;
  while ((1ull << l) < (unsigned long long)d) ++l;                 // ceil(log2 d)
  f.mul = (unsigned)(((1ull << 32) * ((1ull << l) - (unsigned long long)d)) / (unsigned long long)d + 1ull);
  f.shr = l;
  return f;
}
__device__ __forceinline__ int fast_div(int n, const FastDiv& f) {
  return (int)(((unsigned)n + __umulhi((unsigned)n, f.mul)) >> f.shr);      // n < 2^31: the sum cannot wrap
}

// Camera of view v from the ABI's (viewmats, intrins | scalars) description.
struct CamSrc {
  const float* viewmats;
  int stride;
  const float* intrins;  // (V,4) or nullptr
  float fx, fy, cx, cy;
  int c2w_opengl;        // viewmats holds OpenGL camera-to-world matrices (gs_utils.py:32-42 is done here)
  int intr_shared;       // intrins is (1,4): one set of intrinsics for all views (dataset/GS.py:392-394)
};

__device__ __forceinline__ Cam load_cam(const CamSrc& s, int v) {
  Cam c;
  const float* m = s.viewmats + (size_t)v * s.stride;
  if (s.c2w_opengl) {
    // R <- R diag(1,-1,-1); viewmat = [R^T | -R^T T]   (/root/reference/utils/gs_utils.py:32-42)
    const float t0 = __ldg(m + 3), t1 = __ldg(m + 7), t2 = __ldg(m + 11);
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      const float sgn = i == 0 ? 1.0f : -1.0f;
      const float a = sgn * __ldg(m + i), b = sgn * __ldg(m + 4 + i), d = sgn * __ldg(m + 8 + i);
      c.w[3 * i] = a; c.w[3 * i + 1] = b; c.w[3 * i + 2] = d;
      c.p[i] = -(a * t0 + b * t1 + d * t2);
    }
  } else {
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      c.w[3 * i] = __ldg(m + 4 * i); c.w[3 * i + 1] = __ldg(m + 4 * i + 1); c.w[3 * i + 2] = __ldg(m + 4 * i + 2);
      c.p[i] = __ldg(m + 4 * i + 3);
    }
  }
  if (s.intrins) {
    const float* k = s.intrins + (s.intr_shared ? 0 : 4 * v);
    c.fx = __ldg(k); c.fy = __ldg(k + 1); c.cx = __ldg(k + 2); c.cy = __ldg(k + 3);
  } else {
    c.fx = s.fx; c.fy = s.fy; c.cx = s.cx; c.cy = s.cy;
  }
  return c;
}

// All views' cameras, built once per block into shared memory (world->camera matrix incl. the c2w inversion,
// intrinsics, EWA clamp limits, camera centre: kCamFloats floats per view) instead of once per (view, Gaussian)
// pair.  Needs a __syncthreads() before use.
constexpr int kCamFloats = 24;     // w[9] p[3] fx fy cx cy | lim_x lim_y | centre[3] | pad
constexpr int kMaxStagedCams = 64;

__device__ __forceinline__ void stage_cams(const CamSrc& s, const float* __restrict__ campos, int num_views,
                                           int img_w, int img_h, float* __restrict__ smem_cams) {
  for (int v = threadIdx.x; v < num_views; v += blockDim.x) {
    const Cam c = load_cam(s, v);
    float* d = smem_cams + v * kCamFloats;
#pragma unroll
    for (int i = 0; i < 9; ++i) d[i] = c.w[i];
    d[9] = c.p[0]; d[10] = c.p[1]; d[11] = c.p[2];
    d[12] = c.fx; d[13] = c.fy; d[14] = c.cx; d[15] = c.cy;
    float lx, ly;
    cam_limits(c, img_w, img_h, lx, ly);
    d[16] = lx; d[17] = ly;
    const float* m = s.viewmats + (size_t)v * s.stride;
    if (s.c2w_opengl) { d[18] = __ldg(m + 3); d[19] = __ldg(m + 7); d[20] = __ldg(m + 11); }
    else if (campos) { d[18] = __ldg(campos + 3 * v); d[19] = __ldg(campos + 3 * v + 1); d[20] = __ldg(campos + 3 * v + 2); }
    else { d[18] = d[19] = d[20] = 0.f; }
  }
}

__device__ __forceinline__ Cam cam_from_smem(const float* __restrict__ d, float& lim_x, float& lim_y) {
  Cam c;
#pragma unroll
  for (int i = 0; i < 9; ++i) c.w[i] = d[i];
  c.p[0] = d[9]; c.p[1] = d[10]; c.p[2] = d[11];
  c.fx = d[12]; c.fy = d[13]; c.cx = d[14]; c.cy = d[15];
  lim_x = d[16]; lim_y = d[17];
  return c;
}

// Camera centre of view v (for the SH view direction, gs_utils.py:67-69).
__device__ __forceinline__ void load_campos(const CamSrc& s, const float* __restrict__ campos, int v, float* out) {
  if (s.c2w_opengl) {
    const float* m = s.viewmats + (size_t)v * s.stride;
    out[0] = __ldg(m + 3); out[1] = __ldg(m + 7); out[2] = __ldg(m + 11);
  } else {
    out[0] = __ldg(campos + 3 * v); out[1] = __ldg(campos + 3 * v + 1); out[2] = __ldg(campos + 3 * v + 2);
  }
}

// Input-convention flags of the fused multi-view kernels (sb_views_project_fwd / _bwd).
constexpr int kFlagRawScales = 1;      // scales are log-scales: scale = exp(raw)            (gs_utils.py:45)
constexpr int kFlagRawOpacities = 2;   // opacities are logits: opacity = sigmoid(raw)       (gs_utils.py:54)
constexpr int kFlagColorSigmoid = 4;   // sh_degree = -1 only: colour = sigmoid(color_src)   (gs_utils.py:65)
constexpr int kFlagC2wOpenGL = 8;      // viewmats are OpenGL camera-to-world matrices       (gs_utils.py:32-42)
constexpr int kFlagIntrinsShared = 16; // intrins is (1,4), the same fx fy cx cy for every view (dataset/GS.py:392-394)
constexpr int kFlagRawQuats = 32;      // quats are raw: q / |q|, degenerate ones -> (0,0,0,1)  (gs_utils.py:46-51)

__device__ __forceinline__ float sigmoidf_(float x) { return 1.0f / (1.0f + expf(-x)); }

}  // namespace sb
