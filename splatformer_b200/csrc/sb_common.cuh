// Launch-side helpers shared by the .cu translation units of libsplat_b200.so.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/splat_b200.h"
#include "sb_math.cuh"

#define SB_RETURN_LAST() return (int)cudaGetLastError()

namespace sb {

__host__ __device__ inline int ceil_div(int a, int b) { return (a + b - 1) / b; }

// Camera of view v from the ABI's (viewmats, intrins | scalars) description.
struct CamSrc {
  const float* viewmats;
  int stride;
  const float* intrins;  // (V,4) or nullptr
  float fx, fy, cx, cy;
};

__device__ __forceinline__ Cam load_cam(const CamSrc& s, int v) {
  Cam c;
  const float* m = s.viewmats + (size_t)v * s.stride;
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    c.w[3 * i] = __ldg(m + 4 * i); c.w[3 * i + 1] = __ldg(m + 4 * i + 1); c.w[3 * i + 2] = __ldg(m + 4 * i + 2);
    c.p[i] = __ldg(m + 4 * i + 3);
  }
  if (s.intrins) {
    const float* k = s.intrins + 4 * v;
    c.fx = __ldg(k); c.fy = __ldg(k + 1); c.cx = __ldg(k + 2); c.cy = __ldg(k + 3);
  } else {
    c.fx = s.fx; c.fy = s.fy; c.cx = s.cx; c.cy = s.cy;
  }
  return c;
}

}  // namespace sb
