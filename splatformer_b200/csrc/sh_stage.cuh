// Moves a block's rows of an (N, K3) row-major float array between global memory and a padded
// shared-memory layout with float4 accesses (see sh.cu for why).
#pragma once
#include "sb_common.cuh"

namespace sb {

__device__ __forceinline__ int padded_row(int k3) { return k3 | 1; }

// global -> shared.  `src` points at the block's first row; rows * k3 floats are contiguous and the
// span starts 16-byte aligned when (block rows * k3 * 4) % 16 == 0 and the array base is aligned.
// `row` is the shared-memory row stride and `col0` the first column written, so that two global
// arrays (features_dc (N,3) and features_rest (N,K-1,3), /root/reference/utils/gs_utils.py:59-60) can be
// staged side by side into one row without the torch.cat the reference performs per view.
__device__ __forceinline__ void stage_rows_in(float* smem, const float* __restrict__ src, int rows,
                                              int k3, int nthreads, int row, int col0 = 0) {
  smem += col0;
  const int total = rows * k3;
  const int total4 = total >> 2;
  const float4* src4 = reinterpret_cast<const float4*>(src);
  for (int i = threadIdx.x; i < total4; i += nthreads) {
    const float4 v = __ldg(src4 + i);
    const int e = i << 2;
    int g = e / k3, k = e - g * k3;
    const float vals[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      smem[g * row + k] = vals[j];
      if (++k == k3) { k = 0; ++g; }
    }
  }
  for (int e = (total4 << 2) + threadIdx.x; e < total; e += nthreads) {
    const int g = e / k3, k = e - g * k3;
    smem[g * row + k] = __ldg(src + e);
  }
}

// shared -> global
__device__ __forceinline__ void stage_rows_out(const float* smem, float* __restrict__ dst, int rows,
                                               int k3, int nthreads, int row, int col0 = 0) {
  smem += col0;
  const int total = rows * k3;
  const int total4 = total >> 2;
  float4* dst4 = reinterpret_cast<float4*>(dst);
  for (int i = threadIdx.x; i < total4; i += nthreads) {
    const int e = i << 2;
    int g = e / k3, k = e - g * k3;
    float vals[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      vals[j] = smem[g * row + k];
      if (++k == k3) { k = 0; ++g; }
    }
    dst4[i] = make_float4(vals[0], vals[1], vals[2], vals[3]);
  }
  for (int e = (total4 << 2) + threadIdx.x; e < total; e += nthreads) {
    const int g = e / k3, k = e - g * k3;
    dst[e] = smem[g * row + k];
  }
}

}  // namespace sb
