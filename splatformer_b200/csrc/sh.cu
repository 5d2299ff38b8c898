// Spherical-harmonics colour, forward and backward (reference call site
// /root/reference/utils/gs_utils.py:78).  Streaming kernels: 12*K+24 bytes per Gaussian.
//
// Layout problem: coeffs is (N, K, 3) = K*3 consecutive floats per Gaussian.  One thread per
// Gaussian reading its own row would issue stride-(12K)-byte requests.  Instead a block moves
// its 128 rows (one contiguous span of global memory) through shared memory with float4
// accesses and the threads read/write their row from a padded (conflict-free) shared layout.
#include "sh_stage.cuh"

namespace sb {

constexpr int kShBlock = 128;

template <int DEG>
__global__ void __launch_bounds__(kShBlock) sh_fwd_kernel(int num_views, int n, int k3_store,
                                                          const float* __restrict__ dirs,
                                                          const float* __restrict__ coeffs,
                                                          float* __restrict__ colors) {
  extern __shared__ float smem[];
  constexpr int KU = (DEG + 1) * (DEG + 1);
  const int row = padded_row(k3_store);
  const int g0 = blockIdx.x * kShBlock;
  const int nb = min(kShBlock, n - g0);
  stage_rows_in(smem, coeffs + (size_t)g0 * k3_store, nb, k3_store, kShBlock, row);
  __syncthreads();
  const int g = g0 + threadIdx.x;
  if (threadIdx.x >= nb) return;
  float c[KU * 3];
#pragma unroll
  for (int i = 0; i < KU * 3; ++i) c[i] = smem[threadIdx.x * row + i];
  for (int v = 0; v < num_views; ++v) {
    const float* d = dirs + ((size_t)v * n + g) * 3;
    float Y[KU];
    sh_basis(DEG, __ldg(d), __ldg(d + 1), __ldg(d + 2), Y);
    float r = 0.f, gg = 0.f, b = 0.f;
#pragma unroll
    for (int k = 0; k < KU; ++k) { r += Y[k] * c[3 * k]; gg += Y[k] * c[3 * k + 1]; b += Y[k] * c[3 * k + 2]; }
    float* o = colors + ((size_t)v * n + g) * 3;
    o[0] = r; o[1] = gg; o[2] = b;
  }
}

template <int DEG>
__global__ void __launch_bounds__(kShBlock) sh_bwd_kernel(int num_views, int n, int k3_store,
                                                          const float* __restrict__ dirs,
                                                          const float* __restrict__ v_colors,
                                                          float* __restrict__ v_coeffs) {
  extern __shared__ float smem[];
  constexpr int KU = (DEG + 1) * (DEG + 1);
  const int row = padded_row(k3_store);
  const int g0 = blockIdx.x * kShBlock;
  const int nb = min(kShBlock, n - g0);
  const int g = g0 + threadIdx.x;
  if (threadIdx.x < nb) {
    float acc[KU * 3];
#pragma unroll
    for (int i = 0; i < KU * 3; ++i) acc[i] = 0.f;
    for (int v = 0; v < num_views; ++v) {
      const size_t o = ((size_t)v * n + g) * 3;
      float Y[KU];
      sh_basis(DEG, __ldg(dirs + o), __ldg(dirs + o + 1), __ldg(dirs + o + 2), Y);
      const float vr = __ldg(v_colors + o), vg = __ldg(v_colors + o + 1), vb = __ldg(v_colors + o + 2);
#pragma unroll
      for (int k = 0; k < KU; ++k) { acc[3 * k] += Y[k] * vr; acc[3 * k + 1] += Y[k] * vg; acc[3 * k + 2] += Y[k] * vb; }
    }
#pragma unroll
    for (int i = 0; i < KU * 3; ++i) smem[threadIdx.x * row + i] = acc[i];
    for (int i = KU * 3; i < k3_store; ++i) smem[threadIdx.x * row + i] = 0.f;
  }
  __syncthreads();
  stage_rows_out(smem, v_coeffs + (size_t)g0 * k3_store, nb, k3_store, kShBlock, row);
}

template <bool FWD>
static int launch_sh(int num_views, int n, int degree, int degrees_to_use, const float* dirs,
                     const float* in, float* out, cudaStream_t st) {
  if (n < 0 || num_views < 1 || degree < 0 || degree > 4 || degrees_to_use < 0 || degrees_to_use > degree)
    return SB_ERR_BAD_ARG;
  if (n == 0) return 0;
  const int k3 = (degree + 1) * (degree + 1) * 3;
  const size_t smem = (size_t)kShBlock * (k3 | 1) * sizeof(float);
  const dim3 grid(ceil_div(n, kShBlock)), block(kShBlock);
#define SB_SH_CASE(D)                                                                         \
  case D:                                                                                     \
    if (FWD) sh_fwd_kernel<D><<<grid, block, smem, st>>>(num_views, n, k3, dirs, in, out);     \
    else sh_bwd_kernel<D><<<grid, block, smem, st>>>(num_views, n, k3, dirs, in, out);         \
    break;
  switch (degrees_to_use) { SB_SH_CASE(0) SB_SH_CASE(1) SB_SH_CASE(2) SB_SH_CASE(3) SB_SH_CASE(4) }
#undef SB_SH_CASE
  SB_RETURN_LAST();
}

}  // namespace sb

extern "C" int sb_sh_fwd(int num_views, int n, int degree, int degrees_to_use, const float* dirs,
                         const float* coeffs, float* colors, sb_stream_t stream) {
  return sb::launch_sh<true>(num_views, n, degree, degrees_to_use, dirs, coeffs, colors, (cudaStream_t)stream);
}

extern "C" int sb_sh_bwd(int num_views, int n, int degree, int degrees_to_use, const float* dirs,
                         const float* v_colors, float* v_coeffs, sb_stream_t stream) {
  return sb::launch_sh<false>(num_views, n, degree, degrees_to_use, dirs, v_colors, v_coeffs, (cudaStream_t)stream);
}
