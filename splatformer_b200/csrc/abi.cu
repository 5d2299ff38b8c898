// ABI bookkeeping + the small layout kernels around the rasteriser.
#include "raster_common.cuh"

namespace sb {

__global__ void __launch_bounds__(256) pack_kernel(int count, const float* __restrict__ xys,
                                                   const float* __restrict__ conics,
                                                   const float* __restrict__ colors,
                                                   const float* __restrict__ opacity,
                                                   float4* __restrict__ packed) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= count) return;
  const float2 c = __ldg(reinterpret_cast<const float2*>(xys) + i);
  const float ka = __ldg(conics + 3 * i), kb = __ldg(conics + 3 * i + 1), kc = __ldg(conics + 3 * i + 2);
  const float op = __ldg(opacity + i);
  float4 r0, r1, r2;
  pack_record(c.x, c.y, ka, kb, kc, op, __ldg(colors + 3 * i), __ldg(colors + 3 * i + 1), __ldg(colors + 3 * i + 2),
              r0, r1, r2);
  packed[3 * i] = r0; packed[3 * i + 1] = r1; packed[3 * i + 2] = r2;
}

__global__ void __launch_bounds__(256) unpack_grads_kernel(int count, const float4* __restrict__ vp,
                                                           const float* __restrict__ conics,
                                                           const float* __restrict__ opacity,
                                                           float* __restrict__ v_xys,
                                                           float* __restrict__ v_conics,
                                                           float* __restrict__ v_colors,
                                                           float* __restrict__ v_opacity) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= count) return;
  const float4 a = __ldg(vp + 3 * i), b = __ldg(vp + 3 * i + 1), c = __ldg(vp + 3 * i + 2);
  float vx, vy, vca, vcb, vcc, vop;
  moments_to_grads(a, b, __ldg(conics + 3 * i), __ldg(conics + 3 * i + 1), __ldg(conics + 3 * i + 2),
                   __ldg(opacity + i), vx, vy, vca, vcb, vcc, vop);
  reinterpret_cast<float2*>(v_xys)[i] = make_float2(vx, vy);
  v_conics[3 * i] = vca; v_conics[3 * i + 1] = vcb; v_conics[3 * i + 2] = vcc;
  v_opacity[i] = vop;
  v_colors[3 * i] = b.z; v_colors[3 * i + 1] = b.w; v_colors[3 * i + 2] = c.x;
}

// gsplat.utils.compute_cov2d_bounds (upstream _C.compute_cov2d_bounds): conic = inverse of the 2x2
// covariance (a b; b c), radius = ceil(3 sqrt(largest eigenvalue)) with the discriminant floored at
// 0.1 as in the projection kernel; a singular covariance gives conic 0, radius 0.
__global__ void __launch_bounds__(256) cov2d_bounds_kernel(int count, const float* __restrict__ cov2d,
                                                           float* __restrict__ conics, int* __restrict__ radii) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= count) return;
  const float a = __ldg(cov2d + 3 * i), b = __ldg(cov2d + 3 * i + 1), c = __ldg(cov2d + 3 * i + 2);
  // one IEEE rounding per operation, no FMA contraction: same bits as the oracle's float32 chain
  const float det = __fsub_rn(__fmul_rn(a, c), __fmul_rn(b, b));
  float ka = 0.f, kb = 0.f, kc = 0.f;
  int radius = 0;
  if (det != 0.f) {
    const float inv = __fdiv_rn(1.0f, det);
    ka = __fmul_rn(c, inv); kb = __fmul_rn(-b, inv); kc = __fmul_rn(a, inv);
    const float mid = __fmul_rn(0.5f, __fadd_rn(a, c));
    const float disc = __fsqrt_rn(fmaxf(0.1f, __fsub_rn(__fmul_rn(mid, mid), det)));
    const float r = ceilf(__fmul_rn(3.0f, __fsqrt_rn(fmaxf(__fadd_rn(mid, disc), __fsub_rn(mid, disc)))));
    radius = (r >= 0.f && r <= 1073741824.0f) ? (int)r : 0;
  }
  conics[3 * i] = ka; conics[3 * i + 1] = kb; conics[3 * i + 2] = kc;
  radii[i] = radius;
}

}  // namespace sb

extern "C" int sb_cov2d_bounds(int count, const float* cov2d, float* conics, int* radii, sb_stream_t stream) {
  if (count < 0) return SB_ERR_BAD_ARG;
  if (count == 0) return 0;
  sb::cov2d_bounds_kernel<<<sb::ceil_div(count, 256), 256, 0, (cudaStream_t)stream>>>(count, cov2d, conics, radii);
  SB_RETURN_LAST();
}

extern "C" int sb_abi_version(void) { return SB_ABI_VERSION; }

extern "C" const char* sb_error_string(int code) {
  if (code == SB_ERR_BAD_ARG) return "splat_b200: invalid argument";
  if (code == SB_ERR_TEMP_TOO_SMALL) return "splat_b200: scratch buffer too small";
  if (code == SB_ERR_STATIC_SCRATCH_EXCEEDED)
    return "splat_b200: count exceeds the static scratch of the no-allocation entry point "
           "(SB_CUMSUM_STATIC_MAX_COUNT = 16777216 elements); use sb_cumsum_i32 with a caller-supplied buffer";
  return cudaGetErrorString((cudaError_t)code);
}

extern "C" int sb_pack_splats(int count, const float* xys, const float* conics, const float* colors,
                              const float* opacity, float* packed, sb_stream_t stream) {
  if (count < 0) return SB_ERR_BAD_ARG;
  if (count == 0) return 0;
  sb::pack_kernel<<<sb::ceil_div(count, 256), 256, 0, (cudaStream_t)stream>>>(
      count, xys, conics, colors, opacity, reinterpret_cast<float4*>(packed));
  SB_RETURN_LAST();
}

extern "C" int sb_unpack_splat_grads(int count, const float* v_packed, const float* conics,
                                     const float* opacity, float* v_xys, float* v_conics,
                                     float* v_colors, float* v_opacity, sb_stream_t stream) {
  if (count < 0) return SB_ERR_BAD_ARG;
  if (count == 0) return 0;
  sb::unpack_grads_kernel<<<sb::ceil_div(count, 256), 256, 0, (cudaStream_t)stream>>>(
      count, reinterpret_cast<const float4*>(v_packed), conics, opacity, v_xys, v_conics, v_colors, v_opacity);
  SB_RETURN_LAST();
}
