// Micro-benchmark (B200): peer-mask of equal 8-bit digits inside a warp, MATCH.ANY vs eight ballots.
// nvcc -arch=sm_100a -O3 -o /tmp/mv scripts/micro/match_vs_vote.cu && /tmp/mv
#include <cstdio>
#include <cuda_runtime.h>

template <int MODE>
__global__ void __launch_bounds__(256) k(const unsigned* __restrict__ in, unsigned* __restrict__ out, int rounds) {
  const int lane = threadIdx.x & 31;
  unsigned acc = 0;
  unsigned key = in[blockIdx.x * 256 + threadIdx.x];
  for (int r = 0; r < rounds; ++r) {
    const unsigned d = (key >> (r & 7)) & 255u;
    unsigned m;
    if (MODE == 0) {
      m = __match_any_sync(0xffffffffu, d);
    } else {
      m = 0xffffffffu;
#pragma unroll
      for (int b = 0; b < 8; ++b) {
        const unsigned bal = __ballot_sync(0xffffffffu, (d >> b) & 1u);
        m &= ((d >> b) & 1u) ? bal : ~bal;
      }
    }
    acc += __popc(m & ((1u << lane) - 1u)) + (__ffs(m) - 1);
    key = key * 1664525u + 1013904223u + acc;
  }
  out[blockIdx.x * 256 + threadIdx.x] = acc;
}

int main() {
  const int blocks = 148 * 8, rounds = 4096;
  unsigned *in, *out;
  cudaMalloc(&in, blocks * 256 * 4); cudaMalloc(&out, blocks * 256 * 4);
  cudaMemset(in, 0x5a, blocks * 256 * 4);
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  for (int mode = 0; mode < 2; ++mode) {
    for (int rep = 0; rep < 3; ++rep) {
      cudaEventRecord(a);
      if (mode == 0) k<0><<<blocks, 256>>>(in, out, rounds); else k<1><<<blocks, 256>>>(in, out, rounds);
      cudaEventRecord(b); cudaEventSynchronize(b);
      float ms; cudaEventElapsedTime(&ms, a, b);
      const double warp_ops = (double)blocks * 8 * rounds;
      printf("%s: %.3f ms, %.1f cycles per warp-op per SM (at 1.965 GHz)\n", mode == 0 ? "match.any" : "8 ballots", ms,
             ms * 1e-3 * 1.965e9 / (warp_ops / 148));
    }
  }
  return 0;
}
