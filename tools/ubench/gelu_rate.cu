// Rate of the MLP kernels' epilogue arithmetic alone: per thread 16 accumulator values -> exact-erf GELU (gelu2x_8) -> split-bf16
// pairs (split2p), as in epi_mlp_tile, on register data with W warps per SM and nothing else on the machine.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I dlwp-benchmark_b200/csrc -o tools/ubench/gelu_rate tools/ubench/gelu_rate.cu
#include <cstdio>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "tc_mlp.cuh"

using namespace dlwp;
using namespace dlwp::tc;

__global__ void __launch_bounds__(1024, 1) k(int iters, uint32_t* out, long long* cyc) {
  uint32_t r[16];
  for (int e = 0; e < 16; ++e) r[e] = __float_as_uint(0.01f * (float)((threadIdx.x * 16 + e) % 700) - 3.5f);
  uint32_t acc = 0;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    float2 g2[8];
    gelu2x_8(r, g2);
    uint32_t h[8], l[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) split2p(g2[e], h[e], l[e]);
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      acc ^= h[e] + l[e];
      r[2 * e] = r[2 * e] ^ (l[e] & 0x00010000u);      // keeps the next round dependent on this one without changing the range
      r[2 * e + 1] = r[2 * e + 1] ^ (h[e] & 0x00000001u);
    }
  }
  const long long t1 = clock64();
  __syncthreads();
  if (threadIdx.x == 0 && blockIdx.x == 0) cyc[0] = t1 - t0;
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

int main() {
  uint32_t* out;
  long long* cyc;
  cudaMalloc(&out, 148 * 1024 * 4);
  cudaMalloc(&cyc, 8);
  const int iters = 2000;
  for (int warps : {4, 8, 16, 24, 32}) {
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    k<<<148, warps * 32, 200 * 1024>>>(iters, out, cyc);
    long long c = 0;
    if (cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost) != cudaSuccess) {
      printf("failed: %s\n", cudaGetErrorString(cudaGetLastError()));
      return 1;
    }
    const double per_chunk = (double)c / iters;                 // cycles per 16-values-per-thread round of one warp
    const double per_sched = per_chunk / (warps / 4.0);         // per scheduler: cycles per warp-round
    printf("%2d warps / SM: %.1f cycles per warp round (16 values per thread), %.1f per scheduler -> %.2f elements / cycle / SM\n", warps, per_chunk,
           per_sched, 512.0 * warps / per_chunk);
  }
  return 0;
}
