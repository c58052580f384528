// Throughput microbenchmark of tcgen05.mma (kind::f16, bf16 x bf16 -> fp32, cta_group::1, K = 16 per instruction)
// by accumulator shape and shared-memory operand layout: one thread issues a run of UMMAs on fixed operands,
// commits, waits; cycles per instruction = (t1 - t0) / run.  What it answers: is a small-N UMMA bound by the
// tensor pipe (M x N x 16 / 8192 cycles) or by the operand fetch (A and B bytes per instruction), and does the
// SWIZZLE_NONE "chunk-major" layout of this library fetch slower than a 128-byte-swizzled one?
// (Results are not checked: the operands are zero; only the fetch / issue pattern matters.)
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I dlwp-benchmark_b200/csrc -o tools/ubench/umma tools/ubench/umma.cu
#include <cstdio>
#include <cuda_runtime.h>
#include <stdint.h>

#include "tc_common.cuh"

using namespace dlwp::tc;

struct Case {
  int M, N;
  int layout;      // 0: both K-major, no swizzle (chunk-major);  1: B MN-major, no swizzle;  2: both K-major, 128-byte swizzle
  int ksteps;      // distinct K steps cycled through (operand addresses advance like in a real K loop)
};

__device__ __forceinline__ uint64_t desc_sw128(uint32_t saddr) {      // K-major, SWIZZLE_128B: SBO = 1024 (8 rows x 128 B)
  return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)1 << 16) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
}

__global__ void __launch_bounds__(128, 1) k(const Case c, int run, long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_s;
  const int tid = threadIdx.x, warp = tid >> 5;
  for (int i = tid; i < 160 * 1024 / 16; i += 128) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
  if (tid == 0) {
    mbar_init(&bar, 1);
    fence_mbar_init();
  }
  if (warp == 0) {
    tmem_alloc(&tmem_s, 512);
    tmem_relinquish();
  }
  fence_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == 1) {      // converged warp, one elected lane issues (the library's control warps work this way: uniform registers)
    const uint32_t leader = elect_one();
    const uint32_t aA = smem_u32(smem), aB = smem_u32(smem + 64 * 1024);
    uint32_t idesc = idesc_bf16((uint32_t)c.M, (uint32_t)c.N);
    if (c.layout == 1) idesc |= kIdescBMnMajor;
    const uint64_t dA0 = c.layout == 2 ? desc_sw128(aA) : chunk_desc(aA, (uint32_t)c.M, 0);
    const uint64_t dB0 = c.layout == 2 ? desc_sw128(aB) : c.layout == 1 ? smem_desc(aB, 128, 1024) : chunk_desc(aB, (uint32_t)c.N, 0);
    const uint32_t stepA = c.layout == 2 ? 32u : (uint32_t)c.M * 32u, stepB = c.layout == 2 ? 32u : c.layout == 1 ? 256u : (uint32_t)c.N * 32u;
    long long t0 = 0, t1 = 0;
    for (int rep = 0; rep < 2; ++rep) {      // the first repetition warms up
      t0 = clock64();
      for (int i = 0; i < run; i += 4) {
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) {
          const uint32_t kk = ks < c.ksteps ? ks : 0;
          umma_bf16_p(leader, tmem_s, desc_off(dA0, kk * stepA), desc_off(dB0, kk * stepB), idesc, (i | ks) != 0);
        }
      }
      umma_commit_p(leader, &bar);
      mbar_wait(&bar, rep & 1);
      t1 = clock64();
    }
    if (leader) out[0] = t1 - t0;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem_s, 512);
}

int main() {
  long long* d;
  cudaMalloc(&d, 8);
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
  const Case cases[] = {{128, 256, 0, 4}, {128, 128, 0, 4}, {128, 64, 0, 4}, {128, 32, 0, 2}, {128, 16, 0, 1}, {64, 256, 0, 4}, {64, 64, 0, 4},
                        {64, 256, 1, 4}, {128, 64, 1, 4}, {128, 256, 1, 4}, {64, 64, 1, 4}, {64, 128, 1, 4}, {64, 32, 1, 4},
                        {128, 256, 2, 4}, {128, 128, 2, 4}, {128, 64, 2, 4}, {128, 32, 2, 2}, {64, 256, 2, 4}};
  const int run = 512;
  printf("%5s %5s %7s %12s %12s %14s %12s\n", "M", "N", "layout", "cyc/UMMA", "math cyc", "operand bytes", "bytes/cyc");
  for (const Case& c : cases) {
    k<<<1, 128, 160 * 1024>>>(c, run, d);
    long long h = 0;
    cudaError_t e = cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) {
      printf("M=%d N=%d layout=%d: %s\n", c.M, c.N, c.layout, cudaGetErrorString(e));
      return 1;
    }
    const double cyc = (double)h / run, bytes = (c.M + c.N) * 32.0;
    printf("%5d %5d %7s %12.1f %12.1f %14.0f %12.1f\n", c.M, c.N, c.layout == 0 ? "none" : c.layout == 1 ? "none/Bmn" : "sw128", cyc,
           (double)c.M * c.N * 16 / 4096.0, bytes, bytes / cyc);
  }
  return 0;
}
