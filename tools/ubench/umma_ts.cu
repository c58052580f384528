// tcgen05.mma with the A operand in tensor memory (the ".ts" form: D[tmem] (+)= A[tmem] * B[smem]^T): layout check and throughput.
//   Layout assumed (and checked here against a per-thread fp32 sum): row i of A = TMEM lane i; 32-bit column c of the A region
//   holds the bf16 pair (k = 2c in the low half, k = 2c + 1 in the high half); one K = 16 instruction reads 8 columns.
//   Written by the threads with tcgen05.st.32x32b (lane i of warp w writes TMEM lane 32 (w % 4) + i).
// Throughput: a run of UMMAs with A from tensor memory against the same shapes with A from shared memory (umma.cu).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I dlwp-benchmark_b200/csrc -o tools/ubench/umma_ts tools/ubench/umma_ts.cu
#include <cstdio>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "tc_common.cuh"

using namespace dlwp::tc;

__device__ __forceinline__ void umma_ts_p(uint32_t leader, uint32_t d_tmem, uint32_t a_tmem, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p, q;\n\tsetp.ne.b32 p, %4, 0;\n\tsetp.ne.b32 q, %5, 0;\n\t"
      "@q tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(d_tmem), "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"(accumulate), "r"(leader)
      : "memory");
}
__device__ __forceinline__ void tmem_st8(uint32_t taddr, const uint32_t* r) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]),
               "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7])
               : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

__device__ __forceinline__ float aval(int i, int k) { return (float)((i * 7 + k * 3) % 17 - 8) * 0.125f; }
__device__ __forceinline__ float bval(int n, int k) { return (float)((n * 5 + k) % 13 - 6) * 0.25f; }

template <int N>
__global__ void __launch_bounds__(128, 1) k(int run, float* err, long long* cyc) {
  __shared__ __align__(128) uint8_t sB[2 * N * 32 * 2];      // two K steps: [k / 8][n][8] bf16
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_s;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int i = tid; i < N * 32; i += 128) {
    const int n = i / 32, kk = i % 32;
    reinterpret_cast<__nv_bfloat16*>(sB)[(kk >> 3) * (N * 8) + n * 8 + (kk & 7)] = __float2bfloat16(bval(n, kk));
  }
  if (tid == 0) {
    mbar_init(&bar, 1);
    fence_mbar_init();
  }
  if (warp == 0) {
    tmem_alloc(&tmem_s, 256);
    tmem_relinquish();
  }
  fence_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tlane = tmem_s + ((uint32_t)(warp * 32) << 16);
  {      // A: rows = lanes, K = 32 (two K steps): columns 0..15
    uint32_t r[16];
    for (int c = 0; c < 16; ++c) {
      __nv_bfloat162 v = __floats2bfloat162_rn(aval(tid, 2 * c), aval(tid, 2 * c + 1));
      r[c] = *reinterpret_cast<uint32_t*>(&v);
    }
    tmem_st8(tlane, r);
    tmem_st8(tlane + 8, r + 8);
    tmem_st_wait();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t dcol = 64;
  if (warp == 1) {
    const uint32_t leader = elect_one();
    const uint32_t idesc = idesc_bf16(128, N);
    const uint64_t dB = chunk_desc(smem_u32(sB), N, 0);
    // correctness: D = A[:, 0:32] * B[:, 0:32]^T
    umma_ts_p(leader, tmem_s + dcol, tmem_s, dB, idesc, 0);
    umma_ts_p(leader, tmem_s + dcol, tmem_s + 8, desc_off(dB, 2 * N * 16), idesc, 1);
    umma_commit_p(leader, &bar);
    mbar_wait(&bar, 0);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  {
    float emax = 0.f;
    for (int c0 = 0; c0 < N; c0 += 16) {
      uint32_t r[16];
      tmem_ld16(tlane + dcol + c0, r);
      tmem_ld_wait();
      for (int c = 0; c < 16; ++c) {
        float want = 0.f;
        for (int kk = 0; kk < 32; ++kk) want += aval(tid, kk) * bval(c0 + c, kk);
        emax = fmaxf(emax, fabsf(__uint_as_float(r[c]) - want));
      }
    }
    atomicMax(reinterpret_cast<int*>(err), __float_as_int(emax));
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == 1) {      // throughput: A from tensor memory, then A from shared memory (chunk-major zeros would do: reuse sB as A rows)
    const uint32_t leader = elect_one();
    const uint32_t idesc = idesc_bf16(128, N);
    const uint64_t dB = chunk_desc(smem_u32(sB), N, 0);
    long long t0 = 0, t1 = 0;
    for (int rep = 0; rep < 2; ++rep) {
      t0 = clock64();
      for (int i = 0; i < run; i += 2) {
        umma_ts_p(leader, tmem_s + dcol, tmem_s, dB, idesc, i != 0);
        umma_ts_p(leader, tmem_s + dcol, tmem_s + 8, desc_off(dB, 2 * N * 16), idesc, 1);
      }
      umma_commit_p(leader, &bar);
      mbar_wait(&bar, (rep + 1) & 1);
      t1 = clock64();
    }
    if (leader) cyc[0] = t1 - t0;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem_s, 256);
}

template <int N>
int one(float* derr, long long* dcyc) {
  const int run = 512;
  cudaMemset(derr, 0, 4);
  k<N><<<1, 128>>>(run, derr, dcyc);
  float e = -1.f;
  long long c = 0;
  cudaError_t st = cudaMemcpy(&e, derr, 4, cudaMemcpyDeviceToHost);
  cudaMemcpy(&c, dcyc, 8, cudaMemcpyDeviceToHost);
  if (st != cudaSuccess) {
    printf("N=%d: %s\n", N, cudaGetErrorString(st));
    return 1;
  }
  printf("M=128 N=%3d  A in tensor memory: max |D - fp32 sum| = %.3g   %.1f cycles / UMMA (math %.1f)\n", N, e, (double)c / run, 128.0 * N * 16 / 4096.0);
  return 0;
}

int main() {
  float* derr;
  long long* dcyc;
  cudaMalloc(&derr, 4);
  cudaMalloc(&dcyc, 8);
  return one<16>(derr, dcyc) | one<32>(derr, dcyc) | one<64>(derr, dcyc) | one<128>(derr, dcyc);
}
