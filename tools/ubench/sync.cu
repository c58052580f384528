// Latency microbenchmark of the synchronisation primitives the epilogues use (sm_100a), one warp:
// mbarrier.try_wait / test_wait on a completed phase, mbarrier.arrive, fence.proxy.async after shared stores.
#include <cstdio>
#include <cuda_runtime.h>
#include <stdint.h>
__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__global__ void k(long long* out) {
  __shared__ uint64_t bar, bar2;
  __shared__ uint4 buf[64];
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(s32(&bar)));
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 32;" ::"r"(s32(&bar2)));
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(s32(&bar)) : "memory");   // phase 0 complete
  }
  __syncthreads();
  const int N = 64;
  uint32_t ok, acc = 0;
  long long t0 = clock64();
  for (int i = 0; i < N; ++i) {
    asm volatile("{.reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0, 1, 0, p;}" : "=r"(ok) : "r"(s32(&bar)) : "memory");
    acc += ok;
  }
  long long t1 = clock64();
  for (int i = 0; i < N; ++i) {
    asm volatile("{.reg .pred p; mbarrier.test_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0, 1, 0, p;}" : "=r"(ok) : "r"(s32(&bar)) : "memory");
    acc += ok;
  }
  long long t2 = clock64();
  for (int i = 0; i < N; ++i) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(s32(&bar2)) : "memory");
  long long t3 = clock64();
  for (int i = 0; i < N; ++i) {
    buf[(threadIdx.x + i) & 63] = make_uint4(i, acc, i, i);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  long long t4 = clock64();
  for (int i = 0; i < N; ++i) {
    buf[(threadIdx.x + i) & 63] = make_uint4(i, acc, i, i);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(s32(&bar2)) : "memory");
  }
  long long t5 = clock64();
  if (threadIdx.x == 0) {
    out[0] = (t1 - t0) / N; out[1] = (t2 - t1) / N; out[2] = (t3 - t2) / N; out[3] = (t4 - t3) / N; out[4] = (t5 - t4) / N; out[5] = acc + buf[3].x;
  }
}
int main() {
  long long* d; cudaMalloc(&d, 64);
  k<<<1, 32>>>(d); cudaDeviceSynchronize();
  long long h[6]; cudaMemcpy(h, d, 48, cudaMemcpyDeviceToHost);
  printf("try_wait (completed phase): %lld cycles; test_wait: %lld; arrive: %lld; STS.128 + fence.proxy.async: %lld; STS + fence + arrive: %lld  (%s)\n", h[0], h[1], h[2], h[3], h[4], cudaGetErrorString(cudaGetLastError()));
  return 0;
}
