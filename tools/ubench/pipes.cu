// Issue-rate microbenchmark of the instructions the epilogues are made of (sm_100a): cycles per warp instruction
// per SM sub-partition, N independent chains per thread, W warps per sub-partition.
#include <cstdio>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

template <int OP>
__global__ void k(unsigned long long* out, float seed, int iters) {
  float a[8], b[8];
  uint32_t u[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) { a[i] = seed + i + threadIdx.x; b[i] = seed * 0.5f + i; u[i] = threadIdx.x * 77u + i; }
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (OP == 0) { asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(a[i])); }
      if (OP == 1) { __nv_bfloat162 h = __floats2bfloat162_rn(a[i], b[i]); u[i] ^= *reinterpret_cast<uint32_t*>(&h); a[i] = __uint_as_float(u[i]); }   // F2FP + LOP
      if (OP == 2) { float2 x = __ffma2_rn(make_float2(a[i], b[i]), make_float2(1.0001f, 0.9999f), make_float2(0.5f, 0.25f)); a[i] = x.x; b[i] = x.y; }
      if (OP == 3) { u[i] = (u[i] & 0xffff0000u) ^ (u[i] >> 3); }                                                    // LOP3 (+SHF)
      if (OP == 4) { u[i] = __byte_perm(u[i], u[(i + 1) & 7], 0x7632); }                                          // PRMT
      if (OP == 5) { a[i] = fmaf(a[i], 1.0001f, 0.5f); }                                                           // FFMA imm
      if (OP == 6) { a[i] = fmaf(a[i], b[i], a[i]); }                                                              // FFMA reg
      if (OP == 7) { u[i] ^= 0x12345u + u[i]; }   // IADD/LOP
      if (OP == 8) { asm volatile("{ .reg .b32 t; cvt.rn.bf16x2.f32 t, %1, %2; xor.b32 %0, %0, t; }" : "+r"(u[i]) : "f"(a[i]), "f"(b[i])); }   // F2FP with constant inputs
    }
  }
  const long long t1 = clock64();
  float s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += a[i] + b[i] + __uint_as_float(u[i]);
  if (threadIdx.x == 0) out[blockIdx.x] = (unsigned long long)(t1 - t0);
  if (s == 123.456f) out[1000] = 1;
}

template <int OP>
void run(const char* name, unsigned long long* d) {
  for (int warps : {4, 8, 16}) {
    const int iters = 2000;
    k<OP><<<1, warps * 32>>>(d, 1.5f, iters);
    cudaDeviceSynchronize();
    unsigned long long c;
    cudaMemcpy(&c, d, 8, cudaMemcpyDeviceToHost);
    // warp instructions per sub-partition = iters * 8 * (warps / 4)
    printf("%-28s warps/SMSP %d: %.2f cycles per warp instruction per SMSP\n", name, warps / 4, (double)c / (iters * 8.0 * (warps / 4)));
  }
}

int main() {
  unsigned long long* d;
  cudaMalloc(&d, 8192);
  run<0>("MUFU.EX2", d);
  run<1>("F2FP.BF16.PACK_AB + LOP3", d);
  run<8>("F2FP (const inputs) + LOP3", d);
  run<2>("FFMA2 (imm)", d);
  run<3>("LOP3 + SHF", d);
  run<4>("PRMT", d);
  run<5>("FFMA imm", d);
  run<6>("FFMA reg", d);
  run<7>("IADD3 + LOP3", d);
  return 0;
}
