// Self-test of the tcgen05 building blocks (tc_common.cuh): one CTA computes
//   D[M x N] = A[M x K] * B[N x K]^T   (bf16 operands, fp32 accumulate in TMEM)
// and dumps the raw TMEM contents (128 lanes x N columns), so a test can check both the
// arithmetic and where an M = 64 accumulator lands.  Not on the hot path.
#include "common.cuh"
#include "tc_common.cuh"

namespace dlwp {

using namespace tc;

// bmn = 0: B is [N][K] row-major, staged K-major (chunk-major) like A.
// bmn = 1 | 2: B is [K][N] row-major and staged the way the kernels keep an activation tile image
// (rows = k, 16-byte chunks of 8 consecutive n): offset(k, n) = (n / 8) * (K * 16) + k * 16 + (n % 8) * 2,
// read as an MN-major operand: 8 (k) x 8 (n) core matrices of 128 contiguous bytes, K-adjacent core
// matrices 128 B apart, N-adjacent ones K * 16 B apart.  bmn = 1 puts the K stride in LBO and the
// N stride in SBO; bmn = 2 swaps them (kept to pin the encoding on hardware).
__global__ void __launch_bounds__(128) umma_selftest_kernel(const float* __restrict__ A, const float* __restrict__ B,
                                                            float* __restrict__ Draw, int M, int N, int K, int bmn) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_base_s;
  uint8_t* sA = smem;                       // [M rows][K] chunk-major
  uint8_t* sB = smem + (size_t)M * K * 2;   // [N rows][K] chunk-major
  const int tid = threadIdx.x, warp = tid >> 5;

  if (tid == 0) {
    mbar_init(&bar, 1);
    fence_mbar_init();
  }
  if (warp == 0) {
    tmem_alloc(&tmem_base_s, 256);
    tmem_relinquish();
  }
  // operands: fp32 global -> bf16 chunk-major shared
  for (int i = tid; i < M * (K / 8); i += 128) {
    const int r = i % M, kc = i / M;
    uint32_t w[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      __nv_bfloat162 v = __floats2bfloat162_rn(A[(size_t)r * K + kc * 8 + 2 * j], A[(size_t)r * K + kc * 8 + 2 * j + 1]);
      w[j] = *reinterpret_cast<uint32_t*>(&v);
    }
    *reinterpret_cast<uint4*>(sA + (size_t)kc * M * 16 + r * 16) = make_uint4(w[0], w[1], w[2], w[3]);
  }
  if (bmn != 0) {
    for (int i = tid; i < K * (N / 8); i += 128) {
      const int k = i % K, nc = i / K;
      uint32_t w[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        __nv_bfloat162 v = __floats2bfloat162_rn(B[(size_t)k * N + nc * 8 + 2 * j], B[(size_t)k * N + nc * 8 + 2 * j + 1]);
        w[j] = *reinterpret_cast<uint32_t*>(&v);
      }
      *reinterpret_cast<uint4*>(sB + (size_t)nc * K * 16 + k * 16) = make_uint4(w[0], w[1], w[2], w[3]);
    }
  }
  for (int i = tid; bmn == 0 && i < N * (K / 8); i += 128) {
    const int r = i % N, kc = i / N;
    uint32_t w[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      __nv_bfloat162 v = __floats2bfloat162_rn(B[(size_t)r * K + kc * 8 + 2 * j], B[(size_t)r * K + kc * 8 + 2 * j + 1]);
      w[j] = *reinterpret_cast<uint32_t*>(&v);
    }
    *reinterpret_cast<uint4*>(sB + (size_t)kc * N * 16 + r * 16) = make_uint4(w[0], w[1], w[2], w[3]);
  }
  fence_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_base_s;

  if (tid == 0) {
    const uint32_t idesc = idesc_bf16(M, N) | (bmn != 0 ? kIdescBMnMajor : 0u);
    for (int k0 = 0; k0 < K; k0 += 16) {
      uint64_t bdesc;
      if (bmn == 0) bdesc = chunk_desc(smem_u32(sB), N, k0);
      else if (bmn == 1) bdesc = smem_desc(smem_u32(sB) + k0 * 16, 128, K * 16);
      else bdesc = smem_desc(smem_u32(sB) + k0 * 16, K * 16, 128);
      umma_bf16(tmem, chunk_desc(smem_u32(sA), M, k0), bdesc, idesc, k0 > 0);
    }
    umma_commit(&bar);
  }
  mbar_wait(&bar, 0);
  tc_fence_after();

  // dump all 128 lanes
  const int lane_row = warp * 32 + (tid & 31);
  for (int c0 = 0; c0 < N; c0 += 16) {
    uint32_t r[16];
    tmem_ld16(tmem + ((uint32_t)(warp * 32) << 16) + c0, r);
    tmem_ld_wait();
#pragma unroll
    for (int j = 0; j < 16; ++j)
      if (c0 + j < N) Draw[(size_t)lane_row * N + c0 + j] = __uint_as_float(r[j]);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 256);
}

}  // namespace dlwp

using namespace dlwp;

static int selftest(const float* A, const float* B, float* Draw, int M, int N, int K, int bmn, void* stream);

extern "C" int dlwp_selftest_umma(const float* A, const float* B, float* Draw, int M, int N, int K, void* stream) {
  return selftest(A, B, Draw, M, N, K, 0, stream);
}

extern "C" int dlwp_selftest_umma_bmn(const float* A, const float* Bkn, float* Draw, int M, int N, int K, int variant,
                                      void* stream) {
  if (variant != 1 && variant != 2) return fail(DLWP_ERR_INVALID, "dlwp_selftest_umma_bmn: variant must be 1 or 2");
  return selftest(A, Bkn, Draw, M, N, K, variant, stream);
}

static int selftest(const float* A, const float* B, float* Draw, int M, int N, int K, int bmn, void* stream) {
  if (!A || !B || !Draw) return fail(DLWP_ERR_INVALID, "dlwp_selftest_umma: null argument");
  DLWP_REQUIRE(M == 64 || M == 128, "M must be 64 or 128");
  DLWP_REQUIRE(N >= 8 && N <= 256 && N % 8 == 0 && (M == 64 || N % 16 == 0), "N must be a multiple of 8 (M = 64) or 16 (M = 128) in [8, 256]");
  DLWP_REQUIRE(K >= 16 && K <= 128 && K % 16 == 0, "K must be a multiple of 16 in [16, 128]");
  cudaStream_t st = (cudaStream_t)stream;
  const size_t smem = (size_t)(M + N) * K * 2;
  DLWP_CUDA(cudaFuncSetAttribute(umma_selftest_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  umma_selftest_kernel<<<1, 128, smem, st>>>(A, B, Draw, M, N, K, bmn);
  DLWP_LAUNCHED("umma_selftest_kernel");
  return DLWP_OK;
}
