// mix_few_tc_kernel: the per-mode channel mixing of the bf16 kernel set (hidden width 128) for a FEW samples per GPU on tcgen05.
//   Y[m, b, o] = sum_i X[m, b, i] W[m, i, o]   (complex; einsum("bixy,ioxy->boxy"), unet.py:15-17 / neuralop SpectralConv)
// With 1-8 samples the product is bound by reading every mode's 128 x 128 complex weights once (64 KB per mode as bf16,
// 138 MB per layer at 64 x 33 modes), so the weights are the operand that streams: per mode one UMMA chain
//   D[o, (b, re | im)] = A[o, k] * Bop[(b, re | im), k],   M = 128 output channels, N = 16, K = 256 = (re | im part, i)
//   A[o, k < 128] = Re W[i = k, o],  A[o, k >= 128] = Im W[i = k - 128, o]            (packed once per weight commit)
//   Bop[(b, re), k] = (Re x_i | -Im x_i),  Bop[(b, im), k] = (Im x_i | Re x_i)       (built per mode from the spectrum, 8 KB)
// i.e. D[o, (b, re)] = Re y, D[o, (b, im)] = Im y.  Plain bf16 operands, fp32 accumulation (the precision of DLWP_PREC_BF16).
// A persistent CTA per SM takes modes blockIdx.x + k gridDim.x: a producer lane keeps three 64 KB weight images in flight by
// bulk async copies, four warps build the B operand, one lane issues the 16 UMMAs of a mode, four warps take the accumulator
// (lane = output channel) to the bf16 tile images the inverse H-DFT reads (layout: kernels_fp32.cu mix_store_image).
// The CUDA-core forms (mix_dense_fewb16_kernel / mix_dense_kernel) stay for other widths and larger batches.
#include "tc_common.cuh"
#include "tc_wide.cuh"

namespace dlwp {

using namespace tc;

namespace {

constexpr int kMixC = 128;                         // channels in and out
constexpr int kMixWBytes = 2 * kMixC * kMixC * 2;  // one mode's operand image: 128 rows x K = 256 bf16 = 64 KB
constexpr int kMixStages = 3;
constexpr int kMixBBytes = 16 * 256 * 2;           // B operand: 16 rows x K = 256
constexpr int kMixThreads = 10 * 32;               // 4 builder warps, 4 epilogue warps, producer, issuer
enum { XB_WFULL = 0, XB_WEMPTY = 3, XB_BFULL = 6, XB_BEMPTY = 8, XB_DFULL = 10, XB_DEMPTY = 12, XB_N = 14 };
struct MixSmem {
  static constexpr int W = 0, B = W + kMixStages * kMixWBytes, Total = B + 2 * kMixBBytes;
};

__device__ __forceinline__ uint16_t bf16_bits(float v) {
  const __nv_bfloat16 h = __float2bfloat16(v);
  return *reinterpret_cast<const uint16_t*>(&h);
}

// dense mode-major complex weights [modes][i][o] -> chunk-major operand images [modes][k / 8][o][k % 8] bf16
__global__ void mix_few_pack_kernel(const float2* __restrict__ wm, uint8_t* __restrict__ img, long long units) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;      // one 16-byte unit: (mode, K chunk, o)
  if (t >= units) return;
  const int o = (int)(t % kMixC), c = (int)((t / kMixC) % 32);
  const long long m = t / (kMixC * 32);
  uint32_t w[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    uint32_t pr[2];
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      const int k = 8 * c + 2 * j + e, i = k & (kMixC - 1);
      const float2 v = wm[(m * kMixC + i) * kMixC + o];
      pr[e] = bf16_bits(k < kMixC ? v.x : v.y);
    }
    w[j] = pr[0] | (pr[1] << 16);
  }
  *reinterpret_cast<uint4*>(img + t * 16) = make_uint4(w[0], w[1], w[2], w[3]);
}

struct MixFewArgs {
  const float2* X;
  const uint8_t* Wimg;
  uint8_t* ys;
  int modes, B, Kc;
};

__global__ void __launch_bounds__(kMixThreads, 1) mix_few_tc_kernel(const MixFewArgs a) {
  extern __shared__ __align__(128) uint8_t smem[];
  const float2* __restrict__ X = a.X;
  const uint8_t* __restrict__ Wimg = a.Wimg;
  uint8_t* __restrict__ ys = a.ys;
  const int modes = a.modes, B = a.B, Kc = a.Kc;
  pdl_launch_dependents();
  __shared__ uint64_t bars[XB_N];
  __shared__ uint32_t tmem_s;
  const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;
  if (tid == 0) {
    for (int i = 0; i < kMixStages; ++i) {
      mbar_init(&bars[XB_WFULL + i], 1);
      mbar_init(&bars[XB_WEMPTY + i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&bars[XB_BFULL + i], 128);
      mbar_init(&bars[XB_BEMPTY + i], 1);
      mbar_init(&bars[XB_DFULL + i], 1);
      mbar_init(&bars[XB_DEMPTY + i], 128);
    }
    fence_mbar_init();
  }
  if (warp == 9) {
    tmem_alloc(&tmem_s, 32);
    tmem_relinquish();
  }
  // columns of the B operand no sample uses stay zero (both buffers)
  for (int i = tid; i < 2 * kMixBBytes / 16; i += kMixThreads) reinterpret_cast<uint4*>(smem + MixSmem::B)[i] = make_uint4(0u, 0u, 0u, 0u);
  fence_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const int first = blockIdx.x, stride = gridDim.x;

  if (warp == 8) {
    // ---------------- producer lane: the modes' weight images (constant since the weight commit: no dependency wait) ----------------
    const uint32_t leader = elect_one();
    uint32_t k = 0;
    for (int m = first; m < modes; m += stride, ++k) {
      const uint32_t s = k % kMixStages;
      mbar_wait(&bars[XB_WEMPTY + s], ((k / kMixStages) & 1u) ^ 1u);
      mbar_arrive_expect_tx_p(leader, &bars[XB_WFULL + s], kMixWBytes);
      bulk_g2s_p(leader, smem + MixSmem::W + s * kMixWBytes, Wimg + (size_t)m * kMixWBytes, kMixWBytes, &bars[XB_WFULL + s]);
    }
  } else if (warp == 9) {
    // ---------------- issuer lane ----------------
    const uint32_t leader = elect_one();
    const uint32_t idesc = idesc_bf16(128, 16);
    uint32_t k = 0;
    for (int m = first; m < modes; m += stride, ++k) {
      const uint32_t s = k % kMixStages, db = k & 1u;
      mbar_wait(&bars[XB_WFULL + s], (k / kMixStages) & 1u);
      mbar_wait(&bars[XB_BFULL + db], (k >> 1) & 1u);
      mbar_wait(&bars[XB_DEMPTY + db], ((k >> 1) & 1u) ^ 1u);
      tc_fence_after();
      const uint64_t dA = chunk_desc(smem_u32(smem + MixSmem::W + s * kMixWBytes), 128, 0);
      const uint64_t dB = chunk_desc(smem_u32(smem + MixSmem::B + db * kMixBBytes), 16, 0);
#pragma unroll
      for (int ks = 0; ks < 16; ++ks)      // K steps of 16: two chunks of 8 per operand
        umma_bf16_p(leader, tmem_s + 16u * db, desc_off(dA, ks * 2 * 128 * 16), desc_off(dB, ks * 2 * 16 * 16), idesc, ks != 0);
      umma_commit_p(leader, &bars[XB_WEMPTY + s]);
      umma_commit_p(leader, &bars[XB_BEMPTY + db]);
      umma_commit_p(leader, &bars[XB_DFULL + db]);
      __syncwarp();
    }
  } else if (warp < 4) {
    // ---------------- builder warps: thread = input channel i; the mode's spectrum rows -> B operand ----------------
    pdl_wait();      // the spectrum comes from the forward H-DFT kernel
    const int i = tid;
    uint32_t k = 0;
    for (int m = first; m < modes; m += stride, ++k) {
      const uint32_t db = k & 1u;
      float2 x[8];
#pragma unroll
      for (int b = 0; b < 8; ++b) x[b] = b < B ? X[((long long)m * B + b) * kMixC + i] : make_float2(0.f, 0.f);
      mbar_wait(&bars[XB_BEMPTY + db], ((k >> 1) & 1u) ^ 1u);
      // element (row n, k) of the chunk-major operand: ((k / 8) * 16 + n) * 16 + (k % 8) * 2 bytes
      uint8_t* bo = smem + MixSmem::B + db * kMixBBytes + (i >> 3) * 256 + (i & 7) * 2;
#pragma unroll
      for (int b = 0; b < 8; ++b)
        if (b < B) {
          const uint16_t xr = bf16_bits(x[b].x), xi = bf16_bits(x[b].y);
          *reinterpret_cast<uint16_t*>(bo + (2 * b) * 16) = xr;                     // (b, re): k = i
          *reinterpret_cast<uint16_t*>(bo + (2 * b) * 16 + 16 * 256) = xi ^ 0x8000u; //          k = 128 + i: -Im x
          *reinterpret_cast<uint16_t*>(bo + (2 * b + 1) * 16) = xi;                 // (b, im)
          *reinterpret_cast<uint16_t*>(bo + (2 * b + 1) * 16 + 16 * 256) = xr;
        }
      fence_async_smem();
      mbar_arrive(&bars[XB_BFULL + db]);
    }
  } else {
    // ---------------- epilogue warps: thread = output channel o (TMEM lane) ----------------
    pdl_wait();      // the tile images are read by the kernel before the spectrum's producer at the earliest
    const int q4 = warp & 3, o = q4 * 32 + lane;
    uint32_t k = 0;
    for (int m = first; m < modes; m += stride, ++k) {
      const uint32_t db = k & 1u;
      mbar_wait(&bars[XB_DFULL + db], (k >> 1) & 1u);
      tc_fence_after();
      uint32_t r[16];
      tmem_ld16(tmem_s + ((uint32_t)(q4 * 32) << 16) + 16u * db, r);
      tmem_ld_wait();
      tc_fence_before();
      mbar_arrive(&bars[XB_DEMPTY + db]);
      const int row = m / Kc, kx = m - row * Kc;
#pragma unroll
      for (int b = 0; b < 8; ++b)
        if (b < B) {      // per (sample, retained column) a [128 rows = (retained row, re | im)] x [128 channels] chunk-major tile
          uint8_t* p = ys + ((size_t)b * Kc + kx) * 32768 + (size_t)(o >> 3) * 2048 + (size_t)(2 * row) * 16 + (o & 7) * 2;
          *reinterpret_cast<__nv_bfloat16*>(p) = __float2bfloat16(__uint_as_float(r[2 * b]));
          *reinterpret_cast<__nv_bfloat16*>(p + 16) = __float2bfloat16(__uint_as_float(r[2 * b + 1]));
        }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 9) tmem_dealloc(tmem_s, 32);
}

}  // namespace

bool mix_few_tc_covers(int modes, int B, int C, int Kc) { return C == kMixC && B >= 1 && B <= 8 && Kc >= 1 && modes % Kc == 0 && modes / Kc <= 64; }

int launch_mix_few_pack(const float2* wm, uint8_t* Wimg, int modes, cudaStream_t st) {
  const long long units = (long long)modes * 32 * kMixC;
  mix_few_pack_kernel<<<(unsigned)((units + 255) / 256), 256, 0, st>>>(wm, Wimg, units);
  DLWP_LAUNCHED("mix_few_pack_kernel");
  return DLWP_OK;
}

int launch_mix_few_tc(const float2* X, const uint8_t* Wimg, uint8_t* ys, int modes, int B, int Kc, cudaStream_t st) {
  DLWP_REQUIRE(mix_few_tc_covers(modes, B, kMixC, Kc), "mix_few_tc_kernel: %d modes, %d samples, %d columns not covered", modes, B, Kc);
  int sms = 0;
  DLWP_TRY(current_sms(&sms));
  DLWP_TRY(ensure_dynamic_smem((const void*)mix_few_tc_kernel, MixSmem::Total));
  const int grid = modes < sms ? modes : sms;
  const MixFewArgs a{X, Wimg, ys, modes, B, Kc};
  DLWP_CUDA(launch_pdl(mix_few_tc_kernel, dim3(grid), dim3(kMixThreads), (size_t)MixSmem::Total, st, a));
  DLWP_LAUNCHED("mix_few_tc_kernel");
  return DLWP_OK;
}

}  // namespace dlwp
