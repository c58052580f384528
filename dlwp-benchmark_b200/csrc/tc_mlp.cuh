// Pieces shared by the tcgen05 MLP kernels (tc_fno.cu: lifting / projection; tc_afno.cu: the AFNO
// block's LayerNorm + Mlp): CTA organisation, epilogue thread coordinates, the issuer of the two
// chained GEMMs and the GELU epilogue that turns the first accumulator into the second operand.
#pragma once
#include "tc_common.cuh"
#include "tc_fno.cuh"
#include "tc_gelu.cuh"

namespace dlwp {
using namespace tc;

namespace {

constexpr int kImg = kTcImgBytes;            // 8192: one bf16 part of a [128 x 32] tile
constexpr int kTileBytes = 2 * kImg;         // hi + lo
constexpr int kFBytes = 2 * 32 * 128 * 2;     // forward W-DFT operand: [hi | lo] images of 32 rows x 128 pixels

// v = hi + lo for a pair, residual through one packed FMA
__device__ __forceinline__ void split2p(float2 v, uint32_t& hi, uint32_t& lo) {
  __nv_bfloat162 h = __floats2bfloat162_rn(v.x, v.y);
  hi = *reinterpret_cast<uint32_t*>(&h);
  const float2 hf = make_float2(__uint_as_float(hi << 16), __uint_as_float(hi & 0xffff0000u));
  const float2 d = __ffma2_rn(hf, make_float2(-1.0f, -1.0f), v);
  __nv_bfloat162 l = __floats2bfloat162_rn(d.x, d.y);
  lo = *reinterpret_cast<uint32_t*>(&l);
}

}  // namespace

namespace {

constexpr int kEpi = 256;                    // epilogue threads per worker
constexpr int kEpiWarps = kEpi / 32;
constexpr int kCtlWarp = 2 * kEpiWarps;      // first control warp
constexpr int kCta = 2 * kEpi + 2 * 32;      // 576
constexpr int kCtaLayer = kCta + 2 * 32;     // the layer kernel splits control into an issuer and a loader / finisher warp

// epilogue thread coordinates: tile row (= pixel = TMEM lane), column half, TMEM lane quarter
struct EpiId {
  int w, row, half, q4;
  __device__ __forceinline__ explicit EpiId(int tid) {
    const int warp = tid >> 5;
    w = warp / kEpiWarps;
    q4 = warp & 3;
    half = (warp >> 2) & 1;
    row = q4 * 32 + (tid & 31);
  }
};

__device__ __forceinline__ uint32_t full_par(uint32_t u) { return (u >> 1) & 1u; }
__device__ __forceinline__ uint32_t empty_par(uint32_t u) { return ((u >> 1) & 1u) ^ 1u; }
}  // namespace

#ifdef DLWP_TRACE
// clock64 stamps of the MLP epilogue of block 0, worker 0, warp 0 (one copy per translation unit): [chunk step < 64][8]
static __device__ unsigned long long* g_trace_mlp = nullptr;
#define TRE(nc, ev) do { if (g_trace_mlp && blockIdx.x == 0 && threadIdx.x == 0 && (nc) < 64) g_trace_mlp[(nc) * 8 + (ev)] = clock64(); } while (0)
#else
#define TRE(nc, ev) do {} while (0)
#endif

namespace {

// The hidden activations never touch shared memory: the epilogue writes GELU(hidden) back INTO the accumulator columns it
// has just read (each thread owns 16 fp32 columns per 32-column chunk and writes 8 columns of bf16 hi pairs + 8 of lo pairs
// in their place), and the second GEMM takes its A operand from tensor memory.  A2FULL[2 (pass & 1) + chunk] hands a chunk
// to the issuer; the accumulator buffer returns to the first GEMM (D1EMPTY) when the second GEMM has consumed both chunks.
enum { MB_A1FULL = 0, MB_A1EMPTY = 2, MB_D1FULL = 4, MB_D1EMPTY = 6, MB_A2FULL = 8, MB_A2EMPTY = 12, MB_D2FULL = 16,
       MB_D2EMPTY = 18, MB_XTFULL = 20, MB_D3FULL = 21 /* two */, MB_YFREE = 23,
       MB_INFULL = 24 /* two */, MB_INEMPTY = 26 /* two */, MB_PER = 28 };      // IN: fp32 rows bulk-copied for the epilogue

__device__ __forceinline__ void mlp_init_barriers(uint64_t* b, uint32_t a1_full_count) {
  for (int i = 0; i < 2; ++i) {
    mbar_init(&b[MB_A1FULL + i], a1_full_count);
    mbar_init(&b[MB_A1EMPTY + i], 1);
    mbar_init(&b[MB_D1FULL + i], 1);
    mbar_init(&b[MB_D1EMPTY + i], 1);
    mbar_init(&b[MB_A2FULL + i], kEpi);
    mbar_init(&b[MB_A2EMPTY + i], 1);
    mbar_init(&b[MB_A2FULL + 2 + i], kEpi);
    mbar_init(&b[MB_A2EMPTY + 2 + i], 1);
    mbar_init(&b[MB_D2FULL + i], 1);
    mbar_init(&b[MB_D2EMPTY + i], kEpi);
  }
  mbar_init(&b[MB_XTFULL], kEpi);
  mbar_init(&b[MB_D3FULL], 1);
  mbar_init(&b[MB_D3FULL + 1], 1);
  mbar_init(&b[MB_YFREE], 1);
  for (int i = 0; i < 2; ++i) {
    mbar_init(&b[MB_INFULL + i], 1);
    mbar_init(&b[MB_INEMPTY + i], kEpi);
  }
}

struct MlpIssuer {
  uint64_t* b;
  uint32_t aA1;        // two A1 buffers, `a1_stride` bytes apart; lo image at +128*K1*2
  uint32_t a1_stride;
  uint32_t aW1, aW2, tmem;
  int hid, P;
  uint32_t nc;         // chunks issued so far
  uint32_t aOnes, aW1x;   // optional bias term of the first GEMM: Ones[128,16] * W1x[hid,16]^T (0 = none)
  uint32_t leader;        // elect_one() of the (converged) control warp
  uint32_t a1_single;     // 1: one A1 buffer (ring of depth 1) instead of two

  // Accumulator buffer gp & 1 was last the A operand of the second GEMM of pass gp - 2.  Those UMMAs were issued by this same
  // thread before the ones that now overwrite the buffer, and tcgen05.mma instructions of one thread execute in issue order
  // (operand reads at the head of the pipeline, accumulator writes at its tail), so no completion round trip through
  // D1EMPTY is needed: it cost commit -> mbarrier -> wake-up per pass and left the epilogue group waiting for its next
  // accumulator (7 % of the stall samples of the lift and layer+projection kernels).  -DDLWP_D1EMPTY_WAIT restores the wait.
  __device__ __forceinline__ void d1_reuse_wait([[maybe_unused]] int db, [[maybe_unused]] uint32_t gp) {
#ifdef DLWP_D1EMPTY_WAIT
    mbar_wait(&b[MB_D1EMPTY + db], empty_par(gp));
#endif
  }

  template <int K1>
  __device__ __forceinline__ void mma1(uint32_t u, int p, uint32_t gp) {
    const int db = gp & 1, ub = a1_single ? 0 : (u & 1);
    d1_reuse_wait(db, gp);
    if (p == 0) mbar_wait(&b[MB_A1FULL + ub], a1_single ? (u & 1u) : full_par(u));
    tc_fence_after();
    const uint32_t idesc = idesc_bf16(128, 64);
    const uint32_t a_lo = 128 * K1 * 2, w_lo = (uint32_t)hid * K1 * 2;
    const uint64_t dA = chunk_desc(aA1 + ub * a1_stride, 128, 0),
                   dW = smem_desc(aW1 + (uint32_t)p * 64 * 16, (uint32_t)hid * 16, 128);
    uint32_t acc = 0;
#pragma unroll
    for (int term = 0; term < 3; ++term) {
      const uint32_t ao = term == 1 ? a_lo : 0u, wo = term == 2 ? w_lo : 0u;
#pragma unroll
      for (int k0 = 0; k0 < K1; k0 += 16) {
        umma_bf16_p(leader, tmem + 64u * db, desc_off(dA, ao + (k0 >> 3) * 2048), desc_off(dW, wo + (uint32_t)(k0 >> 3) * hid * 16),
                    idesc, acc);
        acc = 1;
      }
    }
    if (aW1x != 0)
      umma_bf16_p(leader, tmem + 64u * db, chunk_desc(aOnes, 128, 0),
                  smem_desc(aW1x + (uint32_t)p * 64 * 16, (uint32_t)hid * 16, 128), idesc, 1);
    umma_commit_p(leader, &b[MB_D1FULL + db]);
    if (p == P - 1) umma_commit_p(leader, &b[MB_A1EMPTY + ub]);
    __syncwarp();
  }

  // The same with the A operand in tensor memory (fno_layerproj_kernel: the layer's output tile never leaves the SM):
  // buffer u & 1 at tA1 + K1 * (u & 1), per K step of 16: [hi 8 columns | lo 8 columns].
  template <int K1>
  __device__ __forceinline__ void mma1_ts(uint32_t tA1, uint32_t u, int p, uint32_t gp) {
    const int db = gp & 1, ub = u & 1;
    d1_reuse_wait(db, gp);
    if (p == 0) mbar_wait(&b[MB_A1FULL + ub], full_par(u));
    tc_fence_after();
    const uint32_t idesc = idesc_bf16(128, 64);
    const uint32_t w_lo = (uint32_t)hid * K1 * 2;
    const uint64_t dW = smem_desc(aW1 + (uint32_t)p * 64 * 16, (uint32_t)hid * 16, 128);
    const uint32_t a1 = tA1 + (uint32_t)K1 * ub;
    uint32_t acc = 0;
#pragma unroll
    for (int term = 0; term < 3; ++term) {
      const uint32_t ao = term == 1 ? 8u : 0u, wo = term == 2 ? w_lo : 0u;
#pragma unroll
      for (int k0 = 0; k0 < K1; k0 += 16) {
        umma_ts_bf16_p(leader, tmem + 64u * db, a1 + (uint32_t)k0 + ao, desc_off(dW, wo + (uint32_t)(k0 >> 3) * hid * 16), idesc, acc);
        acc = 1;
      }
    }
    if (aW1x != 0)
      umma_bf16_p(leader, tmem + 64u * db, chunk_desc(aOnes, 128, 0),
                  smem_desc(aW1x + (uint32_t)p * 64 * 16, (uint32_t)hid * 16, 128), idesc, 1);
    umma_commit_p(leader, &b[MB_D1FULL + db]);
    if (p == P - 1) umma_commit_p(leader, &b[MB_A1EMPTY + ub]);
    __syncwarp();
  }

  template <int N2>
  __device__ __forceinline__ void mma2(uint32_t u, int p) {
    const uint32_t idesc = idesc_bf16(128, N2);
    const uint32_t w_lo = (uint32_t)N2 * hid * 2;
    const uint32_t d2 = tmem + 128u + (N2 > 32 ? 64u : 32u) * (u & 1);
    const uint32_t db = (u * (uint32_t)P + (uint32_t)p) & 1u;
#pragma unroll
    for (int j = 0; j < 2; ++j, ++nc) {
      const int ab = nc & 3;
      mbar_wait(&b[MB_A2FULL + ab], (nc >> 2) & 1u);
      if (p == 0 && j == 0) mbar_wait(&b[MB_D2EMPTY + (u & 1)], empty_par(u));
      tc_fence_after();
      const int hcol0 = p * 64 + 32 * j;
      const uint64_t dW = chunk_desc(aW2, N2, hcol0);
      const uint32_t a2 = tmem + 64u * db + 32u * j;      // [K step 0: hi 8 columns | lo 8][K step 1: hi | lo]
#pragma unroll
      for (int term = 0; term < 3; ++term) {      // hi*hi, lo*hi, hi*lo
        const uint32_t ao = term == 1 ? 8u : 0u, wo = term == 2 ? w_lo : 0u;
#pragma unroll
        for (int kk = 0; kk < 2; ++kk)
          umma_ts_bf16_p(leader, d2, a2 + 16u * kk + ao, desc_off(dW, wo + (uint32_t)kk * 2 * N2 * 16), idesc, (p | j | term | kk) != 0);
      }
      __syncwarp();
    }
#ifdef DLWP_D1EMPTY_WAIT
    umma_commit_p(leader, &b[MB_D1EMPTY + db]);      // the accumulator buffer (now the consumed A operand) goes back to the first GEMM
#endif
    if (p == P - 1) umma_commit_p(leader, &b[MB_D2FULL + (u & 1)]);
  }
};

// epilogue side of one tile's MLP; `after_pass(p)` runs after every pass p.  Each thread turns its 16 hidden columns of its
// pixel per 32-column chunk into split-bf16 operand pairs, in place in tensor memory.
// `bias(p, j, r)` (optional) adds the first layer's bias to the 16 accumulator values of chunk j of pass p, when the
// first GEMM has no spare K column to carry it.
struct NoBias {
  __device__ __forceinline__ void operator()(int, int, uint32_t (&)[16]) const {}
};
template <class Hook, class Bias = NoBias>
__device__ __forceinline__ void epi_mlp_tile(uint64_t* b, const EpiId& id, uint32_t tmem_lane, int P, uint32_t& gp, uint32_t& nc,
                                             Hook after_pass, Bias bias = Bias{}) {
  for (int p = 0; p < P; ++p, ++gp) {
    const int db = gp & 1;
    TRE(nc, 0);
    mbar_wait(&b[MB_D1FULL + db], full_par(gp));
    TRE(nc, 1);
    tc_fence_after();
    // both 16-column chunks of the pass with one tensor-memory round trip
    uint32_t rr[2][16];
    TRE(nc, 2);
    tmem_ld16(tmem_lane + 64u * db, rr[0]);
    tmem_ld16(tmem_lane + 64u * db + 32u, rr[1]);
    tmem_ld_wait();
    TRE(nc, 3);
#pragma unroll
    for (int j = 0; j < 2; ++j, ++nc) {
      uint32_t (&r)[16] = rr[j];
      uint32_t h[8], l[8];
      {
        bias(p, j, r);
        float2 g2[8];             // 2 * gelu: W2's operand image is packed with the factor 1/2
        gelu2x_8(r, g2);
#pragma unroll
        for (int e = 0; e < 8; ++e) split2p(g2[e], h[e], l[e]);
      }
      TRE(nc, 4);
      tmem_st8(tmem_lane + 64u * db + 32u * j, h);
      tmem_st8(tmem_lane + 64u * db + 32u * j + 8u, l);
      if (j == 1) {      // one wait for the stores of both chunks: chunk 0's store latency hides behind chunk 1's arithmetic
        tmem_st_wait();
        TRE(nc, 6);
        tc_fence_before();
        mbar_arrive(&b[MB_A2FULL + ((nc - 1) & 3)]);
        mbar_arrive(&b[MB_A2FULL + (nc & 3)]);
        TRE(nc, 7);
      }
    }
    after_pass(p);
  }
}

}  // namespace

}  // namespace dlwp
