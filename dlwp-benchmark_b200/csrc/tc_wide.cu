// tcgen05 kernels of the FNO step at hidden width 128 in bf16 (see tc_wide.cuh).
//
// tcw_kernel: one persistent, warp-specialised kernel body, instantiated for the DFT GEMMs and the
// layer GEMM of the step (the fused lifting / projection MLPs are tcw_mlp_kernel below):
//   warps 0-7 : epilogue group (TMEM -> registers -> GELU / convert -> shared-memory tile image
//               -> one bulk store; or 16-byte chunk rows straight into the next GEMM's operand images)
//   warp 8    : loader  -- bulk async copies of the streamed operands into a ring of stages
//   warp 9    : issuer  -- tcgen05.mma into one of two TMEM accumulators (256 columns each)
// The three sides meet only on mbarriers (full / empty per stage, acc_full / acc_empty per
// accumulator).  An item is one 128-slot tile (one grid row for TCW_FWD, one (sample, retained
// column) for TCW_HFWD, one (sample, retained column, block of 64 grid rows) for TCW_HINV) and
// consists of one or more K blocks, each one stage of the ring.  Weights stay resident in shared memory; biases are
// added by one extra K = 16 MMA of a ones operand with a [N x 16] bias image (hi, lo columns).
#include "tc_wide.cuh"

#include <cstring>
#include <vector>

#include <cuda_bf16.h>

#include "tc_common.cuh"
#include "tc_gelu.cuh"

namespace dlwp {

using namespace tc;

namespace {

constexpr int kWEpi = 256;                 // epilogue threads
constexpr int kWLoader = kWEpi / 32;       // warp 8
constexpr int kWIssuer = kWLoader + 1;     // warp 9
constexpr int kWCta = kWEpi + 64;          // 320
constexpr int kXTile = 128 * kWideC * 2;   // [128 x 128] bf16 image: 32 KB
constexpr int kHidTile = 128 * kWideHid * 2;   // [128 x 256]: 64 KB
constexpr int kOnes = 128 * 16 * 2;        // ones operand [128 x 16]
constexpr int kMaxStages = 4;

// per-GEMM constants
template <int MODE> struct WT;
template <> struct WT<TCW_FWD> {
  static constexpr int N = 128, NS = 3, RES = 0, BIAS = 0, OUT = 0, NOUT = 1;
  __host__ __device__ static int stage(int) { return 2 * kXTile; }   // forward DFT tile + activation tile
};
template <> struct WT<TCW_HFWD> : WT<TCW_FWD> {};   // forward H-DFT: DFT tile + spectrum image per 64 grid rows
template <> struct WT<TCW_HINV> : WT<TCW_FWD> {};   // inverse H-DFT: DFT tile + mixed-spectrum image
template <> struct WT<TCW_LAYER> {
  static constexpr int N = 128, NS = 2, RES = 128 * 128 * 2, BIAS = 128 * 32, OUT = kXTile, NOUT = 1;
  __host__ __device__ static int stage(int KZ) { return kXTile + 2 * KZ * 256; }   // x, G tile, Z row image
};
template <int MODE>
__host__ __device__ inline int tcw_smem_bytes(int KZ) {
  using T = WT<MODE>;
  return T::RES + T::BIAS + (T::BIAS ? kOnes : 0) + T::NS * T::stage(KZ) + T::NOUT * T::OUT;
}

__device__ __forceinline__ void bulk_wait_read1() { asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory"); }

__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}

// 16 accumulator columns -> two 16-byte chunk rows of the bf16 tile image
template <bool GELU>
__device__ __forceinline__ void cols16_to_image(const uint32_t (&r)[16], uint8_t* img, int row, int c0) {
  uint32_t o[8];
  if (GELU) {
    float2 g[8];
    gelu2x_8(r, g);
#pragma unroll
    for (int e = 0; e < 8; ++e) o[e] = pack_bf16(g[e].x, g[e].y);
  } else {
#pragma unroll
    for (int e = 0; e < 8; ++e) o[e] = pack_bf16(__uint_as_float(r[2 * e]), __uint_as_float(r[2 * e + 1]));
  }
  uint8_t* d = img + (c0 >> 3) * 2048 + row * 16;
  *reinterpret_cast<uint4*>(d) = make_uint4(o[0], o[1], o[2], o[3]);
  *reinterpret_cast<uint4*>(d + 2048) = make_uint4(o[4], o[5], o[6], o[7]);
}

}  // namespace

namespace {

template <int MODE>
__global__ void __launch_bounds__(kWCta, 1) tcw_kernel(const TcwArgs a) {
  using T = WT<MODE>;
  // the three DFT GEMMs contract over the rows of a tile image read in place as an MN-major B operand
  constexpr bool kDft = MODE == TCW_FWD || MODE == TCW_HFWD || MODE == TCW_HINV;
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bars[2 * kMaxStages + 5];
  __shared__ uint32_t tmem_slot;
  pdl_launch_dependents();
  uint64_t* full = bars;
  uint64_t* empty = bars + kMaxStages;
  uint64_t* acc_full = bars + 2 * kMaxStages;
  uint64_t* acc_empty = acc_full + 2;
  uint64_t* res_bar = acc_empty + 2;

  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  const int stage_bytes = T::stage(a.KZ);
  const int nblk = MODE == TCW_FWD ? a.TW : (MODE == TCW_HFWD ? a.HB : 1);
  uint8_t* sRes = smem;
  uint8_t* sBias = sRes + T::RES;
  uint8_t* sOnes = sBias + T::BIAS;
  uint8_t* sStage = sOnes + (T::BIAS ? kOnes : 0);
  uint8_t* sOut = sStage + T::NS * stage_bytes;

  if (tid == 0) {
    for (int s = 0; s < T::NS; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&acc_full[i], 1);
      mbar_init(&acc_empty[i], kWEpi);
    }
    mbar_init(res_bar, 1);
    fence_mbar_init();
  }
  if (T::BIAS && tid < 128) {   // ones operand: columns 0 and 1 (hi and lo parts of the bias) are 1.0
    *reinterpret_cast<uint4*>(sOnes + tid * 16) = make_uint4(0x3F803F80u, 0u, 0u, 0u);
    *reinterpret_cast<uint4*>(sOnes + 2048 + tid * 16) = make_uint4(0u, 0u, 0u, 0u);
    fence_async_smem();
  }
  if (warp == kWIssuer) {
    tmem_alloc(&tmem_slot, 512);
    tmem_relinquish();
    tc_fence_before();
  }
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;

  if (warp == kWLoader) {
    // ------------------------------ loader ------------------------------
    const uint32_t leader = elect_one();
    if (T::RES + T::BIAS > 0) {   // constants: may be fetched before the previous kernel has finished
      mbar_arrive_expect_tx_p(leader, res_bar, (uint32_t)(T::RES + T::BIAS));
      if (T::RES) bulk_g2s_p(leader, sRes, a.b_res, T::RES, res_bar);
      if (T::BIAS) bulk_g2s_p(leader, sBias, a.bias_img, T::BIAS, res_bar);
    }
    pdl_wait();
    uint32_t u = 0;
    for (int item = blockIdx.x; item < a.nitems; item += gridDim.x) {
      for (int blk = 0; blk < nblk; ++blk, ++u) {
        const uint32_t s = u % T::NS;
        uint8_t* st = sStage + s * stage_bytes;
        mbar_wait(&empty[s], ((u / T::NS) & 1u) ^ 1u);
        mbar_arrive_expect_tx_p(leader, &full[s], (uint32_t)stage_bytes);
        if (kDft) {
          // DFT tile: one per tile of the grid row / per block of 64 grid rows; image: this item's block
          const size_t gi = MODE == TCW_HINV ? (size_t)(item % a.HB) : (size_t)blk;
          const size_t ai = MODE == TCW_HINV ? (size_t)(item / a.HB) : (size_t)item * nblk + blk;
          bulk_g2s_p(leader, st, a.g_img + gi * kXTile, kXTile, &full[s]);
          bulk_g2s_p(leader, st + kXTile, a.a_img + ai * kXTile, kXTile, &full[s]);
        } else {   // TCW_LAYER
          const uint32_t zb = (uint32_t)a.KZ * 256u;
          bulk_g2s_p(leader, st, a.a_img + (size_t)item * kXTile, kXTile, &full[s]);
          bulk_g2s_p(leader, st + kXTile, a.g_img + (size_t)(item % a.TW) * zb, zb, &full[s]);
          bulk_g2s_p(leader, st + kXTile + zb, a.z_img + (size_t)(item / a.TW) * zb, zb, &full[s]);
        }
        __syncwarp();
      }
    }
  } else if (warp == kWIssuer) {
    // ------------------------------ issuer ------------------------------
    const uint32_t leader = elect_one();
    pdl_wait();
    if (T::RES + T::BIAS > 0) mbar_wait(res_bar, 0);
    const uint32_t idesc = idesc_bf16(128, T::N) | (kDft ? kIdescBMnMajor : 0u);
    const uint64_t dRes = chunk_desc(smem_u32(sRes), T::N, 0);
    const uint32_t bstep = 2u * T::N * 16u;   // resident operand: bytes per K = 16 step
    uint32_t u = 0, it = 0;
    for (int item = blockIdx.x; item < a.nitems; item += gridDim.x, ++it) {
      const uint32_t acc = it & 1u;
      mbar_wait(&acc_empty[acc], ((it >> 1) & 1u) ^ 1u);
      tc_fence_after();
      const uint32_t d = tmem + acc * 256u;
      for (int blk = 0; blk < nblk; ++blk, ++u) {
        const uint32_t s = u % T::NS;
        const uint32_t st = smem_u32(sStage + s * stage_bytes);
        mbar_wait(&full[s], (u / T::NS) & 1u);
        tc_fence_after();
        const uint64_t dA = chunk_desc(st, 128, 0);
        if (kDft) {
          // B: the activation image read in place as an MN-major operand (pixels = K, 8 x 8 core
          // matrices 128 B apart along pixels, 2048 B along channels)
          const uint64_t dX = smem_desc(st + kXTile, 128, 2048);
#pragma unroll
          for (int j = 0; j < 8; ++j)
            umma_bf16_p(leader, d, desc_off(dA, j * 4096), desc_off(dX, j * 256), idesc, (blk | j) != 0);
        } else {   // TCW_LAYER: skip conv, then the inverse W-DFT, into one accumulator
#pragma unroll
          for (int j = 0; j < 8; ++j) umma_bf16_p(leader, d, desc_off(dA, j * 4096), desc_off(dRes, j * bstep), idesc, j != 0);
          const uint32_t zb = (uint32_t)a.KZ * 256u;
          // Z of this grid row: [KZ rows = (retained column, re | im)] x [128 channels], MN-major (as
          // the inverse H-DFT epilogue stores it): core matrices 128 B apart along K, KZ * 16 B along channels
          const uint64_t dG = chunk_desc(st + kXTile, 128, 0), dZ = smem_desc(st + kXTile + zb, 128, (uint32_t)a.KZ * 16u);
          for (int j = 0; j < (a.KZ >> 4); ++j)
            umma_bf16_p(leader, d, desc_off(dG, j * 4096), desc_off(dZ, j * 256), idesc | kIdescBMnMajor, 1u);
        }
        umma_commit_p(leader, &empty[s]);
      }
      if (T::BIAS)
        umma_bf16_p(leader, d, chunk_desc(smem_u32(sOnes), 128, 0), chunk_desc(smem_u32(sBias), T::N, 0), idesc, 1u);
      umma_commit_p(leader, &acc_full[acc]);
      __syncwarp();
    }
  } else {
    // ------------------------------ epilogue group ------------------------------
    pdl_wait();
    const int q4 = warp & 3, half = warp >> 2;
    const int row = q4 * 32 + lane;                  // accumulator row = TMEM lane
    const uint32_t trow = tmem + ((uint32_t)(q4 * 32) << 16);
    uint32_t it = 0;
    for (int item = blockIdx.x; item < a.nitems; item += gridDim.x, ++it) {
      const uint32_t acc = it & 1u;
      const uint32_t d = trow + acc * 256u;
      uint8_t* ob = sOut + (T::NOUT > 1 ? (it & 1u) * T::OUT : 0u);
      if (T::OUT) {   // the bulk store that last read this buffer must have drained it
        if (tid == 0) {
          if (T::NOUT > 1) bulk_wait_read1();
          else bulk_wait_read0();
        }
        bar_sync(1, kWEpi);
      }
      mbar_wait(&acc_full[acc], (it >> 1) & 1u);
      tc_fence_after();
      if (MODE == TCW_LAYER) {
        const bool gelu = a.gelu != 0;
#pragma unroll 1
        for (int i = 0; i < T::N / 32; ++i) {
          const int c0 = i * 32 + half * 16;
          uint32_t r[16];
          tmem_ld16(d + (uint32_t)c0, r);
          tmem_ld_wait();
          if (gelu) cols16_to_image<true>(r, ob, row, c0);
          else cols16_to_image<false>(r, ob, row, c0);
        }
        tc_fence_before();
        mbar_arrive(&acc_empty[acc]);
        fence_async_smem();
        bar_sync(2, kWEpi);
        if (tid == 0) {
          bulk_s2g(a.out_img + (size_t)item * T::OUT, ob, T::OUT);
          bulk_commit();
        }
      } else if (MODE == TCW_FWD || MODE == TCW_HINV) {
        // bf16 operand images for the next DFT GEMM.  Accumulator row = (index, re | im), columns =
        // channels; every lane stores 16-byte chunk rows (8 channels), re / im lanes side by side.
        //   TCW_FWD : row k = 2 kx + (re | im) of grid row `item`  -> spectrum image (sample, kx, h / 64),
        //             image row 2 (h % 64) + (re | im)
        //   TCW_HINV: row 2 (h % 64) + (re | im) of (sample, kx, h / 64) -> Z image of grid row h,
        //             image row k = 2 kx + (re | im)   (MN-major operand of TCW_LAYER, KZ rows)
        bool valid;
        uint8_t* dst;
        uint32_t cstride;   // bytes between 8-channel chunks
        if (MODE == TCW_FWD) {
          const int b = item / a.H, h = item - b * a.H, kx = row >> 1;
          valid = row < 2 * a.Kc;
          dst = a.out_img + (((size_t)b * a.Kc + kx) * a.HB + (h >> 6)) * kXTile + (size_t)(((h & 63) << 1) | (row & 1)) * 16;
          cstride = 2048;
        } else {
          const int hb = item % a.HB, bk = item / a.HB;
          const int b = bk / a.Kc, kx = bk - b * a.Kc, h = hb * 64 + (row >> 1);
          valid = h < a.H;
          dst = a.out_img + ((size_t)b * a.H + h) * ((size_t)a.KZ * 256) + (size_t)((kx << 1) | (row & 1)) * 16;
          cstride = (uint32_t)a.KZ * 16u;
        }
        if (MODE == TCW_HINV || q4 * 32 < 2 * a.Kc) {
#pragma unroll 1
          for (int i = 0; i < 4; ++i) {
            const int c0 = half * 64 + i * 16;
            uint32_t r[16];
            tmem_ld16(d + (uint32_t)c0, r);
            tmem_ld_wait();
            if (valid) {
              uint32_t o[8];
#pragma unroll
              for (int e = 0; e < 8; ++e)
                o[e] = pack_bf16(__uint_as_float(r[2 * e]) * a.scale, __uint_as_float(r[2 * e + 1]) * a.scale);
              uint8_t* p = dst + (size_t)(c0 >> 3) * cstride;
              *reinterpret_cast<uint4*>(p) = make_uint4(o[0], o[1], o[2], o[3]);
              *reinterpret_cast<uint4*>(p + cstride) = make_uint4(o[4], o[5], o[6], o[7]);
            }
          }
        }
        tc_fence_before();
        mbar_arrive(&acc_empty[acc]);
      } else if (MODE == TCW_HFWD) {
        // accumulator row 2 r + (re | im) of (sample, kx), columns = channels -> X [R, Kc, B, 128] complex fp32
        // (what mix_dense reads).  Lane pairs swap halves of each 16-column block so that every lane
        // stores whole complex values: the even lane keeps channels c0..c0+7, the odd lane c0+8..c0+15.
        if (q4 * 32 < 2 * a.R) {
          const int b = item / a.Kc, kx = item - b * a.Kc;
          const int rr = row >> 1, odd = lane & 1;
          const bool valid = row < 2 * a.R;
          float2* dst = a.spec + (((size_t)rr * a.Kc + kx) * a.B + b) * kWideC;
#pragma unroll 1
          for (int i = 0; i < 4; ++i) {
            const int c0 = half * 64 + i * 16;
            uint32_t r[16];
            tmem_ld16(d + (uint32_t)c0, r);
            tmem_ld_wait();
            float2 o[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const float mine = __uint_as_float(odd ? r[8 + j] : r[j]);
              const float send = __uint_as_float(odd ? r[j] : r[8 + j]);
              const float got = __shfl_xor_sync(0xffffffffu, send, 1);
              o[j] = odd ? make_float2(got, mine) : make_float2(mine, got);
            }
            if (valid) {
              float4* p = reinterpret_cast<float4*>(dst + c0 + odd * 8);
#pragma unroll
              for (int j = 0; j < 4; ++j) p[j] = make_float4(o[2 * j].x, o[2 * j].y, o[2 * j + 1].x, o[2 * j + 1].y);
            }
          }
        }
        tc_fence_before();
        mbar_arrive(&acc_empty[acc]);
      }
    }
    if (T::OUT && tid == 0) bulk_wait0();
  }
  __syncthreads();
  if (warp == kWIssuer) tmem_dealloc(tmem, 512);
}

// ---- fused pointwise MLP: out = W2 gelu(W1 x + b1) + b2 with the 256-wide hidden kept on chip ------
//   KIND 0 (lifting):    x = [128 x 32] input images,       out = [128 x 128] activation images
//   KIND 1 (projection): x = [128 x 128] activation images, out = NCHW fp32 frame (+ residual frame)
// Roles: warps 0-15 first epilogue (accumulator 1 -> 2 gelu -> bf16 hidden tile in shared memory, handed
// to the issuer in four 64-column chunks so that the second GEMM runs under the GELU of the later
// chunks), warps 16-19 second epilogue (accumulator 2 -> output), warp 20 loader, warp 21 issuer.
// Accumulator 2 lives in the first columns of accumulator 1's buffer: those columns have been read
// by the time the chunks that cover them were handed over.  Two buffers, so that the first GEMM of
// the next tile runs while this tile is in the epilogues.
constexpr int kMlpEpi1Warps = 16;                       // four warps per TMEM lane quarter, 16 of every 64 columns each
constexpr int kMlpEpi1 = kMlpEpi1Warps * 32;
constexpr int kMlpEpi2 = 128;
constexpr int kMlpLoader = kMlpEpi1Warps + 4, kMlpIssuer = kMlpLoader + 1;
constexpr int kMlpCta = (kMlpIssuer + 1) * 32;

template <int KIND> struct MT;
template <> struct MT<0> {
  static constexpr int K1 = 32, N2 = 128, NS = 4, W1B = 256 * 32 * 2, B1B = 0, W2B = 128 * 256 * 2, B2B = 128 * 32, OUT = kXTile;
};
template <> struct MT<1> {
  static constexpr int K1 = 128, N2 = 16, NS = 2, W1B = 256 * 128 * 2, B1B = 256 * 32, W2B = 16 * 256 * 2, B2B = 16 * 32, OUT = 0;
};
template <int KIND>
constexpr int mlp_smem_bytes() {
  using T = MT<KIND>;
  return T::W1B + T::B1B + T::W2B + T::B2B + kOnes + T::NS * 128 * T::K1 * 2 + kHidTile + T::OUT;
}

template <int KIND>
__global__ void __launch_bounds__(kMlpCta, 1) tcw_mlp_kernel(const TcwArgs a) {
  using T = MT<KIND>;
  constexpr int kStage = 128 * T::K1 * 2;
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bars[2 * kMaxStages + 2 + 2 + 4 + 4 + 2 + 1];
  __shared__ uint32_t tmem_slot;
  pdl_launch_dependents();
  uint64_t* full = bars;
  uint64_t* empty = full + kMaxStages;
  uint64_t* d1_full = empty + kMaxStages;
  uint64_t* acc_empty = d1_full + 2;
  uint64_t* h_full = acc_empty + 2;
  uint64_t* h_empty = h_full + 4;
  uint64_t* d2_full = h_empty + 4;
  uint64_t* res_bar = d2_full + 2;

  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  uint8_t* sW1 = smem;
  uint8_t* sB1 = sW1 + T::W1B;
  uint8_t* sW2 = sB1 + T::B1B;
  uint8_t* sB2 = sW2 + T::W2B;
  uint8_t* sOnes = sB2 + T::B2B;
  uint8_t* sStage = sOnes + kOnes;
  uint8_t* sH = sStage + T::NS * kStage;
  uint8_t* sOut = sH + kHidTile;

  if (tid == 0) {
    for (int s = 0; s < T::NS; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&d1_full[i], 1);
      mbar_init(&acc_empty[i], kMlpEpi2);
      mbar_init(&d2_full[i], 1);
    }
    for (int q = 0; q < 4; ++q) {
      mbar_init(&h_full[q], kMlpEpi1);
      mbar_init(&h_empty[q], 1);
    }
    mbar_init(res_bar, 1);
    fence_mbar_init();
  }
  if (tid < 128) {
    *reinterpret_cast<uint4*>(sOnes + tid * 16) = make_uint4(0x3F803F80u, 0u, 0u, 0u);
    *reinterpret_cast<uint4*>(sOnes + 2048 + tid * 16) = make_uint4(0u, 0u, 0u, 0u);
    fence_async_smem();
  }
  if (warp == kMlpIssuer) {
    tmem_alloc(&tmem_slot, 512);
    tmem_relinquish();
    tc_fence_before();
  }
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;

  if (warp == kMlpLoader) {
    const uint32_t leader = elect_one();
    mbar_arrive_expect_tx_p(leader, res_bar, (uint32_t)(T::W1B + T::B1B + T::W2B + T::B2B));
    bulk_g2s_p(leader, sW1, a.b_res, T::W1B, res_bar);
    if (T::B1B) bulk_g2s_p(leader, sB1, a.bias_img, T::B1B, res_bar);
    bulk_g2s_p(leader, sW2, a.w2_img, T::W2B, res_bar);
    bulk_g2s_p(leader, sB2, a.bias2_img, T::B2B, res_bar);
    pdl_wait();
    uint32_t u = 0;
    for (int item = blockIdx.x; item < a.nitems; item += gridDim.x, ++u) {
      const uint32_t s = u % T::NS;
      mbar_wait(&empty[s], ((u / T::NS) & 1u) ^ 1u);
      mbar_arrive_expect_tx_p(leader, &full[s], (uint32_t)kStage);
      bulk_g2s_p(leader, sStage + s * kStage, a.a_img + (size_t)item * kStage, kStage, &full[s]);
      __syncwarp();
    }
  } else if (warp == kMlpIssuer) {
    const uint32_t leader = elect_one();
    pdl_wait();
    mbar_wait(res_bar, 0);
    const uint32_t idesc1 = idesc_bf16(128, 256), idesc2 = idesc_bf16(128, T::N2);
    const uint64_t dW1 = chunk_desc(smem_u32(sW1), 256, 0), dW2 = chunk_desc(smem_u32(sW2), T::N2, 0);
    const uint64_t dH = chunk_desc(smem_u32(sH), 128, 0), dOnes = chunk_desc(smem_u32(sOnes), 128, 0);
    const uint32_t w2step = 2u * T::N2 * 16u;
    const int n_mine = a.nitems > (int)blockIdx.x ? (a.nitems - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
    // first GEMM of tile `it` (K1 / 16 MMAs of N = 256, + bias)
    auto gemm1 = [&](uint32_t it) {
      const uint32_t acc = it & 1u, s = it % T::NS;
      mbar_wait(&acc_empty[acc], ((it >> 1) & 1u) ^ 1u);
      mbar_wait(&full[s], (it / T::NS) & 1u);
      tc_fence_after();
      const uint64_t dA = chunk_desc(smem_u32(sStage + s * kStage), 128, 0);
      const uint32_t d = tmem + acc * 256u;
#pragma unroll
      for (int j = 0; j < T::K1 / 16; ++j) umma_bf16_p(leader, d, desc_off(dA, j * 4096), desc_off(dW1, j * 8192), idesc1, j != 0);
      if (T::B1B) umma_bf16_p(leader, d, dOnes, chunk_desc(smem_u32(sB1), 256, 0), idesc1, 1u);
      umma_commit_p(leader, &empty[s]);
      umma_commit_p(leader, &d1_full[acc]);
    };
    if (n_mine > 0) gemm1(0);
    for (uint32_t it = 0; it < (uint32_t)n_mine; ++it) {
      if (it + 1 < (uint32_t)n_mine) gemm1(it + 1);
      // second GEMM of tile `it`, chunk by chunk as the hidden tile arrives; its accumulator covers
      // columns [0, N2) of this tile's first accumulator, which chunks 0 .. ceil(N2 / 64) - 1 have drained
      const uint32_t d = tmem + (it & 1u) * 256u, par = it & 1u;
      constexpr int kFirst = T::N2 > 64 ? 2 : 1;
#pragma unroll
      for (int q = 0; q < kFirst; ++q) mbar_wait(&h_full[q], par);
      tc_fence_after();
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        if (q >= kFirst) {
          mbar_wait(&h_full[q], par);
          tc_fence_after();
        }
#pragma unroll
        for (int j = 0; j < 4; ++j)
          umma_bf16_p(leader, d, desc_off(dH, (q * 4 + j) * 4096), desc_off(dW2, (q * 4 + j) * w2step), idesc2, (q | j) != 0);
        if (q == 3) umma_bf16_p(leader, d, dOnes, chunk_desc(smem_u32(sB2), T::N2, 0), idesc2, 1u);
        umma_commit_p(leader, &h_empty[q]);
      }
      umma_commit_p(leader, &d2_full[it & 1u]);
      __syncwarp();
    }
  } else if (warp < kMlpEpi1Warps) {
    // ---- first epilogue: 2 gelu(accumulator 1) -> hidden tile, 64 columns per hand-over ----
    pdl_wait();
    const int q4 = warp & 3, part = warp >> 2, row = q4 * 32 + lane;
    const uint32_t trow = tmem + ((uint32_t)(q4 * 32) << 16);
    uint32_t it = 0;
    for (int item = blockIdx.x; item < a.nitems; item += gridDim.x, ++it) {
      const uint32_t d = trow + (it & 1u) * 256u;
      mbar_wait(&d1_full[it & 1u], (it >> 1) & 1u);
      tc_fence_after();
      uint32_t r[2][16];
      tmem_ld16(d + (uint32_t)(part * 16), r[0]);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        tmem_ld_wait();
        if (q < 3) tmem_ld16(d + (uint32_t)((q + 1) * 64 + part * 16), r[(q + 1) & 1]);   // next chunk's columns, under this chunk's GELU
        mbar_wait(&h_empty[q], (it & 1u) ^ 1u);   // the second GEMM of the previous tile has read this chunk
        cols16_to_image<true>(r[q & 1], sH, row, q * 64 + part * 16);
        tc_fence_before();
        fence_async_smem();
        mbar_arrive(&h_full[q]);
      }
    }
  } else {
    // ---- second epilogue: accumulator 2 -> output ----
    pdl_wait();
    const int q4 = warp & 3, row = q4 * 32 + lane, t2 = tid - kMlpEpi1;
    const uint32_t trow = tmem + ((uint32_t)(q4 * 32) << 16);
    uint32_t it = 0;
    for (int item = blockIdx.x; item < a.nitems; item += gridDim.x, ++it) {
      const uint32_t acc = it & 1u, d = trow + acc * 256u;
      if (KIND == 0) {
        if (t2 == 0) bulk_wait_read0();
        bar_sync(3, kMlpEpi2);
        mbar_wait(&d2_full[acc], (it >> 1) & 1u);
        tc_fence_after();
#pragma unroll 1
        for (int i = 0; i < 8; ++i) {
          uint32_t r[16];
          tmem_ld16(d + (uint32_t)(i * 16), r);
          tmem_ld_wait();
          cols16_to_image<false>(r, sOut, row, i * 16);
        }
        tc_fence_before();
        mbar_arrive(&acc_empty[acc]);
        fence_async_smem();
        bar_sync(4, kMlpEpi2);
        if (t2 == 0) {
          bulk_s2g(a.out_img + (size_t)item * kXTile, sOut, kXTile);
          bulk_commit();
        }
      } else {
        const int w = (item % a.TW) * 128 + row, rid = item / a.TW;
        const int b = rid / a.H, h = rid - b * a.H;
        const bool pvalid = w < a.W;
        const size_t plane = (size_t)a.H * a.W, px = (size_t)h * a.W + w;
        float lastv[16];
#pragma unroll
        for (int j = 0; j < 16; ++j)
          lastv[j] = (pvalid && a.last && j < a.Cg) ? a.last[(size_t)b * a.last_bstride + j * plane + px] : 0.f;
        mbar_wait(&d2_full[acc], (it >> 1) & 1u);
        tc_fence_after();
        uint32_t r[16];
        tmem_ld16(d, r);
        tmem_ld_wait();
        tc_fence_before();
        mbar_arrive(&acc_empty[acc]);
        if (pvalid) {
#pragma unroll
          for (int j = 0; j < 16; ++j)
            if (j < a.Cg) a.out[(size_t)b * a.out_bstride + j * plane + px] = __uint_as_float(r[j]) + lastv[j];
        }
      }
    }
    if (KIND == 0 && t2 == 0) bulk_wait0();
  }
  __syncthreads();
  if (warp == kMlpIssuer) tmem_dealloc(tmem, 512);
}

// ---- layout kernels ---------------------------------------------------------------------------
// x_t from the NCHW sources -> [128 slots x 32] bf16 input images: columns 0..Cin-1 hi parts,
// columns Cin and 15 ones (they meet the hi and lo parts of the first bias in the weight image),
// columns 16..16+Cin-1 lo parts.
__global__ void __launch_bounds__(128) wide_gather_kernel(const WideGatherArgs a) {
  pdl_launch_dependents();
  pdl_wait();
  // a CTA walks over tiles (one CTA per tile was bound by the rate at which 17 k CTAs of 128 threads are dispatched)
  for (int tile = blockIdx.x; tile < a.ntiles; tile += gridDim.x) {
  const int s = threadIdx.x;
  const int w = (tile % a.TW) * 128 + s, rid = tile / a.TW;
  const int b = rid / a.H, h = rid - b * a.H;
  const size_t plane = (size_t)a.H * a.W, px = (size_t)h * a.W + w;
  float x[14];
#pragma unroll
  for (int c = 0; c < 14; ++c) {
    x[c] = 0.f;
    if (c < a.Cin && w < a.W) {
      int g = 0, ch = c;
      while (ch >= a.grp[g].nch) ch -= a.grp[g++].nch;   // uniform: which source group holds channel c
      x[c] = a.grp[g].base[(size_t)b * a.grp[g].bstride + ch * plane + px];
    }
  }
  uint32_t v[16];   // bf16 pairs: columns (2 i, 2 i + 1)
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const float x0 = 2 * i < 14 ? x[2 * i < 14 ? 2 * i : 0] : 0.f, x1 = 2 * i + 1 < 14 ? x[2 * i + 1 < 14 ? 2 * i + 1 : 0] : 0.f;
    const float h0 = (2 * i == a.Cin || 2 * i == 15) ? 1.f : x0, h1 = (2 * i + 1 == a.Cin || 2 * i + 1 == 15) ? 1.f : x1;
    const __nv_bfloat162 hi = __floats2bfloat162_rn(h0, h1);
    const float2 hf = __bfloat1622float2(hi);
    const __nv_bfloat162 lo = __floats2bfloat162_rn(2 * i < a.Cin ? x0 - hf.x : 0.f, 2 * i + 1 < a.Cin ? x1 - hf.y : 0.f);
    v[i] = *reinterpret_cast<const uint32_t*>(&hi);
    v[8 + i] = *reinterpret_cast<const uint32_t*>(&lo);
  }
  uint8_t* img = a.in_img + (size_t)tile * 8192 + s * 16;   // [128 x 32] bf16 image
#pragma unroll
  for (int j = 0; j < 4; ++j) *reinterpret_cast<uint4*>(img + j * 2048) = make_uint4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
  }
}

}  // namespace

// ---- launchers ---------------------------------------------------------------------------------
template <int MODE>
static int launch_tcw_mode(const TcwArgs& a, cudaStream_t st, const char* name) {
  int sms = 0;
  DLWP_TRY(current_sms(&sms));
  const int smem = tcw_smem_bytes<MODE>(a.KZ);
  DLWP_REQUIRE(smem <= 226 * 1024, "%s: %d bytes of shared memory needed", name, smem);
  DLWP_TRY(ensure_dynamic_smem((const void*)tcw_kernel<MODE>, smem));
  const int grid = a.nitems < sms ? a.nitems : sms;
  void (*kern)(TcwArgs) = tcw_kernel<MODE>;
  DLWP_CUDA(launch_pdl(kern, dim3((unsigned)grid), dim3(kWCta), (size_t)smem, st, a));
  DLWP_LAUNCHED(name);
  return DLWP_OK;
}

int launch_tcw(int mode, const TcwArgs& a, cudaStream_t st) {
  DLWP_REQUIRE(a.nitems > 0 && a.TW > 0 && a.H > 0 && a.W > 0, "tcw: empty problem");
  DLWP_REQUIRE(a.KZ % 16 == 0 && a.KZ >= 16 && a.KZ <= 96 && 2 * a.Kc <= a.KZ, "tcw: %d retained columns do not fit", a.Kc);
  switch (mode) {
    case TCW_FWD: return launch_tcw_mode<TCW_FWD>(a, st, "tcw_fwd_wdft");
    case TCW_HFWD: return launch_tcw_mode<TCW_HFWD>(a, st, "tcw_fwd_hdft");
    case TCW_HINV: return launch_tcw_mode<TCW_HINV>(a, st, "tcw_inv_hdft");
    case TCW_LAYER: return launch_tcw_mode<TCW_LAYER>(a, st, "tcw_layer");
  }
  return fail(DLWP_ERR_INVALID, "tcw: unknown mode %d", mode);
}

template <int KIND>
static int launch_tcw_mlp_kind(const TcwArgs& a, cudaStream_t st, const char* name) {
  int sms = 0;
  DLWP_TRY(current_sms(&sms));
  constexpr int smem = mlp_smem_bytes<KIND>();
  static_assert(smem <= 226 * 1024, "fused MLP kernel: shared memory");
  DLWP_TRY(ensure_dynamic_smem((const void*)tcw_mlp_kernel<KIND>, smem));
  const int grid = a.nitems < sms ? a.nitems : sms;
  void (*kern)(TcwArgs) = tcw_mlp_kernel<KIND>;
  DLWP_CUDA(launch_pdl(kern, dim3((unsigned)grid), dim3(kMlpCta), (size_t)smem, st, a));
  DLWP_LAUNCHED(name);
  return DLWP_OK;
}

int launch_tcw_mlp(int kind, const TcwArgs& a, cudaStream_t st) {
  DLWP_REQUIRE(a.nitems > 0 && a.TW > 0 && a.H > 0 && a.W > 0 && a.w2_img && a.bias2_img, "tcw_mlp: incomplete arguments");
  if (kind == 0) return launch_tcw_mlp_kind<0>(a, st, "tcw_lift");
  if (kind == 1) return launch_tcw_mlp_kind<1>(a, st, "tcw_proj");
  return fail(DLWP_ERR_INVALID, "tcw_mlp: unknown kind %d", kind);
}

int launch_wide_gather(const WideGatherArgs& a, cudaStream_t st) {
  DLWP_REQUIRE(a.Cin >= 1 && a.Cin <= 14 && a.ngrp >= 1 && a.ngrp <= 4, "wide_gather: %d input channels", a.Cin);
  void (*kern)(WideGatherArgs) = wide_gather_kernel;
  int sms = 0;
  DLWP_TRY(current_sms(&sms));
  const int grid = a.ntiles < 16 * sms ? a.ntiles : 16 * sms;      // 16 CTAs of 128 threads per SM
  DLWP_CUDA(launch_pdl(kern, dim3((unsigned)grid), dim3(128), 0, st, a));
  DLWP_LAUNCHED("wide_gather_kernel");
  return DLWP_OK;
}

// ---- host packing --------------------------------------------------------------------------------
template <class S>
static void pack_image_t(const S* m, int rows, int cols, int ld, int R, int K, uint8_t* out) {
  std::memset(out, 0, (size_t)R * K * 2);
  for (int r = 0; r < rows; ++r)
    for (int k = 0; k < cols; ++k) {
      const __nv_bfloat16 v = __float2bfloat16((float)m[(size_t)r * ld + k]);
      std::memcpy(out + (size_t)(k >> 3) * R * 16 + (size_t)r * 16 + (k & 7) * 2, &v, 2);
    }
}
void tcw_pack_image(const float* m, int rows, int cols, int ld, int R, int K, uint8_t* out) {
  pack_image_t(m, rows, cols, ld, R, K, out);
}
void tcw_pack_image_f64(const double* m, int rows, int cols, int ld, int R, int K, uint8_t* out) {
  pack_image_t(m, rows, cols, ld, R, K, out);
}

}  // namespace dlwp
