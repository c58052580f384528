// afno_mid_tc_kernel: the retained-column part of AFNO2D (fourcastnet.py:85-124) on tcgen05 tensor cores:
//   X[ky, kx, c]  = sum_h Fh[ky, h] A[h, kx, c]                      forward DFT along H           (rfft2, :85)
//   O1 = ReLU(X W1 + b1), O2 = O1 W2 + b2 per channel block           block-diagonal complex MLP    (:96-119)
//   Y = softshrink(O2) on re and im                                                                  (:121)
//   Z[h, kx, c]   = sum_ky Gh[h, ky] Y[ky, kx, c]                     inverse DFT along H           (irfft2, :124)
// as four chained real GEMMs on (re | im)-stacked indices, every product in split bf16 (a_hi b_hi + a_lo b_hi +
// a_hi b_lo, fp32 accumulation in tensor memory).  The fp32 FMA form of the same mathematics is afno_mid_kernel
// (afno_fused.cu); it stays for the geometries this kernel does not cover.
//
// Work item = 4 consecutive retained columns of the flattened (sample, kx) axis = 128 spectral tokens (column, ky).
//
//   GEMM 1  D1[(re|im, ky), (col, c)] = Fs[64, K = 2H] * Aimg                 M = 64, N = 256, K = 64, k = 2h + (re|im)
//           Aimg arrives from afno_fwd_tc_kernel as a ready MN-major B operand (K rows of 8 channels).
//   E 1     lane = (re|im, ky) writes its 16-byte rows of the token-major MLP operand X[tok = (col, ky)][k],
//           k = (block, re|im, i): one 128-token K-major A operand, 16 chunks of 8 k.
//   GEMM 2a per channel block: D2[tok, (re|im, o)] = X[tok, K block] * W1s[(re|im, o), K block]     M = 128, N = K = 2 bs
//   E 2a    + b1, ReLU, split -> the same operand rows in place (O1)
//   GEMM 2b / E 2b   the same with W2s, + b2, softshrink -> MN-major B operand of GEMM 3: K rows k = (re|im, ky) of 8 channels
//   GEMM 3  D3[(h, re|im), (col, c)] = Gs[64, K = 64] * Yimg                  M = 64, N = 256, K = 64
//   E 3     lane = (h, re|im) = one K row of the Z' operand image afno_inv_tc_kernel reads (MN-major: K rows
//           (grid row of the tile, kx, re|im) of 8 channels): 16-byte stores, re and im rows adjacent (whole sectors).
//
// The four operands of an item overlay each other in one 64 KB buffer (each is dead when the next is written);
// two buffers: the next item's A images are in flight while this item computes.
#include <cstring>
#include <vector>

#include "tc_afno.cuh"
#include "tc_common.cuh"
#include "tc_fno.cuh"

namespace dlwp {

using namespace tc;

#ifdef DLWP_TRACE
__device__ unsigned long long* g_trace_amid = nullptr;      // [item < 4][24] clock64 stamps of block 0: control warp 0..11, epilogue warp 0 at 12..23
#define TRM(it, ev) do { if (g_trace_amid && blockIdx.x == 0 && (threadIdx.x & 31) == 0 && (it) < 4) g_trace_amid[(it) * 24 + (ev)] = clock64(); } while (0)
#else
#define TRM(it, ev) do {} while (0)
#endif

namespace {

constexpr int kEpiWarps = 16;
constexpr int kThreads = (kEpiWarps + 1) * 32;      // 16 epilogue warps + 1 control warp
constexpr int kCols = 4;                            // retained columns per item
constexpr int kPart = 32768;                        // hi (or lo) half of an item's operand buffer
constexpr int kBuf = 2 * kPart;

struct Smem {      // byte offsets
  static constexpr int F = 0, G = F + 16384, W1 = G + 16384, W2 = W1 + 16384, Bias = W2 + 16384, Buf = Bias + 1024,
                       Total = Buf + 2 * kBuf;
};

// (D2AFULL / HFULL: one barrier per group of 32 accumulator columns = the channel block(s) warp share j works on: the first
//  layer's epilogue of a group and the second layer's UMMAs of that group run under the other groups' first-layer UMMAs)
enum { MB_CONST = 0, MB_AFULL, MB_D1FULL = MB_AFULL + 2, MB_XFULL, MB_D2AFULL, MB_HFULL = MB_D2AFULL + 4, MB_D2BFULL = MB_HFULL + 4, MB_YFULL,
       MB_D3FULL, MB_N };
constexpr uint32_t kD2Col = 256;

__device__ __forceinline__ void split2p(float2 v, uint32_t& hi, uint32_t& lo) {
  __nv_bfloat162 h = __floats2bfloat162_rn(v.x, v.y);
  hi = *reinterpret_cast<uint32_t*>(&h);
  const float2 hf = make_float2(__uint_as_float(hi << 16), __uint_as_float(hi & 0xffff0000u));
  const float2 d = __ffma2_rn(hf, make_float2(-1.0f, -1.0f), v);
  __nv_bfloat162 l = __floats2bfloat162_rn(d.x, d.y);
  lo = *reinterpret_cast<uint32_t*>(&l);
}

// eight fp32 accumulator values -> one 16-byte row of the hi image and one of the lo image
__device__ __forceinline__ void split8_store(const uint32_t* v, uint8_t* dst, uint32_t lo_off) {
  uint32_t hi[4], lo[4];
#pragma unroll
  for (int e = 0; e < 4; ++e) split2p(make_float2(__uint_as_float(v[2 * e]), __uint_as_float(v[2 * e + 1])), hi[e], lo[e]);
  *reinterpret_cast<uint4*>(dst) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
  *reinterpret_cast<uint4*>(dst + lo_off) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
}

__device__ __forceinline__ float softshrink1(float v, float lambda) { return v > lambda ? v - lambda : (v < -lambda ? v + lambda : 0.f); }

}  // namespace

__global__ void __launch_bounds__(kThreads, 1) afno_mid_tc_kernel(const AfnoMid2Args a) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ uint64_t bars[MB_N];
  __shared__ uint32_t tmem_s;
  pdl_launch_dependents();
  const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;
  const int nitems = (a.ncols + kCols - 1) / kCols, stride = gridDim.x, first = blockIdx.x;
  const int bs = a.bs, nblk = kAfnoE / bs;

  if (tid == 0) {
    mbar_init(&bars[MB_CONST], 1);
    mbar_init(&bars[MB_AFULL], 1);
    mbar_init(&bars[MB_AFULL + 1], 1);
    mbar_init(&bars[MB_D1FULL], 1);
    mbar_init(&bars[MB_XFULL], 512);
    for (int g = 0; g < 4; ++g) {
      mbar_init(&bars[MB_D2AFULL + g], 1);
      mbar_init(&bars[MB_HFULL + g], 128);
    }
    mbar_init(&bars[MB_D2BFULL], 1);
    mbar_init(&bars[MB_YFULL], 512);
    mbar_init(&bars[MB_D3FULL], 1);
    fence_mbar_init();
  }
  if (warp == kEpiWarps) {
    tmem_alloc(&tmem_s, 512);
    tmem_relinquish();
  }
  // columns beyond the last one of a ragged final item are computed on whatever the buffers hold: keep that finite
  for (int i = tid; i < 2 * kBuf / 16; i += kThreads) reinterpret_cast<uint4*>(smem + Smem::Buf)[i] = make_uint4(0, 0, 0, 0);
  if (tid < 2 * kAfnoE) {      // biases, in accumulator column order (block, re|im, o): constants, not written by the predecessor
    reinterpret_cast<float*>(smem + Smem::Bias)[tid] = a.b1[tid];
    reinterpret_cast<float*>(smem + Smem::Bias)[2 * kAfnoE + tid] = a.b2[tid];
  }
  fence_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t wbytes = (uint32_t)(nblk * 2 * (2 * bs) * (2 * bs) * 2);      // per layer: nblk x [hi | lo] x 2bs rows x 2bs k
  if (tid == 0) {      // constant operands: requested before the dependency wait
    mbar_arrive_expect_tx(&bars[MB_CONST], 2u * 16384u + 2u * wbytes);
    bulk_g2s(smem + Smem::F, a.Fimg, 16384u, &bars[MB_CONST]);
    bulk_g2s(smem + Smem::G, a.Gimg, 16384u, &bars[MB_CONST]);
    bulk_g2s(smem + Smem::W1, a.W1img, wbytes, &bars[MB_CONST]);
    bulk_g2s(smem + Smem::W2, a.W2img, wbytes, &bars[MB_CONST]);
  }
  if (warp == kEpiWarps) TRM(0, 0);
  pdl_wait();      // Aimg comes from the previous kernel; Zimg is read by the kernel before it
  if (warp == kEpiWarps) TRM(0, 1);
  const uint32_t tmem = tmem_s;

  if (warp == kEpiWarps) {
    // ------------------------------ control warp: image loads and every tcgen05.mma ------------------------------
    const uint32_t leader = elect_one();
    auto load = [&](int item, int s) {      // hi and lo images of the item's columns: two copies of up to 32 KB
      const int c0 = item * kCols, nc = a.ncols - c0 < kCols ? a.ncols - c0 : kCols;
      const uint32_t bytes = (uint32_t)nc * 8192u;
      mbar_arrive_expect_tx_p(leader, &bars[MB_AFULL + s], 2u * bytes);
      bulk_g2s_p(leader, smem + Smem::Buf + s * kBuf, a.Aimg + (size_t)c0 * 8192, bytes, &bars[MB_AFULL + s]);
      bulk_g2s_p(leader, smem + Smem::Buf + s * kBuf + kPart, a.Aimg + ((size_t)a.ncols + c0) * 8192, bytes, &bars[MB_AFULL + s]);
    };
    if (first < nitems) load(first, 0);
    mbar_wait(&bars[MB_CONST], 0);
    const uint32_t aF = smem_u32(smem + Smem::F), aG = smem_u32(smem + Smem::G), aW1 = smem_u32(smem + Smem::W1), aW2 = smem_u32(smem + Smem::W2);
    const uint32_t idesc13 = idesc_bf16(64, 256) | kIdescBMnMajor, idesc2 = idesc_bf16(128, (uint32_t)(2 * bs));
    const uint32_t wblk = (uint32_t)(2 * (2 * bs) * (2 * bs) * 2), wlo = wblk / 2;      // one block's [hi | lo] images; offset of lo
    const uint32_t wchunk = (uint32_t)(2 * bs * 16);                                      // K-chunk stride of a block image
    const uint64_t dF = chunk_desc(aF, 64, 0), dG = chunk_desc(aG, 64, 0);
    auto gemm1 = [&](uint32_t n) {      // item n (its images are in buffer n & 1) -> D1
      mbar_wait(&bars[MB_AFULL + (n & 1u)], (n >> 1) & 1u);
      TRM(n, 2);
      tc_fence_after();
      const uint64_t dA = smem_desc(smem_u32(smem + Smem::Buf + (n & 1u) * kBuf), 128, 1024);
#pragma unroll
      for (int term = 0; term < 3; ++term) {      // hi*hi, lo*hi, hi*lo
        const uint32_t fo = term == 1 ? 8192u : 0u, ao = term == 2 ? (uint32_t)kPart : 0u;
#pragma unroll
        for (int k0 = 0; k0 < 64; k0 += 16)
          umma_bf16_p(leader, tmem, desc_off(dF, fo + (uint32_t)(k0 >> 3) * 1024u), desc_off(dA, ao + (uint32_t)k0 * 16u), idesc13, (term | k0) != 0);
      }
      umma_commit_p(leader, &bars[MB_D1FULL]);
      TRM(n, 3);
    };
    if (first < nitems) {
      if (first + stride < nitems) load(first + stride, 1);
      gemm1(0);
    }
    uint32_t n = 0;
    for (int item = first; item < nitems; item += stride, ++n) {
      const uint32_t par = n & 1u, s = n & 1u;
      const uint32_t aB = smem_u32(smem + Smem::Buf + s * kBuf);
      // GEMM 2a / 2b: one small GEMM per channel block (K = N = 2 bs) on the token-major operand, in groups of 32 columns
      const int bpg = nblk / 4;      // channel blocks per group: 1 (bs = 16) or 2 (bs = 8)
      for (int layer = 0; layer < 2; ++layer) {
        if (layer == 0) {
          mbar_wait(&bars[MB_XFULL], par);
          TRM(n, 4);
          tc_fence_after();
        }
        const uint32_t aW = layer == 0 ? aW1 : aW2;
        for (int g = 0; g < 4; ++g) {
          if (layer == 1) {      // the group's K chunks have been rewritten by its first-layer epilogue
            mbar_wait(&bars[MB_HFULL + g], par);
            if (g == 0) TRM(n, 6);
            tc_fence_after();
          }
          for (int blk = g * bpg; blk < (g + 1) * bpg; ++blk) {
            const uint64_t dX = chunk_desc(aB, 128, (uint32_t)(blk * 2 * bs)), dW = smem_desc(aW + (uint32_t)blk * wblk, wchunk, 128);
            const uint32_t d2 = tmem + kD2Col + (uint32_t)(blk * 2 * bs);
#pragma unroll
            for (int term = 0; term < 3; ++term) {      // X_hi W_hi, X_lo W_hi, X_hi W_lo
              const uint32_t xo = term == 1 ? (uint32_t)kPart : 0u, wo = term == 2 ? wlo : 0u;
#pragma unroll
              for (int k0 = 0; k0 < 32; k0 += 16)      // (unrolled with a uniform guard: bs is 8 or 16)
                if (k0 < 2 * bs)
                  umma_bf16_p(leader, d2, desc_off(dX, xo + (uint32_t)(k0 >> 3) * 2048u), desc_off(dW, wo + (uint32_t)(k0 >> 3) * wchunk), idesc2, (term | k0) != 0);
            }
          }
          if (layer == 0) umma_commit_p(leader, &bars[MB_D2AFULL + g]);
        }
        if (layer == 1) umma_commit_p(leader, &bars[MB_D2BFULL]);
        TRM(n, 5 + 2 * layer);
      }
      // GEMM 3 (its accumulator takes the columns of D2, dead by now, and the 128 after them)
      mbar_wait(&bars[MB_YFULL], par);
      TRM(n, 8);
      tc_fence_after();
      {
        const uint64_t dY = smem_desc(aB, 128, 1024);
#pragma unroll
        for (int term = 0; term < 3; ++term) {      // Gs_hi Y_hi, Gs_lo Y_hi, Gs_hi Y_lo
          const uint32_t go = term == 1 ? 8192u : 0u, yo = term == 2 ? (uint32_t)kPart : 0u;
#pragma unroll
          for (int k0 = 0; k0 < 64; k0 += 16)
            umma_bf16_p(leader, tmem + kD2Col, desc_off(dG, go + (uint32_t)(k0 >> 3) * 1024u), desc_off(dY, yo + (uint32_t)k0 * 16u), idesc13, (term | k0) != 0);
        }
        umma_commit_p(leader, &bars[MB_D3FULL]);
        TRM(n, 9);
      }
      // the next item's GEMM 1 runs under this item's last epilogue (D1 was read long ago; the next GEMM 2a, which shares
      // columns with D3, is ordered behind E 3 by the epilogue warps' arrival on XFULL)
      if (item + stride < nitems) gemm1(n + 1);
      if (item + 2 * stride < nitems) {      // this item's buffer is free when GEMM 3 has read Y
        mbar_wait(&bars[MB_D3FULL], par);
        load(item + 2 * stride, (int)s);
      }
      __syncwarp();
    }
  } else {
    // ------------------------------ epilogue warps: quarter q of the TMEM lanes, share j of the columns ------------------------------
    const int q = warp & 3, j = warp >> 2;
    const uint32_t tq = tmem + ((uint32_t)(q * 32) << 16);
    const float* sb1 = reinterpret_cast<const float*>(smem + Smem::Bias) + j * 32;
    const float* sb2 = sb1 + 2 * kAfnoE;
    const float lambda = a.lambda;
    const int cshift = bs == 16 ? 1 : 0;      // channel octets per (block, re|im) group: 2 or 1
    uint32_t n = 0;
    for (int item = first; item < nitems; item += stride, ++n) {
      const uint32_t par = n & 1u;
      uint8_t* buf = smem + Smem::Buf + (n & 1u) * kBuf;
      {
        // ---- E 1: D1 row (re|im, ky) (M = 64: lanes < 16 of quarter q hold rows 16 q + lane), column j of the item: its 64 channels ----
        mbar_wait(&bars[MB_D1FULL], par);
        if (warp == 0) TRM(n, 12);
        tc_fence_after();
        const int row = 16 * q + lane, reim = row >> 5, ky = row & 31;
        uint8_t* dst = buf + (j * 32 + ky) * 16;
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          uint32_t v[32];
          tmem_ld32(tq + (uint32_t)(j * 64 + half * 32), v);
          tmem_ld_wait();
          if (lane < 16) {
#pragma unroll
            for (int oc = 0; oc < 4; ++oc) {      // channel octet 4 half + oc -> K chunk (block, re|im, octet in block)
              const int o8 = 4 * half + oc;
              const int chunk = ((o8 >> cshift) << (cshift + 1)) + (reim << cshift) + (o8 & cshift);
              split8_store(v + 8 * oc, dst + chunk * 2048, kPart);
            }
          }
        }
        tc_fence_before();
        fence_async_smem();
        mbar_arrive(&bars[MB_XFULL]);
        if (warp == 0) TRM(n, 13);
      }
      // ---- E 2a / E 2b: lane = token 32 q + lane, accumulator columns 32 j .. 32 j + 31 = k of the next operand ----
      {
        mbar_wait(&bars[MB_D2AFULL + j], par);
        if (warp == 0) TRM(n, 14);
        tc_fence_after();
        uint32_t v[32];
        tmem_ld32(tq + kD2Col + (uint32_t)(j * 32), v);
        tmem_ld_wait();
#pragma unroll
        for (int e = 0; e < 32; ++e) v[e] = __float_as_uint(fmaxf(__uint_as_float(v[e]) + sb1[e], 0.f));
        uint8_t* dst = buf + (j * 4) * 2048 + (q * 32 + lane) * 16;
#pragma unroll
        for (int oc = 0; oc < 4; ++oc) split8_store(v + 8 * oc, dst + oc * 2048, kPart);
        tc_fence_before();
        fence_async_smem();
        mbar_arrive(&bars[MB_HFULL + j]);
        if (warp == 0) TRM(n, 15);
      }
      {
        mbar_wait(&bars[MB_D2BFULL], par);
        if (warp == 0) TRM(n, 16);
        tc_fence_after();
        uint32_t v[32];
        tmem_ld32(tq + kD2Col + (uint32_t)(j * 32), v);
        tmem_ld_wait();
#pragma unroll
        for (int e = 0; e < 32; ++e) v[e] = __float_as_uint(softshrink1(__uint_as_float(v[e]) + sb2[e], lambda));
        // columns (block, re|im, o) -> channel octet of the item column q, K row (re|im, ky = lane)
#pragma unroll
        for (int oc = 0; oc < 4; ++oc) {
          const int g = 4 * j + oc;      // group of 8 accumulator columns
          const int reim = bs == 16 ? (g >> 1) & 1 : g & 1;
          const int c8 = bs == 16 ? ((g >> 2) << 1) + (g & 1) : g >> 1;
          split8_store(v + 8 * oc, buf + (q * 8 + c8) * 1024 + (reim * 32 + lane) * 16, kPart);
        }
        tc_fence_before();
        fence_async_smem();
        mbar_arrive(&bars[MB_YFULL]);
        if (warp == 0) TRM(n, 17);
      }
      {
        // ---- E 3: D3 row (h, re|im) = 16 q + lane (lanes < 16), column j of the item -> K row of the Z' image of its tile ----
        mbar_wait(&bars[MB_D3FULL], par);
        if (warp == 0) TRM(n, 18);
        tc_fence_after();
        const int row = 16 * q + lane, h = row >> 1, reim = row & 1;
        const int col = item * kCols + j, b = col / a.Kc, kx = col - b * a.Kc;
        const bool live = lane < 16 && col < a.ncols;
        const int grow = b * a.H + h, tile = grow / a.RT, h2 = grow - tile * a.RT;
        const size_t zpart = (size_t)a.KZ * 128;      // 8 channel octets x KZ rows x 16 bytes
        uint8_t* dst = a.Zimg + (size_t)tile * (2 * zpart) + (size_t)((h2 * a.Kc + kx) * 2 + reim) * 16;
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          uint32_t v[32];
          tmem_ld32(tq + kD2Col + (uint32_t)(j * 64 + half * 32), v);
          tmem_ld_wait();
          if (live) {
#pragma unroll
            for (int oc = 0; oc < 4; ++oc) split8_store(v + 8 * oc, dst + (size_t)(4 * half + oc) * (a.KZ * 16), (uint32_t)zpart);
          }
        }
        tc_fence_before();
        if (warp == 0) TRM(n, 19);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == kEpiWarps) tmem_dealloc(tmem, 512);
}

bool afno_mid_tc_covers(int H, int R, int bs) { return H == 32 && R >= 1 && R <= 32 && (bs == 8 || bs == 16); }

int launch_afno_mid_tc(const AfnoMid2Args& a, cudaStream_t st) {
  int sms = 0;
  DLWP_TRY(current_sms(&sms));
  DLWP_REQUIRE(afno_mid_tc_covers(a.H, a.R, a.bs) && a.ncols >= 1 && a.Kc >= 1 && a.RT >= 1 && a.H % a.RT == 0 && a.RT * 2 * a.Kc <= a.KZ,
               "afno_mid_tc_kernel: unsupported geometry H=%d R=%d bs=%d", a.H, a.R, a.bs);
  DLWP_TRY(ensure_dynamic_smem((const void*)afno_mid_tc_kernel, Smem::Total));
  const int nitems = (a.ncols + kCols - 1) / kCols;
  const int grid = nitems < sms ? nitems : sms;
  DLWP_CUDA(launch_pdl(afno_mid_tc_kernel, dim3((unsigned)grid), dim3(kThreads), (size_t)Smem::Total, st, a));
  DLWP_LAUNCHED("afno_mid_tc_kernel");
  return DLWP_OK;
}

// Fs [64 rows (re|im, ky)][K = 2H, k = 2h + (re|im)]:  X = Fh A:  Xr = Fr Ar - Fi Ai,  Xi = Fi Ar + Fr Ai
// Gs [64 rows 2h + (re|im)][K = 64, k = (re|im) 32 + ky]:  Z = Gh Y likewise.    Fh [R, H], Gh [H, R] complex (plan.cu: build_bases)
void afno_mid_tc_pack_bases(const double* Fr, const double* Fi, const double* Gr, const double* Gi, int H, int R, uint8_t* Fimg, uint8_t* Gimg) {
  std::vector<double> F((size_t)64 * 64, 0.0), G((size_t)64 * 64, 0.0);
  for (int ky = 0; ky < R; ++ky)
    for (int h = 0; h < H; ++h) {
      const double fr = Fr[ky * H + h], fi = Fi[ky * H + h], gr = Gr[h * R + ky], gi = Gi[h * R + ky];
      F[(size_t)ky * 64 + 2 * h] = fr;
      F[(size_t)ky * 64 + 2 * h + 1] = -fi;
      F[(size_t)(32 + ky) * 64 + 2 * h] = fi;
      F[(size_t)(32 + ky) * 64 + 2 * h + 1] = fr;
      G[(size_t)(2 * h) * 64 + ky] = gr;
      G[(size_t)(2 * h) * 64 + 32 + ky] = -gi;
      G[(size_t)(2 * h + 1) * 64 + ky] = gi;
      G[(size_t)(2 * h + 1) * 64 + 32 + ky] = gr;
    }
  tc_pack_images_f64(F.data(), 64, 64, 64, 64, 64, Fimg);
  tc_pack_images_f64(G.data(), 64, 64, 64, 64, 64, Gimg);
}

// One layer of the block MLP, w [2, nb, bs, bs] (planes real / imaginary, fourcastnet.py:73-76), b [2, nb, bs]  ->
// per block the [hi | lo] images of Ws[(re|im, o)][k = (re|im, i)] and the bias in column order (block, re|im, o):
//   out_re[o] = sum_i x_re[i] w0[i][o] - x_im[i] w1[i][o],   out_im[o] = sum_i x_im[i] w0[i][o] + x_re[i] w1[i][o]
void afno_mid_tc_pack_layer(const float* w, const float* b, int nb, int bs, uint8_t* Wimg, float* bias) {
  const int n2 = 2 * bs;
  std::vector<float> m((size_t)n2 * n2);
  for (int k = 0; k < nb; ++k) {
    const float* w0 = w + (size_t)k * bs * bs;
    const float* w1 = w + (size_t)(nb + k) * bs * bs;
    for (int o = 0; o < bs; ++o)
      for (int i = 0; i < bs; ++i) {
        m[(size_t)o * n2 + i] = w0[i * bs + o];
        m[(size_t)o * n2 + bs + i] = -w1[i * bs + o];
        m[(size_t)(bs + o) * n2 + i] = w1[i * bs + o];
        m[(size_t)(bs + o) * n2 + bs + i] = w0[i * bs + o];
      }
    tc_pack_images(m.data(), n2, n2, n2, n2, n2, Wimg + (size_t)k * 2 * n2 * n2 * 2);
    for (int o = 0; o < bs; ++o) {
      bias[k * n2 + o] = b[k * bs + o];
      bias[k * n2 + bs + o] = b[(nb + k) * bs + o];
    }
  }
}

}  // namespace dlwp

#ifdef DLWP_TRACE
extern "C" int dlwp_debug_set_trace_afno_mid(unsigned long long* p) {
  return cudaMemcpyToSymbol(dlwp::g_trace_amid, &p, sizeof(p)) == cudaSuccess ? 0 : -2;
}
#endif
