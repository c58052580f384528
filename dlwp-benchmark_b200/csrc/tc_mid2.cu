// fno_mid_tc_kernel: the retained-column part of a spectral layer on tcgen05 tensor cores:
//   X[b, r, kx, i] = sum_h  Fh[r, h] * A[b, h, kx, i]            forward truncated DFT along H
//   Y[b, r, kx, o] = sum_i  X[b, r, kx, i] * W[i, o, r, kx]      einsum("bixy,ioxy->boxy"), unet.py:15-17 / neuralop SpectralConv
//   Z[b, h, kx, o] = sum_r  Gh[h, r] * Y[b, r, kx, o]            inverse DFT along H
// as three chained real GEMMs on (re | im)-stacked indices, every product in split bf16 (a_hi b_hi + a_lo b_hi + a_hi b_lo,
// fp32 accumulation in tensor memory) like the rest of the fp32-parity kernel set.  The fp32 FMA form of the same
// mathematics is fno_mid_kernel (tc_mid.cu): it stays for the geometries this kernel does not cover.
//
// Work item = (retained column kx, NB consecutive samples); CTA = one kx, a strided share of its sample groups; the
// 96 KB of mode-weight images of the column stay in shared memory for the CTA's lifetime.
//
//   GEMM 1  D1[(re | im | -im | re, r), (b, i)] = Fs[row, k] * A[k, (b, i)]      M = 64, N = 32 NB, K = 2H,  k = 2h + (re|im)
//           A arrives from the lift / layer kernels as ready MN-major operand tiles (tc_fno.cuh).  The four row blocks
//           land in the four TMEM lane quarters: lane = retained row r, quarter = which copy of X it holds.
//   E 1     lane r owns the rows of ITS mode in the operand images of GEMM 2,  P = [Xr | -Xi]  and  Q = [Xi | Xr]
//           (rows = (P | Q, sample, mode % 4), K = 64, hi and lo parts): quarter 0 writes Xr into P, 1 Xi into Q,
//           2 -Xi into P, 3 Xr into Q, as 16-byte rows.
//   GEMM 2  per group of 4 modes:  D2[(mode, o), (P | Q, b, mode')] = Wm[(mode, o), K] * [P ; Q][(.., b, mode'), K]
//           Wm = [Wr^T | Wi^T];  M = 128, N = 8 NB, K = 64; the blocks mode = mode' are used:
//           columns (P, b): Yr = Xr Wr - Xi Wi,  columns (Q, b): Yi = Xi Wr + Xr Wi.
//   E 2     quarter = mode % 4, lane = o: writes (Yr, Yi) pairs into the operand image of GEMM 3: rows (b, o), K = 2r + (re|im).
//   GEMM 3  D3[(h, re|im), (b, o)] = Gs[(h, re|im), K] * Yimg[(b, o), K]              M = 2H (64 or 128), N = 32 NB, K = 32
//   E 3     lane = K row (h, re|im) of the Z' operand image fno_layer_kernel reads (MN-major: K rows of 8 channels,
//           layout in tc_mid.cu), columns = the 32 channels of a sample: 16-byte stores, 64 contiguous bytes per tile.
//
// X overlays the A tiles (dead once GEMM 1 has completed); the next item's tiles are requested as soon as GEMM 2 has
// read X.  The three accumulators of an item share tensor-memory columns (the stages of an item run one after the other).
#include <cstring>
#include <vector>

#include "tc_common.cuh"
#include "tc_fno.cuh"

namespace dlwp {

using namespace tc;

#ifdef DLWP_TRACE
__device__ unsigned long long* g_trace_mid2 = nullptr;      // [item < 4][24] clock64 stamps of block (0, 0): control warp 0..11, epilogue warp 0 at 12..23
#define TR2(it, ev) do { if (g_trace_mid2 && blockIdx.x == 0 && blockIdx.y == 0 && (threadIdx.x & 31) == 0 && (it) < 4) g_trace_mid2[(it) * 24 + (ev)] = clock64(); } while (0)
#else
#define TR2(it, ev) do {} while (0)
#endif

namespace {

constexpr int kEpiWarps = 16;
constexpr int kMidThreads = (kEpiWarps + 1) * 32;      // 16 epilogue warps + 1 control warp
constexpr int kWGroup = 32768;                         // weight images of 4 modes: hi 16 KB | lo 16 KB (128 rows x 64 K)

struct Mid2Smem {      // byte offsets
  int W, F, G, A, Y, Total, NG, f_part, g_part, a_part, x_stride, y_part;
  __host__ __device__ Mid2Smem(int H, int R, int NB) {
    NG = (R + 3) / 4;                      // groups of 4 modes
    f_part = 64 * 2 * H * 2;               // one part of Fs: 64 rows x 2H
    g_part = (2 * H < 64 ? 64 : 2 * H) * 32 * 2;      // one part of Gs: max(2H, 64) rows x 32
    a_part = NB * H * 128;                 // hi (or lo) tiles of NB samples
    x_stride = 8 * 16 * NB * 16 + 64;      // one group's [P ; Q] images: 8 K chunks x (8 NB hi rows + 8 NB lo rows); + 64: bank spread
    y_part = NB * 32 * 64;                 // one part of the GEMM 3 operand: 32 NB rows x 32 K
    W = 0;
    F = W + NG * kWGroup;
    G = F + 2 * f_part;
    A = G + 2 * g_part;
    int ax = 2 * a_part > NG * x_stride ? 2 * a_part : NG * x_stride;
    ax = (ax + 127) & ~127;
    Y = A + ax;
    Total = Y + 2 * y_part;
  }
};

// (D3FULL: one barrier per half of the item's samples when NB = 8: the last epilogue of samples 0..3 runs under the UMMAs of 4..7)
// (D2FULL: one barrier per group of 4 modes: the epilogue of a group runs under the UMMAs of the next ones)
enum { MB_CONST = 0, MB_AFULL, MB_D1FULL, MB_XFULL, MB_D2FULL, MB_YFULL = MB_D2FULL + 4, MB_D3FULL, MB_E3DONE = MB_D3FULL + 2, MB_N };
constexpr uint32_t kD2Col = 256;

// v = hi + lo for a pair (hi = bf16(v), lo = bf16(v - hi)), residual through one packed FMA
__device__ __forceinline__ void split2p(float2 v, uint32_t& hi, uint32_t& lo) {
  __nv_bfloat162 h = __floats2bfloat162_rn(v.x, v.y);
  hi = *reinterpret_cast<uint32_t*>(&h);
  const float2 hf = make_float2(__uint_as_float(hi << 16), __uint_as_float(hi & 0xffff0000u));
  const float2 d = __ffma2_rn(hf, make_float2(-1.0f, -1.0f), v);
  __nv_bfloat162 l = __floats2bfloat162_rn(d.x, d.y);
  lo = *reinterpret_cast<uint32_t*>(&l);
}

// c[4 n + q] for a warp-uniform q without dynamic register indexing
__device__ __forceinline__ uint32_t pick4(const uint32_t* c, int n, int q) {
  const uint32_t lo = q & 1 ? c[4 * n + 1] : c[4 * n], hi = q & 1 ? c[4 * n + 3] : c[4 * n + 2];
  return q & 2 ? hi : lo;
}

template <int NB>
__global__ void __launch_bounds__(kMidThreads, 1) fno_mid_tc_kernel(const TcMid2Args a) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ uint64_t bars[MB_N];
  __shared__ uint32_t tmem_s;
  pdl_launch_dependents();
  const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;
  if (warp == kEpiWarps) TR2(0, 0);
  const int H = a.H, R = a.R, kx = blockIdx.x, chunk = blockIdx.y, nchunks = gridDim.y;
  const Mid2Smem L(H, R, NB);
  const int ngroups = (a.B + NB - 1) / NB;

  if (tid == 0) {
    mbar_init(&bars[MB_CONST], 1);
    mbar_init(&bars[MB_AFULL], 1);
    mbar_init(&bars[MB_D1FULL], 1);
    mbar_init(&bars[MB_XFULL], 512);
    for (int g = 0; g < 4; ++g) mbar_init(&bars[MB_D2FULL + g], 1);
    mbar_init(&bars[MB_YFULL], 512);
    mbar_init(&bars[MB_D3FULL], 1);
    mbar_init(&bars[MB_D3FULL + 1], 1);
    mbar_init(&bars[MB_E3DONE], 512);
    fence_mbar_init();
  }
  if (warp == kEpiWarps) {
    tmem_alloc(&tmem_s, 512);
    tmem_relinquish();
  }
  // the K slots of the GEMM 3 operand beyond 2R are never written: they stay zero (their Gs columns are zero too)
  for (int i = tid; i < (2 * L.y_part) / 16; i += kMidThreads) reinterpret_cast<uint4*>(smem + L.Y)[i] = make_uint4(0, 0, 0, 0);
  fence_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (tid == 0) {      // constant operands (packed at weight-commit / plan-build time): requested before the dependency wait
    mbar_arrive_expect_tx(&bars[MB_CONST], (uint32_t)(L.NG * kWGroup + 2 * L.f_part + 2 * L.g_part));
    for (int g = 0; g < L.NG; ++g)
      bulk_g2s(smem + L.W + g * kWGroup, a.Wimg + ((size_t)kx * L.NG + g) * kWGroup, kWGroup, &bars[MB_CONST]);
    bulk_g2s(smem + L.F, a.Fimg, (uint32_t)(2 * L.f_part), &bars[MB_CONST]);
    bulk_g2s(smem + L.G, a.Gimg, (uint32_t)(2 * L.g_part), &bars[MB_CONST]);
  }
  if (warp == kEpiWarps) TR2(0, 1);
  pdl_wait();      // the spectrum tiles come from the previous kernel; Zimg is read by it
  if (warp == kEpiWarps) TR2(0, 2);
  const uint32_t tmem = tmem_s;
  const size_t tile_b = (size_t)H * 128;
  constexpr int CS = 16 * NB * 16;      // K-chunk stride of a group's [P ; Q] images (8 NB hi rows, 8 NB lo rows)

  if (warp == kEpiWarps) {
    // ------------------------------ control warp: tile loads and every tcgen05.mma ------------------------------
    const uint32_t leader = elect_one();
    const uint64_t pol_in = a.l2hint ? l2_evict_first() : 0;
    auto load = [&](int grp) {      // one bulk copy per (sample, hi | lo), from 2 nb lanes
      const int b0 = (a.rev ? ngroups - 1 - grp : grp) * NB, nb = a.B - b0 < NB ? a.B - b0 : NB;
      mbar_arrive_expect_tx_p(leader, &bars[MB_AFULL], (uint32_t)(2 * nb * tile_b));
      __syncwarp();
      const int part = lane >> 3, bl = lane & 7;
      if (lane < 16 && bl < nb)      // (dead after this read: first in line for eviction from the L2)
        bulk_g2s_hp(1u, smem + L.A + part * L.a_part + bl * tile_b, a.A + ((size_t)(kx * 2 + part) * a.B + b0 + bl) * tile_b, (uint32_t)tile_b,
                    &bars[MB_AFULL], pol_in);
      __syncwarp();
    };
    if (chunk < ngroups) load(chunk);
    mbar_wait(&bars[MB_CONST], 0);
    TR2(0, 3);
    const uint32_t aW = smem_u32(smem + L.W), aF = smem_u32(smem + L.F), aG = smem_u32(smem + L.G), aA = smem_u32(smem + L.A),
                   aY = smem_u32(smem + L.Y);
    const int M3 = 2 * H < 64 ? 64 : 2 * H;      // rows of Gs (H = 16: the upper 32 rows are zero)
    constexpr int NHV = NB == 8 ? 2 : 1, NBH = NB / NHV;      // GEMM 3 in halves of the item's samples (NB = 8); samples per half
    const uint32_t idesc1 = idesc_bf16(64, NB * 32) | kIdescBMnMajor, idesc2 = idesc_bf16(128, 8 * NB),
                   idesc3 = idesc_bf16((uint32_t)M3, NBH * 32);
    uint32_t n = 0;
    for (int grp = chunk; grp < ngroups; grp += nchunks, ++n) {
      const uint32_t par = n & 1u;
      // GEMM 1
      mbar_wait(&bars[MB_AFULL], par);
      TR2(n, 4);
      tc_fence_after();
      {
        const uint64_t dF = chunk_desc(aF, 64, 0), dA = smem_desc(aA, 128, (uint32_t)(2 * H * 16));
        uint32_t acc = 0;
#pragma unroll
        for (int term = 0; term < 3; ++term) {      // hi*hi, lo*hi, hi*lo
          const uint32_t fo = term == 1 ? (uint32_t)L.f_part : 0u, ao = term == 2 ? (uint32_t)L.a_part : 0u;
          // (fully unrolled with a uniform guard: with a run-time trip count the descriptors of every instruction went
          //  through general registers and R2UR moves, ~140 cycles per UMMA instead of ~40)
#pragma unroll
          for (int k0 = 0; k0 < 128; k0 += 16) {
            if (k0 < 2 * H) {
              umma_bf16_p(leader, tmem, desc_off(dF, fo + (uint32_t)(k0 >> 3) * 1024u), desc_off(dA, ao + (uint32_t)k0 * 16u), idesc1, acc);
              acc = 1;
            }
          }
        }
        umma_commit_p(leader, &bars[MB_D1FULL]);
      }
      // GEMM 2: four modes per accumulator (rows (mode, o) x columns (P | Q, b, mode'): the blocks mode = mode' are used)
      TR2(n, 5);
      mbar_wait(&bars[MB_XFULL], par);
      TR2(n, 6);
      tc_fence_after();
      for (int g = 0; g < L.NG; ++g) {
        const uint64_t dW = chunk_desc(aW + (uint32_t)g * kWGroup, 128, 0), dX = smem_desc(aA + (uint32_t)(g * L.x_stride), (uint32_t)CS, 128);
        const uint32_t d2 = tmem + kD2Col + (uint32_t)(g * 8 * NB);
#pragma unroll
        for (int term = 0; term < 3; ++term) {
          const uint32_t wo = term == 1 ? 16384u : 0u, xo = term == 2 ? (uint32_t)(8 * NB * 16) : 0u;
#pragma unroll
          for (int k0 = 0; k0 < 64; k0 += 16)
            umma_bf16_p(leader, d2, desc_off(dW, wo + (uint32_t)(k0 >> 3) * 2048u), desc_off(dX, xo + (uint32_t)(k0 >> 3) * CS), idesc2, (term | k0) != 0);
        }
        umma_commit_p(leader, &bars[MB_D2FULL + g]);
      }
      // X is dead once GEMM 2 has completed (the last group's commit covers the earlier ones): the next item's tiles may overwrite it
      TR2(n, 7);
      mbar_wait(&bars[MB_D2FULL + L.NG - 1], par);
      TR2(n, 8);
      if (grp + nchunks < ngroups) load(grp + nchunks);
      // GEMM 3
      mbar_wait(&bars[MB_YFULL], par);
      TR2(n, 9);
      tc_fence_after();
      {
        const uint64_t dG = chunk_desc(aG, (uint32_t)M3, 0);
#pragma unroll
        for (int hv = 0; hv < NHV; ++hv) {      // rows (b, o) of the operand: the second half of the samples starts NBH * 32 rows on
          const uint64_t dY = smem_desc(aY + (uint32_t)(hv * NBH * 512), (uint32_t)(NB * 512), 128);
#pragma unroll
          for (int term = 0; term < 3; ++term) {      // Gs_hi Y_hi, Gs_lo Y_hi, Gs_hi Y_lo
            const uint32_t go = term == 1 ? (uint32_t)L.g_part : 0u, yo = term == 2 ? (uint32_t)L.y_part : 0u;
#pragma unroll
            for (int k0 = 0; k0 < 32; k0 += 16)
              umma_bf16_p(leader, tmem + (uint32_t)(hv * NBH * 32), desc_off(dG, go + (uint32_t)(k0 >> 3) * (uint32_t)(M3 * 16)),
                          desc_off(dY, yo + (uint32_t)(k0 >> 3) * (NB * 512)), idesc3, (term | k0) != 0);
          }
          umma_commit_p(leader, &bars[MB_D3FULL + hv]);
        }
      }
      TR2(n, 10);
      mbar_wait(&bars[MB_E3DONE], par);      // D3 shares its columns with the next item's D1
      TR2(n, 11);
      __syncwarp();
    }
  } else {
    // ------------------------------ epilogue warps: quarter q of the TMEM lanes, share j of the columns ------------------------------
    const int q = warp & 3, j = warp >> 2;
    const uint32_t tq = tmem + ((uint32_t)(q * 32) << 16);
    uint32_t n = 0;
    for (int grp = chunk; grp < ngroups; grp += nchunks, ++n) {
      const uint32_t par = n & 1u;
      const int b0 = (a.rev ? ngroups - 1 - grp : grp) * NB;      // rev: the sample groups from the last one down (tc_plan.cu)
      {
        // ---- E 1: D1 row (copy q of X, r = lane) -> rows of the [P ; Q] operand images of mode r, NB / 4 samples per warp ----
        constexpr int MB = NB > 4 ? 2 : 1;      // warp j: samples j and j + 4 of the item (those below NB)
        mbar_wait(&bars[MB_D1FULL], par);
        if (warp == 0) TR2(n, 12);
        tc_fence_after();
        uint32_t v[MB][32];
#pragma unroll
        for (int t = 0; t < MB; ++t) tmem_ld32(tq + (uint32_t)(((j + 4 * t) % NB) * 32), v[t]);
        tmem_ld_wait();
        if (warp == 0) TR2(n, 18);
        if (lane < R && j < NB) {
          // quarter 0: Xr -> P, K chunks 0..3;  1: Xi -> Q, 0..3;  2: -Xi -> P, 4..7;  3: Xr -> Q, 4..7.   Row: ((P | Q) NB + b) 4 + mode % 4
          uint8_t* img = smem + L.A + (lane >> 2) * L.x_stride + (lane & 3) * 16 + (q >> 1) * 4 * CS + (q & 1) * NB * 64;
#pragma unroll
          for (int t = 0; t < MB; ++t) {
            uint8_t* d = img + (j + 4 * t) * 64;
            uint32_t hi[16], lo[16];
#pragma unroll
            for (int e = 0; e < 16; ++e) split2p(make_float2(__uint_as_float(v[t][2 * e]), __uint_as_float(v[t][2 * e + 1])), hi[e], lo[e]);
#pragma unroll
            for (int c = 0; c < 4; ++c) {
              *reinterpret_cast<uint4*>(d + c * CS) = make_uint4(hi[4 * c], hi[4 * c + 1], hi[4 * c + 2], hi[4 * c + 3]);
              *reinterpret_cast<uint4*>(d + c * CS + 8 * NB * 16) = make_uint4(lo[4 * c], lo[4 * c + 1], lo[4 * c + 2], lo[4 * c + 3]);
            }
          }
        }
        if (warp == 0) TR2(n, 21);
        tc_fence_before();
        fence_async_smem();
        if (warp == 0) TR2(n, 22);
        mbar_arrive(&bars[MB_XFULL]);
        if (warp == 0) TR2(n, 13);
      }
      // ---- E 2: warp (q, j): mode r = 4 j + q, lane = o: its columns (P | Q, b, q) of group j -> (Yr, Yi) pairs of the GEMM 3 operand ----
      if (j < L.NG) {
        mbar_wait(&bars[MB_D2FULL + j], par);
        if (warp == 0) TR2(n, 14);
        tc_fence_after();
        uint32_t c[8 * NB];
#pragma unroll
        for (int t = 0; t < 8 * NB; t += 16) tmem_ld16(tq + kD2Col + (uint32_t)(j * 8 * NB + t), c + t);
        tmem_ld_wait();
        if (4 * j + q < R) {
          uint8_t* dst = smem + L.Y + j * (NB * 512) + lane * 16 + q * 4;
#pragma unroll
          for (int b = 0; b < NB; ++b) {
            uint32_t h, l;
            split2p(make_float2(__uint_as_float(pick4(c, b, q)), __uint_as_float(pick4(c, NB + b, q))), h, l);
            *reinterpret_cast<uint32_t*>(dst + b * 512) = h;
            *reinterpret_cast<uint32_t*>(dst + b * 512 + L.y_part) = l;
          }
        }
      }
      tc_fence_before();
      fence_async_smem();
      mbar_arrive(&bars[MB_YFULL]);
      if (warp == 0) TR2(n, 15);
      // ---- E 3: D3 row = K row (h, re|im) of the Z' images; warp (q, j): rows of quarter q, NB / 4 samples: 32 channels each ----
      {
        constexpr int MB = NB > 4 ? 2 : 1;      // warp j: samples j and j + 4 of the item (those below NB)
        // M = 64: accumulator row 16 q + lane (lanes < 16);  M = 128: row 32 q + lane
        const int row = H == 64 ? 32 * q + lane : 16 * q + lane;
        const bool live = (H == 64 || lane < 16) && row < 2 * H;
        const int h = row >> 1, reim = row & 1;
        const size_t zpart = (size_t)a.KZ * 64;
        const int rts = a.RT == 4 ? 2 : a.RT == 2 ? 1 : 0;      // RT is 1, 2 or 4
        const int k = kx * 2 * a.RT + (h & (a.RT - 1)) * 2 + reim;
#pragma unroll
        for (int t = 0; t < MB; ++t) {      // (NB = 8: round t = half t of GEMM 3: samples 0..3 are stored under the UMMAs of 4..7)
          mbar_wait(&bars[MB_D3FULL + (NB == 8 ? t : 0)], par);
          if (warp == 0 && t == 0) TR2(n, 16);
          tc_fence_after();
          uint32_t z[32];
          tmem_ld32(tq + (uint32_t)(((j + 4 * t) % NB) * 32), z);
          tmem_ld_wait();
          if (warp == 0 && t == 0) TR2(n, 19);
          const int b = b0 + j + 4 * t;
          if (live && j < NB && b < a.B) {
            uint32_t hi[16], lo[16];
#pragma unroll
            for (int e = 0; e < 16; ++e) split2p(make_float2(__uint_as_float(z[2 * e]), __uint_as_float(z[2 * e + 1])), hi[e], lo[e]);
            uint8_t* dst = a.Zimg + (size_t)((uint32_t)(b * H + h) >> rts) * (2 * zpart) + (size_t)k * 16;
#pragma unroll
            for (int c = 0; c < 4; ++c) {      // channel groups
              *reinterpret_cast<uint4*>(dst + (size_t)c * (a.KZ * 16)) = make_uint4(hi[4 * c], hi[4 * c + 1], hi[4 * c + 2], hi[4 * c + 3]);
              *reinterpret_cast<uint4*>(dst + (size_t)c * (a.KZ * 16) + zpart) = make_uint4(lo[4 * c], lo[4 * c + 1], lo[4 * c + 2], lo[4 * c + 3]);
            }
          }
        }
      }
      if (warp == 0) TR2(n, 20);
      tc_fence_before();
      mbar_arrive(&bars[MB_E3DONE]);
      if (warp == 0) TR2(n, 17);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == kEpiWarps) tmem_dealloc(tmem, 512);
}

// mode-major fp32 weights -> [kx][group of 4 modes][hi 16 KB | lo 16 KB] images: rows (mode % 4, o), K = 64:
// k < 32: Re W[i = k][o], else Im W[i = k - 32][o].  (Modes beyond R in the last group stay zero: the buffer is cleared once.)
__global__ void mid2_pack_kernel(const float2* __restrict__ wm, uint8_t* __restrict__ out, int R, int Kc) {
  const int total = R * Kc * 32 * 64, NG = (R + 3) / 4;
  for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < total; t += gridDim.x * blockDim.x) {
    const int k = t & 63, o = (t >> 6) & 31, mm = t >> 11;      // mm = kx * R + r
    const int kx = mm / R, r = mm - kx * R;
    const float2 w = wm[((size_t)(r * Kc + kx) * 32 + (k & 31)) * 32 + o];
    const float v = k < 32 ? w.x : w.y;
    const __nv_bfloat16 h = __float2bfloat16_rn(v);
    const __nv_bfloat16 l = __float2bfloat16_rn(v - __bfloat162float(h));
    uint8_t* img = out + ((size_t)kx * NG + (r >> 2)) * kWGroup + (k >> 3) * 2048 + ((r & 3) * 32 + o) * 16 + (k & 7) * 2;
    *reinterpret_cast<__nv_bfloat16*>(img) = h;
    *reinterpret_cast<__nv_bfloat16*>(img + 16384) = l;
  }
}

}  // namespace

// largest number of samples per item this geometry has shared memory for (0: not covered)
int tc_mid2_samples(int H, int R) {
  if ((H != 16 && H != 32 && H != 64) || R < 1 || R > 16) return 0;
  for (int nb : {8, 4, 2}) {
    if (nb * H > 256) continue;      // the A tiles of an item: at most 64 KB
    if (Mid2Smem(H, R, nb).Total <= 226 * 1024) return nb;
  }
  return 0;
}

// Samples per item for a launch on B samples: the stages of an item run one after the other, so an item's time grows only
// mildly with its sample count (measured on C4: ~6 k / 7.5 k / 10.7 k cycles for 2 / 4 / 8 samples) -- small batches
// are cut into more, smaller items so that they cover the machine.
static int mid2_pick(int H, int R, int B, int Kc, int sms) {
  const int nb_max = tc_mid2_samples(H, R);
  int ctas_per_col = sms / Kc;
  if (ctas_per_col < 1) ctas_per_col = 1;
  int best = 0;
  double best_cost = 0.0;
  for (int nb : {8, 4, 2}) {
    if (nb > nb_max) continue;
    const int groups = (B + nb - 1) / nb, rounds = (groups + ctas_per_col - 1) / ctas_per_col;
    const double cost = rounds * (nb == 8 ? 10.7 : nb == 4 ? 7.5 : 6.0);
    if (best == 0 || cost < best_cost) {
      best = nb;
      best_cost = cost;
    }
  }
  return best;
}

int launch_tc_mid2(const TcMid2Args& a, cudaStream_t st) {
  int sms = 0;
  DLWP_TRY(current_sms(&sms));
  DLWP_REQUIRE(tc_mid2_samples(a.H, a.R) != 0 && a.Kc >= 1 && a.B >= 1, "fno_mid_tc_kernel: unsupported geometry H=%d R=%d", a.H, a.R);
  const int nb = mid2_pick(a.H, a.R, a.B, a.Kc, sms);
  const Mid2Smem L(a.H, a.R, nb);
  const int ngroups = (a.B + nb - 1) / nb;
  int nchunks = sms / a.Kc;
  if (nchunks < 1) nchunks = 1;
  if (nchunks > ngroups) nchunks = ngroups;
  const dim3 grid((unsigned)a.Kc, (unsigned)nchunks);
  void (*kern)(TcMid2Args) = nb == 8 ? fno_mid_tc_kernel<8> : nb == 4 ? fno_mid_tc_kernel<4> : fno_mid_tc_kernel<2>;
  DLWP_TRY(ensure_dynamic_smem((const void*)kern, L.Total));
  DLWP_CUDA(launch_pdl(kern, grid, dim3(kMidThreads), (size_t)L.Total, st, a));
  DLWP_LAUNCHED("fno_mid_tc_kernel");
  return DLWP_OK;
}

int launch_tc_mid2_pack(const float2* wm, uint8_t* Wimg, int R, int Kc, cudaStream_t st) {
  const int total = R * Kc * 32 * 64;
  mid2_pack_kernel<<<(total + 255) / 256, 256, 0, st>>>(wm, Wimg, R, Kc);
  DLWP_LAUNCHED("mid2_pack_kernel");
  return DLWP_OK;
}

// Fs [64 rows][2H], four blocks of 16 rows (r < 16): block 0 = Re X: k = 2h: Re Fh[r, h], k = 2h + 1: -Im Fh[r, h];
// block 1 = Im X: Im Fh, Re Fh; block 2 = -Im X; block 3 = Re X again.
// Gs [max(2H, 64) rows][32]: row 2h = (h, re): k = 2r: Re Gh[h, r], k = 2r + 1: -Im Gh[h, r]; row 2h + 1 = (h, im): Im Gh, Re Gh.
void tc_mid2_pack_bases(const double* Fr, const double* Fi, const double* Gr, const double* Gi, int H, int R, uint8_t* Fimg, uint8_t* Gimg) {
  const int M3 = 2 * H < 64 ? 64 : 2 * H;
  std::vector<double> F((size_t)64 * 2 * H, 0.0), G((size_t)M3 * 32, 0.0);
  for (int r = 0; r < R; ++r)
    for (int h = 0; h < H; ++h) {
      const double fr = Fr[r * H + h], fi = Fi[r * H + h];
      const size_t c = (size_t)2 * h;
      F[(size_t)r * 2 * H + c] = fr;
      F[(size_t)r * 2 * H + c + 1] = -fi;
      F[(size_t)(16 + r) * 2 * H + c] = fi;
      F[(size_t)(16 + r) * 2 * H + c + 1] = fr;
      F[(size_t)(32 + r) * 2 * H + c] = -fi;
      F[(size_t)(32 + r) * 2 * H + c + 1] = -fr;
      F[(size_t)(48 + r) * 2 * H + c] = fr;
      F[(size_t)(48 + r) * 2 * H + c + 1] = -fi;
      G[(size_t)(2 * h) * 32 + 2 * r] = Gr[h * R + r];
      G[(size_t)(2 * h) * 32 + 2 * r + 1] = -Gi[h * R + r];
      G[(size_t)(2 * h + 1) * 32 + 2 * r] = Gi[h * R + r];
      G[(size_t)(2 * h + 1) * 32 + 2 * r + 1] = Gr[h * R + r];
    }
  tc_pack_images_f64(F.data(), 64, 2 * H, 2 * H, 64, 2 * H, Fimg);
  tc_pack_images_f64(G.data(), M3, 32, 32, M3, 32, Gimg);
}

}  // namespace dlwp

#ifdef DLWP_TRACE
extern "C" int dlwp_debug_set_trace_mid2(unsigned long long* p) {
  return cudaMemcpyToSymbol(dlwp::g_trace_mid2, &p, sizeof(p)) == cudaSuccess ? 0 : -2;
}
#endif
