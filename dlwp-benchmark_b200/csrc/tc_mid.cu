// fno_mid_kernel: everything a spectral layer does on the retained columns, in one launch:
//   X[b, r, kx, i] = sum_h  Fh[r, h] * A[b, h, kx, i]            forward truncated DFT along H
//   Y[b, r, kx, o] = sum_i  X[b, r, kx, i] * W[i, o, r, kx]      einsum("bixy,ioxy->boxy"), unet.py:15-17
//   Z[b, h, kx, o] = sum_r  Gh[h, r] * Y[b, r, kx, o]            inverse DFT along H
// A is what the forward W-DFT of the lift / layer kernels left in HBM (planar fp32
// [B][Kc][H][re|im][32]: one contiguous slab per (sample, column), one bulk copy each);
// Z leaves as the split-bf16 operand image the next fno_layer_kernel feeds to tcgen05.mma
// without touching it (layout below).  fp32 FMA arithmetic: 3 x 12 k complex MACs per (sample,
// column) is 0.26 GFMA per layer at the reference sizes -- far below the cost of staging these
// skinny products for the tensor cores.
//
// Work split: CTA = one retained column kx x a contiguous share of the samples; the mode
// weights W[:, :, :, kx] (R x 32 x 32 complex, 96 KB at R = 12) are bulk-copied into SHARED
// memory once per CTA, so they are read from L2 exactly once per CTA and the threads keep their
// registers for deep load prefetch in the DFT phases.  Per pass of up to 4 samples:
//   phase 1  item = (sample, 4 rows r): lane = input channel; A rows stream from shared memory
//            (bulk-copied, double buffered), Fh is read as warp-broadcast 16-byte loads
//   phase 2  warp r: lane = output channel, all samples of the pass at once (one weight load
//            serves 4 samples); X is read as warp-broadcast 16-byte loads
//   phase 3  item = (sample, 8 grid rows h): lane = output channel; split to bf16 hi / lo and
//            stored into the operand image
//
// Z operand image of one 128-pixel tile (RT grid rows), K index k = (kx * RT + h2) * 2 + (re | im):
//   byte offset = part * (KZ * 64) + (c / 8) * (KZ * 16) + k * 16 + (c % 8) * 2,   part 0 = hi, 1 = lo
// i.e. an MN-major UMMA operand (K rows of 8 channels = 16 bytes, four channel groups): the 2 RT rows of one retained
// column are 32 RT contiguous bytes per channel group, so every store of a (tile, column) is made of whole 32-byte
// sectors (the K-major form interleaved the columns' bytes inside every 16-byte row: four partial writes per sector).
#include "tc_common.cuh"
#include "tc_fno.cuh"

namespace dlwp {

using namespace tc;

#ifdef DLWP_TRACE
__device__ unsigned long long* g_trace_mid = nullptr;
#define TRM(p, ev) do { if (g_trace_mid && blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 32 && (p) < 8) g_trace_mid[(p) * 8 + (ev)] = clock64(); } while (0)
#else
#define TRM(p, ev) do {} while (0)
#endif

namespace {


__device__ __forceinline__ void cmac(float2& acc, float gx, float gy, float2 v) {
  acc.x = fmaf(gx, v.x, acc.x);
  acc.x = fmaf(-gy, v.y, acc.x);
  acc.y = fmaf(gx, v.y, acc.y);
  acc.y = fmaf(gy, v.x, acc.y);
}

// RT = grid rows per 128-pixel tile, kNB = samples per pass (4, or 2 when the slabs are large)
template <int RT, int kNB>
__global__ void __launch_bounds__(512, 1) fno_mid_kernel(const TcMidArgs a) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ uint64_t bars[3];
  pdl_launch_dependents();
  TRM(0, 0);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, nw = blockDim.x >> 5;
  const int H = a.H, R = a.R, Rp = (R + 3) & ~3, Kc = a.Kc, B = a.B;
  const int kx = blockIdx.x, chunk = blockIdx.y, nchunks = gridDim.y;
  // this CTA's contiguous share of the samples (balanced to within one), in passes of up to kNB
  const int per = B / nchunks, rem = B % nchunks;
  const int b_lo = chunk * per + (chunk < rem ? chunk : rem), b_hi = b_lo + per + (chunk < rem ? 1 : 0);

  float* sA = reinterpret_cast<float*>(smem);              // [2][kNB][H][re|im][32]
  float2* sX = reinterpret_cast<float2*>(smem) + 2 * kNB * H * 32;   // [kNB][Rp][32]
  float2* sY = sX + kNB * Rp * 32;                          // [kNB][R][32]
  float2* sF = sY + kNB * R * 32;                           // [H][Rp]   Fh transposed, pad rows zero
  float2* sG = sF + H * Rp;                                 // [R][H]    Gh transposed
  float2* sW = sG + R * H;                                  // [R][32 i][32 o] mode weights of this column

  if (tid == 0) {
    mbar_init(&bars[0], 1);
    mbar_init(&bars[1], 1);
    mbar_init(&bars[2], 1);
    fence_mbar_init();
  }
  __syncthreads();
  if (warp == 0) {   // weights of this column: one 8 KB bulk copy per retained row
    if (lane == 0) mbar_arrive_expect_tx(&bars[2], (uint32_t)(R * 8192));
    __syncwarp();
    if (lane < R) bulk_g2s(sW + lane * 1024, a.Wm + (size_t)(lane * Kc + kx) * 1024, 8192, &bars[2]);
  }
  auto issue_load = [&](int b0, int s) {   // warp 0: the (sample, kx) slabs of samples b0.. -> stage s
    const int nb = b_hi - b0 < kNB ? b_hi - b0 : kNB;
    if (lane == 0) mbar_arrive_expect_tx(&bars[s], (uint32_t)(nb * H * 256));
    __syncwarp();
    if (lane < nb)
      bulk_g2s(sA + (size_t)(s * kNB + lane) * H * 64, a.A + ((size_t)(b0 + lane) * Kc + kx) * H * 64, (uint32_t)(H * 256),
               &bars[s]);
  };
  for (int i = tid; i < H * Rp; i += blockDim.x) {
    const int h = i / Rp, r = i - h * Rp;
    sF[i] = r < R ? a.Fh[r * H + h] : make_float2(0.f, 0.f);
  }
  for (int i = tid; i < R * H; i += blockDim.x) {
    const int r = i / H, h = i - r * H;
    sG[i] = a.Gh[h * R + r];
  }
  __syncthreads();
  TRM(0, 1);
  pdl_wait();   // the spectrum below comes from the previous kernel; Zimg is read by it
  TRM(0, 2);
  if (warp == 0 && b_lo < b_hi) issue_load(b_lo, 0);

  uint32_t p = 0;
  for (int b0 = b_lo; b0 < b_hi; b0 += kNB, ++p) {
    const int s = p & 1;
    const int nb = b_hi - b0 < kNB ? b_hi - b0 : kNB;
    if (warp == 0 && b0 + kNB < b_hi) issue_load(b0 + kNB, s ^ 1);
    TRM(p + 1, 0);
    mbar_wait(&bars[s], (p >> 1) & 1u);
    TRM(p + 1, 1);

    // ---- phase 1: forward H-DFT, 4 retained rows per item ----
    const int ngrp = Rp >> 2;
    for (int item = warp; item < nb * ngrp; item += nw) {
      const int bl = item / ngrp, rg = item - bl * ngrp;
      const float* Ab = sA + (size_t)(s * kNB + bl) * H * 64 + lane;
      const float4* F4 = reinterpret_cast<const float4*>(sF) + rg * 2;
      float2 acc[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[j] = make_float2(0.f, 0.f);
#pragma unroll 8
      for (int h = 0; h < H; ++h) {
        const float2 v = make_float2(Ab[h * 64], Ab[h * 64 + 32]);
        const float4 f01 = F4[h * (Rp >> 1)], f23 = F4[h * (Rp >> 1) + 1];
        cmac(acc[0], f01.x, f01.y, v);
        cmac(acc[1], f01.z, f01.w, v);
        cmac(acc[2], f23.x, f23.y, v);
        cmac(acc[3], f23.z, f23.w, v);
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) sX[(bl * Rp + rg * 4 + j) * 32 + lane] = acc[j];
    }
    __syncthreads();
    TRM(p + 1, 2);

    // ---- phase 2: per-mode channel mixing; warp r serves every sample of the pass ----
    if (p == 0) mbar_wait(&bars[2], 0);
    for (int r = warp; r < R; r += nw) {
      const float2* Wr = sW + r * 1024 + lane;
      float2 y[kNB];
#pragma unroll
      for (int bl = 0; bl < kNB; ++bl) y[bl] = make_float2(0.f, 0.f);
#pragma unroll 4
      for (int i2 = 0; i2 < 16; ++i2) {
        const float2 w0 = Wr[(2 * i2) * 32], w1 = Wr[(2 * i2 + 1) * 32];
#pragma unroll
        for (int bl = 0; bl < kNB; ++bl) {
          const float4 x = reinterpret_cast<const float4*>(sX + (bl * Rp + r) * 32)[i2];
          cmac(y[bl], x.x, x.y, w0);
          cmac(y[bl], x.z, x.w, w1);
        }
      }
#pragma unroll
      for (int bl = 0; bl < kNB; ++bl)
        if (bl < nb) sY[(bl * R + r) * 32 + lane] = y[bl];
    }
    __syncthreads();
    TRM(p + 1, 3);

    // ---- phase 3: inverse H-DFT, 8 grid rows per item, split-bf16 operand image out ----
    const int nhg = H >> 3;
    for (int item = warp; item < nb * nhg; item += nw) {
      const int bl = item / nhg, hg = item - bl * nhg;
      const int b = b0 + bl;
      const float2* Yb = sY + bl * R * 32 + lane;
      float2 z[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) z[j] = make_float2(0.f, 0.f);
#pragma unroll 4
      for (int r = 0; r < R; ++r) {
        const float2 y = Yb[r * 32];
        const float4* G4 = reinterpret_cast<const float4*>(sG + r * H + hg * 8);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float4 g = G4[j];
          cmac(z[2 * j], g.x, g.y, y);
          cmac(z[2 * j + 1], g.z, g.w, y);
        }
      }
      uint32_t hi[8], lo[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) split2(z[j].x, z[j].y, hi[j], lo[j]);
      // K rows of this column: k = kx * 2 RT + h2 * 2 + (re|im); lane = channel: 2 bytes per K row (8 lanes fill a 16-byte row)
      const size_t tile0 = ((size_t)b * H + hg * 8) / RT;
      const size_t part = (size_t)a.KZ * 64;
      uint8_t* dst = a.Zimg + tile0 * (2 * part) + (size_t)(lane >> 3) * (a.KZ * 16) + (size_t)(kx * 2 * RT) * 16 + (lane & 7) * 2;
#pragma unroll
      for (int j = 0; j < 8; ++j) {      // grid row hg * 8 + j: tile j / RT, row j % RT of it
        uint8_t* d = dst + (size_t)(j / RT) * (2 * part) + (size_t)((j % RT) * 2) * 16;
        *reinterpret_cast<uint16_t*>(d) = (uint16_t)(hi[j] & 0xffffu);
        *reinterpret_cast<uint16_t*>(d + 16) = (uint16_t)(hi[j] >> 16);
        *reinterpret_cast<uint16_t*>(d + part) = (uint16_t)(lo[j] & 0xffffu);
        *reinterpret_cast<uint16_t*>(d + part + 16) = (uint16_t)(lo[j] >> 16);
      }
    }
    TRM(p + 1, 4);
    // no barrier here: the next pass writes sX only after every warp left phase 2 (second
    // __syncthreads above) and sY only after the next pass's first __syncthreads
  }
}

}  // namespace

static size_t mid_smem(int H, int R, int nb) {
  const int Rp = (R + 3) & ~3;
  return (size_t)(2 * nb * H * 32 + nb * Rp * 32 + nb * R * 32 + H * Rp + R * H + R * 1024) * sizeof(float2);
}
static int mid_samples(int H, int R) { return mid_smem(H, R, kTcMidSamples) <= 220 * 1024 ? kTcMidSamples : 2; }
size_t tc_mid_smem_bytes(int H, int R) { return mid_smem(H, R, mid_samples(H, R)); }

int launch_tc_mid(const TcMidArgs& a, cudaStream_t st) {
  int sms = 0;
  DLWP_TRY(current_sms(&sms));
  DLWP_REQUIRE(a.R >= 1 && a.R <= 16 && a.H % 8 == 0 && a.Kc >= 1, "fno_mid_kernel: unsupported geometry R=%d H=%d", a.R, a.H);
  const int nbs = mid_samples(a.H, a.R);
  const size_t smem = mid_smem(a.H, a.R, nbs);
  DLWP_REQUIRE(smem <= 220 * 1024, "fno_mid_kernel: %zu bytes of shared memory needed", smem);
  // one CTA per SM at most; every CTA takes a contiguous, balanced share of the samples
  int nchunks = sms / a.Kc;
  if (nchunks < 1) nchunks = 1;
  if (nchunks > a.B) nchunks = a.B;
  const dim3 grid((unsigned)a.Kc, (unsigned)nchunks);
  const int threads = 512;   // 16 warps: warp r < R owns mode row r in phase 2, all 16 share phases 1 and 3
  void (*kern)(TcMidArgs) = nullptr;
  int slot = 0;
#define DLWP_MID_PICK(rt, nb, idx) \
  if (a.RT == rt && nbs == nb) {   \
    kern = fno_mid_kernel<rt, nb>; \
    slot = idx;                    \
  }
  DLWP_MID_PICK(1, 4, 0) DLWP_MID_PICK(2, 4, 1) DLWP_MID_PICK(4, 4, 2) DLWP_MID_PICK(1, 2, 3) DLWP_MID_PICK(2, 2, 4)
  DLWP_MID_PICK(4, 2, 5)
#undef DLWP_MID_PICK
  if (!kern) return fail(DLWP_ERR_INVALID, "fno_mid_kernel: %d grid rows per tile", a.RT);
  (void)slot;
  DLWP_TRY(ensure_dynamic_smem((const void*)kern, (int)smem));
  DLWP_CUDA(launch_pdl(kern, grid, dim3(threads), smem, st, a));
  DLWP_LAUNCHED("fno_mid_kernel");
  return DLWP_OK;
}

}  // namespace dlwp

#ifdef DLWP_TRACE
extern "C" int dlwp_debug_set_trace_mid(unsigned long long* p) {
  return cudaMemcpyToSymbol(dlwp::g_trace_mid, &p, sizeof(p)) == cudaSuccess ? 0 : -2;
}
#endif
