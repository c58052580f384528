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
//   GEMM 1  D1[(re|im, r), (b, i)] = Fs[(re|im, r), k] * A[k, (b, i)]      M = 64 (32 rows used), N = 32 NB, K = 2H
//           k = 2h + (re|im).  A arrives from the lift / layer kernels as ready MN-major operand tiles (tc_fno.cuh).
//           Rows (re, r) land in TMEM lanes 0..15, rows (im, r) in lanes 32..47: lane = retained row r.
//   E 1     lane r of the re (im) quarter owns the operand images of ITS mode: it writes Xr (Xi) of every sample as
//           16-byte rows of   P = [Xr | -Xi]   and   Q = [Xi | Xr]   (rows = samples, K = 64), hi and lo parts.
//   GEMM 2  per mode r:  D2[o, (P | Q, b)] = Wm[o, K] * [P ; Q][b, K],  Wm = [Wr^T | Wi^T]       M = 64 (32 used), N = 2 NB, K = 64
//           columns b: Yr = Xr Wr - Xi Wi,  columns NB + b: Yi = Xi Wr + Xr Wi.  All R accumulators side by side in TMEM.
//   E 2     lane o writes (Yr, Yi) pairs into the operand image of GEMM 3: rows (b, o), K = 2r + (re|im).
//   GEMM 3  per 4 samples:  D3[(b, o), (h, re|im)] = Yimg[(b, o), K] * Gs[(h, re|im), K]          M = 128, N = 2H, K = 32
//   E 3     lane (b, o): split to bf16 hi / lo and store into the Z' operand image fno_layer_kernel reads (tc_mid.cu).
//
// X overlays the A tiles (dead once GEMM 1 has completed); the next item's tiles are requested as soon as GEMM 2 has
// read X.  D3 shares its tensor-memory columns with D1.
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

constexpr int kMidThreads = 9 * 32;      // 8 epilogue warps + 1 control warp
constexpr int kWImg = 8192;              // one mode's weight images: hi 4 KB | lo 4 KB (32 rows x 64 K)

struct Mid2Smem {      // byte offsets
  int W, F, G, A, Y, Total, f_part, g_part, a_part, x_stride, y_part;
  __host__ __device__ Mid2Smem(int H, int R, int NB) {
    f_part = 32 * 2 * H * 2;               // one part of Fs: 32 rows x 2H
    g_part = 2 * H * 32 * 2;               // one part of Gs: 2H rows x 32
    a_part = NB * H * 128;                 // hi (or lo) tiles of NB samples
    x_stride = 8 * 4 * NB * 16 + 16;       // one mode's [P ; Q] images (hi rows, lo rows): 8 K chunks x 4 NB rows; + 16: bank spread
    y_part = NB * 32 * 64;                 // one part of the GEMM 3 operand: 32 NB rows x 32 K
    W = 0;
    F = W + R * kWImg;
    G = F + 2 * f_part;
    A = G + 2 * g_part;
    int ax = 2 * a_part > R * x_stride ? 2 * a_part : R * x_stride;
    ax = (ax + 127) & ~127;
    Y = A + ax;
    Total = Y + 2 * y_part + 1024;         // + room behind the last image for the rows an M = 64 operand reads past a 32-row image
  }
};

enum { MB_CONST = 0, MB_AFULL, MB_D1FULL, MB_XFULL, MB_D2FULL, MB_YFULL, MB_D3FULL, MB_E3DONE, MB_N };
constexpr uint32_t kD2Col = 256;

template <int NB>
__global__ void __launch_bounds__(kMidThreads, 1) fno_mid_tc_kernel(const TcMid2Args a) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ uint64_t bars[MB_N];
  __shared__ uint32_t tmem_s;
  pdl_launch_dependents();
  const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;
  if (warp == 8) TR2(0, 0);
  const int H = a.H, R = a.R, kx = blockIdx.x, chunk = blockIdx.y, nchunks = gridDim.y;
  const Mid2Smem L(H, R, NB);
  const int ngroups = (a.B + NB - 1) / NB;

  if (tid == 0) {
    mbar_init(&bars[MB_CONST], 1);
    mbar_init(&bars[MB_AFULL], 1);
    mbar_init(&bars[MB_D1FULL], 1);
    mbar_init(&bars[MB_XFULL], 128);
    mbar_init(&bars[MB_D2FULL], 1);
    mbar_init(&bars[MB_YFULL], 128);
    mbar_init(&bars[MB_D3FULL], 1);
    mbar_init(&bars[MB_E3DONE], 256);
    fence_mbar_init();
  }
  if (warp == 8) {
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
    mbar_arrive_expect_tx(&bars[MB_CONST], (uint32_t)(R * kWImg + 2 * L.f_part + 2 * L.g_part));
    for (int r = 0; r < R; ++r)
      bulk_g2s(smem + L.W + r * kWImg, a.Wimg + ((size_t)kx * R + r) * kWImg, kWImg, &bars[MB_CONST]);
    bulk_g2s(smem + L.F, a.Fimg, (uint32_t)(2 * L.f_part), &bars[MB_CONST]);
    bulk_g2s(smem + L.G, a.Gimg, (uint32_t)(2 * L.g_part), &bars[MB_CONST]);
  }
  if (warp == 8) TR2(0, 1);
  pdl_wait();      // the spectrum tiles come from the previous kernel; Zimg is read by it
  if (warp == 8) TR2(0, 2);
  const uint32_t tmem = tmem_s;
  const size_t tile_b = (size_t)H * 128;

  if (warp == 8) {
    // ------------------------------ control warp: tile loads and every tcgen05.mma ------------------------------
    const uint32_t leader = elect_one();
    auto load = [&](int grp) {
      const int b0 = grp * NB, nb = a.B - b0 < NB ? a.B - b0 : NB;
      const uint32_t bytes = (uint32_t)(nb * tile_b);
      mbar_arrive_expect_tx_p(leader, &bars[MB_AFULL], 2 * bytes);
      bulk_g2s_p(leader, smem + L.A, a.A + ((size_t)(kx * 2) * a.B + b0) * tile_b, bytes, &bars[MB_AFULL]);
      bulk_g2s_p(leader, smem + L.A + L.a_part, a.A + ((size_t)(kx * 2 + 1) * a.B + b0) * tile_b, bytes, &bars[MB_AFULL]);
    };
    if (chunk < ngroups) load(chunk);
    mbar_wait(&bars[MB_CONST], 0);
    TR2(0, 3);
    const uint32_t aW = smem_u32(smem + L.W), aF = smem_u32(smem + L.F), aG = smem_u32(smem + L.G), aA = smem_u32(smem + L.A),
                   aY = smem_u32(smem + L.Y);
    const uint32_t idesc1 = idesc_bf16(64, NB * 32) | kIdescBMnMajor, idesc2 = idesc_bf16(64, 2 * NB),
                   idesc3 = idesc_bf16(128, 2 * H);
    uint32_t n = 0;
    for (int grp = chunk; grp < ngroups; grp += nchunks, ++n) {
      const uint32_t par = n & 1u;
      // GEMM 1
      mbar_wait(&bars[MB_AFULL], par);
      TR2(n, 4);
      tc_fence_after();
      {
        const uint64_t dF = chunk_desc(aF, 32, 0), dA = smem_desc(aA, 128, (uint32_t)(2 * H * 16));
        uint32_t acc = 0;
#pragma unroll
        for (int term = 0; term < 3; ++term) {      // hi*hi, lo*hi, hi*lo
          const uint32_t fo = term == 1 ? (uint32_t)L.f_part : 0u, ao = term == 2 ? (uint32_t)L.a_part : 0u;
          for (int k0 = 0; k0 < 2 * H; k0 += 16) {
            umma_bf16_p(leader, tmem, desc_off(dF, fo + (uint32_t)(k0 >> 3) * 512u), desc_off(dA, ao + (uint32_t)k0 * 16u), idesc1, acc);
            acc = 1;
          }
        }
        umma_commit_p(leader, &bars[MB_D1FULL]);
      }
      // GEMM 2: one accumulator per mode
      TR2(n, 5);
      mbar_wait(&bars[MB_XFULL], par);
      TR2(n, 6);
      tc_fence_after();
      for (int r = 0; r < R; ++r) {
        const uint64_t dW = chunk_desc(aW + (uint32_t)r * kWImg, 32, 0),
                       dX = smem_desc(aA + (uint32_t)(r * L.x_stride), (uint32_t)(4 * NB * 16), 128);
        const uint32_t d2 = tmem + kD2Col + (uint32_t)(r * 2 * NB);
#pragma unroll
        for (int term = 0; term < 3; ++term) {
          const uint32_t wo = term == 1 ? 4096u : 0u, xo = term == 2 ? (uint32_t)(2 * NB * 16) : 0u;
#pragma unroll
          for (int k0 = 0; k0 < 64; k0 += 16)
            umma_bf16_p(leader, d2, desc_off(dW, wo + (uint32_t)(k0 >> 3) * 512u), desc_off(dX, xo + (uint32_t)(k0 >> 3) * (4 * NB * 16)), idesc2,
                        (term | k0) != 0);
        }
      }
      umma_commit_p(leader, &bars[MB_D2FULL]);
      // X is dead once GEMM 2 has completed: the next item's tiles may overwrite it
      TR2(n, 7);
      mbar_wait(&bars[MB_D2FULL], par);
      TR2(n, 8);
      if (grp + nchunks < ngroups) load(grp + nchunks);
      // GEMM 3
      mbar_wait(&bars[MB_YFULL], par);
      TR2(n, 9);
      tc_fence_after();
      for (int gi = 0; gi < NB / 4; ++gi) {
        const uint64_t dY = smem_desc(aY + (uint32_t)(gi * 128 * 16), (uint32_t)(NB * 512), 128), dG = chunk_desc(aG, (uint32_t)(2 * H), 0);
        const uint32_t d3 = tmem + (uint32_t)(gi * 2 * H);
#pragma unroll
        for (int term = 0; term < 3; ++term) {
          const uint32_t yo = term == 1 ? (uint32_t)L.y_part : 0u, go = term == 2 ? (uint32_t)L.g_part : 0u;
#pragma unroll
          for (int k0 = 0; k0 < 32; k0 += 16)
            umma_bf16_p(leader, d3, desc_off(dY, yo + (uint32_t)(k0 >> 3) * (NB * 512)), desc_off(dG, go + (uint32_t)(k0 >> 3) * (uint32_t)(2 * H * 16)),
                        idesc3, (term | k0) != 0);
        }
      }
      umma_commit_p(leader, &bars[MB_D3FULL]);
      TR2(n, 10);
      mbar_wait(&bars[MB_E3DONE], par);      // D3 shares its columns with the next item's D1
      TR2(n, 11);
      __syncwarp();
    }
  } else {
    // ------------------------------ epilogue warps ------------------------------
    const int q = warp & 3, j = warp >> 2;
    const uint32_t tq = tmem + ((uint32_t)(q * 32) << 16);
    const int CS = 4 * NB * 16;      // K-chunk stride of the [P ; Q] images
    uint32_t n = 0;
    for (int grp = chunk; grp < ngroups; grp += nchunks, ++n) {
      const uint32_t par = n & 1u;
      const int b0 = grp * NB;
      if (q < 2) {
        // ---- E 1: D1 row (re | im, r = lane) -> the [P ; Q] operand images of mode r ----
        mbar_wait(&bars[MB_D1FULL], par);
        if (warp == 0) TR2(n, 12);
        tc_fence_after();
        for (int bl = j * (NB / 2); bl < (j + 1) * (NB / 2); ++bl) {
          uint32_t v[32];
          tmem_ld32(tq + (uint32_t)(bl * 32), v);
          tmem_ld_wait();
          if (lane < R) {
            uint32_t hi[16], lo[16];
#pragma unroll
            for (int e = 0; e < 16; ++e) split2(__uint_as_float(v[2 * e]), __uint_as_float(v[2 * e + 1]), hi[e], lo[e]);
            uint8_t* img = smem + L.A + lane * L.x_stride;
            // q = 0 (Xr): P part 0, Q part 1;   q = 1 (Xi): Q part 0, P part 1 negated
            uint8_t* d0 = img + (q == 0 ? bl : NB + bl) * 16;                 // K chunks 0..3
            uint8_t* d1 = img + 4 * CS + (q == 0 ? NB + bl : bl) * 16;        // K chunks 4..7
            const uint32_t sg = q == 0 ? 0u : 0x80008000u;
#pragma unroll
            for (int g = 0; g < 4; ++g) {
              *reinterpret_cast<uint4*>(d0 + g * CS) = make_uint4(hi[4 * g], hi[4 * g + 1], hi[4 * g + 2], hi[4 * g + 3]);
              *reinterpret_cast<uint4*>(d0 + g * CS + 2 * NB * 16) = make_uint4(lo[4 * g], lo[4 * g + 1], lo[4 * g + 2], lo[4 * g + 3]);
              *reinterpret_cast<uint4*>(d1 + g * CS) = make_uint4(hi[4 * g] ^ sg, hi[4 * g + 1] ^ sg, hi[4 * g + 2] ^ sg, hi[4 * g + 3] ^ sg);
              *reinterpret_cast<uint4*>(d1 + g * CS + 2 * NB * 16) =
                  make_uint4(lo[4 * g] ^ sg, lo[4 * g + 1] ^ sg, lo[4 * g + 2] ^ sg, lo[4 * g + 3] ^ sg);
            }
          }
        }
        tc_fence_before();
        fence_async_smem();
        mbar_arrive(&bars[MB_XFULL]);
        if (warp == 0) TR2(n, 13);
        // ---- E 2: D2 row o = 16 q + lane of every mode -> (Yr, Yi) pairs of the GEMM 3 operand ----
        mbar_wait(&bars[MB_D2FULL], par);
        if (warp == 0) TR2(n, 14);
        tc_fence_after();
        const int o = 16 * q + lane;
        for (int r = j; r < R; r += 2) {
          uint32_t c[16];
          if (NB == 8) tmem_ld16(tq + kD2Col + (uint32_t)(r * 16), c);
          else tmem_ld8(tq + kD2Col + (uint32_t)(r * 8), c);
          tmem_ld_wait();
          if (lane < 16) {
            uint8_t* dst = smem + L.Y + (r >> 2) * (NB * 512) + o * 16 + (r & 3) * 4;
#pragma unroll
            for (int b = 0; b < NB; ++b) {
              uint32_t h, l;
              split2(__uint_as_float(c[b]), __uint_as_float(c[NB + b]), h, l);
              *reinterpret_cast<uint32_t*>(dst + b * 512) = h;
              *reinterpret_cast<uint32_t*>(dst + b * 512 + L.y_part) = l;
            }
          }
        }
        tc_fence_before();
        fence_async_smem();
        mbar_arrive(&bars[MB_YFULL]);
        if (warp == 0) TR2(n, 15);
      }
      // ---- E 3: D3 row (sample q of the group, channel lane) -> Z' operand images ----
      mbar_wait(&bars[MB_D3FULL], par);
      if (warp == 0) TR2(n, 16);
      tc_fence_after();
      const size_t zpart = (size_t)a.KZ * 64;
      const int k0 = kx * 2 * a.RT;
      for (int gi = 0; gi < NB / 4; ++gi) {
        const int b = b0 + gi * 4 + q;
        for (int c0 = 0; c0 < H; c0 += 16) {
          uint32_t z[16];
          tmem_ld16(tq + (uint32_t)(gi * 2 * H + j * H + c0), z);
          tmem_ld_wait();
          if (b < a.B) {
            uint32_t hi[8], lo[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) split2(__uint_as_float(z[2 * e]), __uint_as_float(z[2 * e + 1]), hi[e], lo[e]);
            const int h0 = (j * H + c0) >> 1;      // 8 consecutive grid rows
            const size_t tile0 = ((size_t)b * H + h0) / a.RT;
            uint8_t* dst = a.Zimg + tile0 * (2 * zpart) + (size_t)(k0 >> 3) * 512 + lane * 16 + (k0 & 7) * 2;
            if (a.RT == 1) {
#pragma unroll
              for (int t = 0; t < 8; ++t) {
                *reinterpret_cast<uint32_t*>(dst + (size_t)t * (2 * zpart)) = hi[t];
                *reinterpret_cast<uint32_t*>(dst + (size_t)t * (2 * zpart) + zpart) = lo[t];
              }
            } else if (a.RT == 2) {
#pragma unroll
              for (int t = 0; t < 4; ++t) {
                *reinterpret_cast<uint2*>(dst + (size_t)t * (2 * zpart)) = make_uint2(hi[2 * t], hi[2 * t + 1]);
                *reinterpret_cast<uint2*>(dst + (size_t)t * (2 * zpart) + zpart) = make_uint2(lo[2 * t], lo[2 * t + 1]);
              }
            } else {
#pragma unroll
              for (int t = 0; t < 2; ++t) {
                *reinterpret_cast<uint4*>(dst + (size_t)t * (2 * zpart)) = make_uint4(hi[4 * t], hi[4 * t + 1], hi[4 * t + 2], hi[4 * t + 3]);
                *reinterpret_cast<uint4*>(dst + (size_t)t * (2 * zpart) + zpart) = make_uint4(lo[4 * t], lo[4 * t + 1], lo[4 * t + 2], lo[4 * t + 3]);
              }
            }
          }
        }
      }
      tc_fence_before();
      mbar_arrive(&bars[MB_E3DONE]);
      if (warp == 0) TR2(n, 17);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 8) tmem_dealloc(tmem, 512);
}

// mode-major fp32 weights -> [kx][r][hi | lo] images: rows o (32), K = 64: k < 32: Re W[i = k][o], else Im W[i = k - 32][o]
__global__ void mid2_pack_kernel(const float2* __restrict__ wm, uint8_t* __restrict__ out, int R, int Kc) {
  const int total = R * Kc * 32 * 64;
  for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < total; t += gridDim.x * blockDim.x) {
    const int k = t & 63, o = (t >> 6) & 31, m = t >> 11;      // m = kx * R + r
    const int kx = m / R, r = m - kx * R;
    const float2 w = wm[((size_t)(r * Kc + kx) * 32 + (k & 31)) * 32 + o];
    const float v = k < 32 ? w.x : w.y;
    const __nv_bfloat16 h = __float2bfloat16_rn(v);
    const __nv_bfloat16 l = __float2bfloat16_rn(v - __bfloat162float(h));
    uint8_t* img = out + (size_t)m * kWImg + (k >> 3) * 512 + o * 16 + (k & 7) * 2;
    *reinterpret_cast<__nv_bfloat16*>(img) = h;
    *reinterpret_cast<__nv_bfloat16*>(img + 4096) = l;
  }
}

}  // namespace

int tc_mid2_samples(int H, int R) {
  if (H < 16 || H > 64 || H % 16 != 0 || R < 1 || R > 16) return 0;
  for (int nb : {8, 4}) {
    if (nb * 32 > 256 || (nb / 4) * 2 * H > 256) continue;      // D1 / D3 columns
    if (Mid2Smem(H, R, nb).Total <= 226 * 1024) return nb;
  }
  return 0;
}

int launch_tc_mid2(const TcMid2Args& a, cudaStream_t st) {
  const int nb = tc_mid2_samples(a.H, a.R);
  DLWP_REQUIRE(nb != 0 && a.Kc >= 1 && a.B >= 1, "fno_mid_tc_kernel: unsupported geometry H=%d R=%d", a.H, a.R);
  int sms = 0;
  DLWP_TRY(current_sms(&sms));
  const Mid2Smem L(a.H, a.R, nb);
  const int ngroups = (a.B + nb - 1) / nb;
  int nchunks = sms / a.Kc;
  if (nchunks < 1) nchunks = 1;
  if (nchunks > ngroups) nchunks = ngroups;
  const dim3 grid((unsigned)a.Kc, (unsigned)nchunks);
  void (*kern)(TcMid2Args) = nb == 8 ? fno_mid_tc_kernel<8> : fno_mid_tc_kernel<4>;
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

// Fs [32 rows][2H]: row r = (re, r): k = 2h: Re Fh[r, h], k = 2h + 1: -Im Fh[r, h]; row 16 + r = (im, r): Im Fh, Re Fh.
// Gs [2H rows][32]: row 2h = (h, re): k = 2r: Re Gh[h, r], k = 2r + 1: -Im Gh[h, r]; row 2h + 1 = (h, im): Im Gh, Re Gh.
void tc_mid2_pack_bases(const double* Fr, const double* Fi, const double* Gr, const double* Gi, int H, int R, uint8_t* Fimg, uint8_t* Gimg) {
  std::vector<double> F((size_t)32 * 2 * H, 0.0), G((size_t)2 * H * 32, 0.0);
  for (int r = 0; r < R; ++r)
    for (int h = 0; h < H; ++h) {
      F[(size_t)r * 2 * H + 2 * h] = Fr[r * H + h];
      F[(size_t)r * 2 * H + 2 * h + 1] = -Fi[r * H + h];
      F[(size_t)(16 + r) * 2 * H + 2 * h] = Fi[r * H + h];
      F[(size_t)(16 + r) * 2 * H + 2 * h + 1] = Fr[r * H + h];
      G[(size_t)(2 * h) * 32 + 2 * r] = Gr[h * R + r];
      G[(size_t)(2 * h) * 32 + 2 * r + 1] = -Gi[h * R + r];
      G[(size_t)(2 * h + 1) * 32 + 2 * r] = Gi[h * R + r];
      G[(size_t)(2 * h + 1) * 32 + 2 * r + 1] = Gr[h * R + r];
    }
  tc_pack_images_f64(F.data(), 32, 2 * H, 2 * H, 32, 2 * H, Fimg);
  tc_pack_images_f64(G.data(), 2 * H, 32, 32, 2 * H, 32, Gimg);
}

}  // namespace dlwp

#ifdef DLWP_TRACE
extern "C" int dlwp_debug_set_trace_mid2(unsigned long long* p) {
  return cudaMemcpyToSymbol(dlwp::g_trace_mid2, &p, sizeof(p)) == cudaSuccess ? 0 : -2;
}
#endif
