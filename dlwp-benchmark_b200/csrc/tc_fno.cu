// tcgen05 kernels of the FNO / TFNO step (see tc_fno.cuh for the data layout and the split).
#include "tc_fno.cuh"

#include <cstring>

#include "tc_common.cuh"

namespace dlwp {

using namespace tc;

namespace {

constexpr int kImg = kTcImgBytes;            // 8192: one bf16 part of a [128 x 32] tile
constexpr int kTileBytes = 2 * kImg;         // hi + lo
constexpr int kLboXT = 528;                  // chunk stride of the transposed tile: 512 + 16 keeps the
                                             // 4-byte transposing stores of a warp on 32 distinct banks
constexpr int kXTBytes = 16 * kLboXT;        // one bf16 part of the [32 ch x 128 px] transposed tile
constexpr int kStageLd = 33;
constexpr int kStageBytes = 2 * 32 * kStageLd * 4;

// exact-erf GELU (F.gelu default) through Abramowitz-Stegun 7.1.26 (|erf error| <= 1.5e-7):
// two MUFU ops and ~13 FMA-pipe ops per element instead of erff's branches
__device__ __forceinline__ float gelu_fast(float x) {
  const float z = fabsf(x) * 0.70710678118654752f;
  float t, e;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(fmaf(0.3275911f, z, 1.0f)));
  float p = fmaf(1.061405429f, t, -1.453152027f);
  p = fmaf(p, t, 1.421413741f);
  p = fmaf(p, t, -0.284496736f);
  p = fmaf(p, t, 0.254829592f);
  p *= t;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(z * z * -1.4426950408889634f));
  const float erf_abs = fmaf(-p, e, 1.0f);
  const float hx = 0.5f * x;
  return fmaf(hx, copysignf(erf_abs, x), hx);
}

// ---- common tail of the lift and layer kernels ------------------------------------------------
// v[32] = the 32 output channels of this thread's pixel (tile row = tid).  Writes the tile
// image to HBM and, when `Aout` is set, the forward truncated W-DFT of the tile:
//   D3[(part, h2, kx'), c] = sum_{px} F''[(part, h2, kx'), px] * v[px, c]
// as one more M=64 UMMA whose B operand is the tile transposed through shared memory.
__device__ __forceinline__ void finish_tile(const float (&v)[32], int tid, int tile, const TcGeom& g, uint8_t* y_img,
                                            float2* Aout, uint8_t* sF, uint8_t* sXT, float* stage, uint32_t tmem_d3,
                                            uint64_t* bar, uint32_t& phase) {
  {
    uint32_t hi[16], lo[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) split2(v[2 * j], v[2 * j + 1], hi[j], lo[j]);
    uint8_t* dst = y_img + (size_t)tile * kTileBytes + tid * 16;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      *reinterpret_cast<uint4*>(dst + q * 2048) = make_uint4(hi[4 * q], hi[4 * q + 1], hi[4 * q + 2], hi[4 * q + 3]);
      *reinterpret_cast<uint4*>(dst + kImg + q * 2048) =
          make_uint4(lo[4 * q], lo[4 * q + 1], lo[4 * q + 2], lo[4 * q + 3]);
    }
  }
  if (Aout == nullptr) return;

  // transposed tile [channel][pixel]: lanes pair up so that each store holds two adjacent pixels
  const int odd = tid & 1, me = tid & ~1;
  uint8_t* xt = sXT + (me >> 3) * kLboXT + (me & 7) * 2;
#pragma unroll
  for (int j = 0; j < 16; ++j) {
    const float send = odd ? v[2 * j] : v[2 * j + 1];
    const float recv = __shfl_xor_sync(0xffffffffu, send, 1);
    const float a = odd ? recv : v[2 * j];
    const float b = odd ? v[2 * j + 1] : recv;
    uint32_t h, l;
    split2(a, b, h, l);
    const int row = 2 * j + odd;
    *reinterpret_cast<uint32_t*>(xt + row * 16) = h;
    *reinterpret_cast<uint32_t*>(xt + kXTBytes + row * 16) = l;
  }
  fence_async_smem();
  tc_fence_before();
  __syncthreads();
  if (tid == 0) {
    tc_fence_after();
    const uint32_t idesc = idesc_bf16(64, 32);
    const uint32_t aF = smem_u32(sF), aXT = smem_u32(sXT);
#pragma unroll
    for (int part = 0; part < 2; ++part)
#pragma unroll
      for (int k0 = 0; k0 < 128; k0 += 16)
        umma_bf16(tmem_d3, chunk_desc(aF, 64, k0), smem_desc(aXT + part * kXTBytes + (k0 >> 3) * kLboXT, kLboXT, 128),
                  idesc, (part | k0) != 0);
    umma_commit(bar);
  }
  mbar_wait(bar, phase);
  phase ^= 1;
  tc_fence_after();

  // an M=64 accumulator keeps row r in TMEM lane 32 * (r / 16) + r % 16: warp w holds rows 16w..16w+15
  const int warp = tid >> 5, lane = tid & 31;
  uint32_t r[32];
  tmem_ld32(tmem_d3 + ((uint32_t)(warp * 32) << 16), r);
  tmem_ld_wait();
  if (lane < 16) {
    float* s = stage + ((warp >> 1) * 32 + (warp & 1) * 16 + lane) * kStageLd;
#pragma unroll
    for (int c = 0; c < 32; ++c) s[c] = __uint_as_float(r[c]);
  }
  __syncthreads();
  const int n = g.RT * g.Kc * 32;
  for (int e = tid; e < n; e += 128) {
    const int c = e & 31, q = e >> 5;
    const int kx = q % g.Kc, h2 = q / g.Kc;
    const int idx = h2 * g.KXP + 2 * kx;
    const float re = stage[idx * kStageLd + c] + stage[(32 + idx) * kStageLd + c];
    const float im = stage[(idx + 1) * kStageLd + c] + stage[(33 + idx) * kStageLd + c];
    Aout[((size_t)(tile * g.RT + h2) * g.Kc + kx) * 32 + c] = make_float2(re, im);
  }
}

}  // namespace

// =============================================================================================
// fno_layer_kernel: y = act( irfft_W(Z) + skip(x) + bias ) on 128-pixel tiles
//   D[128 px, 32 ch] = G'[128, 32] * Z'^T  +  X[128, 32] * Ws^T      (12 UMMAs, split-bf16)
// G' is block-diagonal over the grid rows of the tile; Z' holds (Re, Im) of the retained
// columns of those rows for every channel.
// =============================================================================================
namespace {
struct LayerSmem {
  static constexpr int G = 0, F = G + 2 * kImg, Ws = F + 64 * 128 * 2, X = Ws + 4096, Z = X + kTileBytes,
                       XT = Z + 4096, Stage = XT + 2 * kXTBytes, Bias = Stage + kStageBytes, Total = Bias + 128;
};
}  // namespace

__global__ void __launch_bounds__(128) fno_layer_kernel(const TcLayerArgs a) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ uint64_t bar_const, bar_x, bar_mma;
  __shared__ uint32_t tmem_s;
  uint8_t *sG = smem + LayerSmem::G, *sF = smem + LayerSmem::F, *sWs = smem + LayerSmem::Ws, *sX = smem + LayerSmem::X,
          *sZ = smem + LayerSmem::Z, *sXT = smem + LayerSmem::XT;
  float* stage = reinterpret_cast<float*>(smem + LayerSmem::Stage);
  float* sBias = reinterpret_cast<float*>(smem + LayerSmem::Bias);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const TcGeom g = a.g;

  if (tid == 0) {
    mbar_init(&bar_const, 1);
    mbar_init(&bar_x, 1);
    mbar_init(&bar_mma, 1);
    fence_mbar_init();
  }
  if (warp == 0) {
    tmem_alloc(&tmem_s, 64);
    tmem_relinquish();
  }
  for (int i = tid; i < 1024; i += 128) reinterpret_cast<uint32_t*>(sZ)[i] = 0u;   // pad columns stay zero
  if (tid < 32) sBias[tid] = a.bias ? a.bias[tid] : 0.f;
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_d = tmem_s, tmem_d3 = tmem_s + 32;
  if (tid == 0) {
    mbar_arrive_expect_tx(&bar_const, 2 * kImg + 64 * 128 * 2 + 4096);
    bulk_g2s(sG, a.Gimg, 2 * kImg, &bar_const);
    bulk_g2s(sF, a.Fimg, 64 * 128 * 2, &bar_const);
    bulk_g2s(sWs, a.Wsimg, 4096, &bar_const);
  }
  mbar_wait(&bar_const, 0);
  uint32_t ph_x = 0, ph_mma = 0;

  for (int tile = blockIdx.x; tile < g.ntiles; tile += gridDim.x) {
    if (tid == 0) {
      mbar_arrive_expect_tx(&bar_x, kTileBytes);
      bulk_g2s(sX, a.x_img + (size_t)tile * kTileBytes, kTileBytes, &bar_x);
    }
    // Z'[ch, (h2, kx')] <- Z[(row, kx), ch] (fp32 complex), split into bf16 hi / lo
    for (int q = warp; q < g.RT * g.Kc; q += 4) {
      const int h2 = q / g.Kc, kx = q - h2 * g.Kc;
      const float2 z = a.Z[((size_t)(tile * g.RT + h2) * g.Kc + kx) * 32 + lane];
      uint32_t h, l;
      split2(z.x, z.y, h, l);
      const int k = h2 * g.KXP + 2 * kx;
      const int off = (k >> 3) * 512 + lane * 16 + (k & 7) * 2;
      *reinterpret_cast<uint32_t*>(sZ + off) = h;
      *reinterpret_cast<uint32_t*>(sZ + 2048 + off) = l;
    }
    fence_async_smem();
    mbar_wait(&bar_x, ph_x);
    ph_x ^= 1;
    tc_fence_before();
    __syncthreads();
    if (tid == 0) {
      tc_fence_after();
      const uint32_t idesc = idesc_bf16(128, 32);
      const uint32_t aG = smem_u32(sG), aZ = smem_u32(sZ), aX = smem_u32(sX), aW = smem_u32(sWs);
      uint32_t acc = 0;
#pragma unroll
      for (int term = 0; term < 3; ++term) {      // hi*hi, lo*hi, hi*lo
        const uint32_t ao = term == 1 ? kImg : 0, bo = term == 2 ? 2048 : 0;
#pragma unroll
        for (int k0 = 0; k0 < 32; k0 += 16) {
          umma_bf16(tmem_d, chunk_desc(aG + ao, 128, k0), chunk_desc(aZ + bo, 32, k0), idesc, acc);
          acc = 1;
          umma_bf16(tmem_d, chunk_desc(aX + ao, 128, k0), chunk_desc(aW + bo, 32, k0), idesc, acc);
        }
      }
      umma_commit(&bar_mma);
    }
    mbar_wait(&bar_mma, ph_mma);
    ph_mma ^= 1;
    tc_fence_after();

    float v[32];
    {
      uint32_t r[32];
      tmem_ld32(tmem_d + ((uint32_t)(warp * 32) << 16), r);
      tmem_ld_wait();
#pragma unroll
      for (int c = 0; c < 32; ++c) {
        const float t = __uint_as_float(r[c]) + sBias[c];
        v[c] = a.gelu ? gelu_fast(t) : t;
      }
    }
    finish_tile(v, tid, tile, g, a.y_img, a.Aout, sF, sXT, stage, tmem_d3, &bar_mma, ph_mma);
    tc_fence_before();
    __syncthreads();
  }
  if (warp == 0) tmem_dealloc(tmem_s, 64);
}

int launch_tc_layer(const TcLayerArgs& a, cudaStream_t st) {
  static int sms = 0;
  if (!sms) {
    int dev = 0;
    DLWP_CUDA(cudaGetDevice(&dev));
    DLWP_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    DLWP_CUDA(cudaFuncSetAttribute(fno_layer_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, LayerSmem::Total));
  }
  const int grid = a.g.ntiles < 2 * sms ? a.g.ntiles : 2 * sms;
  fno_layer_kernel<<<grid, 128, LayerSmem::Total, st>>>(a);
  DLWP_LAUNCHED("fno_layer_kernel");
  return DLWP_OK;
}

// =============================================================================================
// Fused 2-layer pointwise MLP on one 128-pixel tile (neuralop MLP: 1x1 conv, GELU, 1x1 conv)
//   D1[128, 128] = A1 * W1[pass]^T              (hidden columns, 128 at a time, in TMEM)
//   A2 = split(GELU(D1 + b1))                   (32 columns at a time, double-buffered in SMEM)
//   D2[128, N2] += A2 * W2[:, cols]^T
// The hidden activations never leave the SM.
// =============================================================================================
namespace {

struct MlpBars {
  uint64_t* d1;
  uint64_t* a2;        // [2]
  uint32_t ph_d1;
  uint32_t ph_a2[2];
};

template <int K1, int N2, bool kBias1>
__device__ __forceinline__ void mlp_tile(int tid, uint32_t aA1, uint32_t aW1, int hid, const float* sB1, uint32_t aW2,
                                         uint8_t* sA2, uint32_t tmem_d1, uint32_t tmem_d2, MlpBars& bars) {
  const int warp = tid >> 5;
  const uint32_t a1_lo = aA1 + 128 * K1 * 2, w1_lo = aW1 + (uint32_t)hid * K1 * 2, w2_lo = aW2 + (uint32_t)N2 * hid * 2;
  const uint32_t aA2 = smem_u32(sA2);
  int cg = 0;     // chunk counter within this tile
  for (int p = 0; p < hid / 128; ++p) {
    if (tid == 0) {
      tc_fence_after();
      const uint32_t idesc = idesc_bf16(128, 128);
      uint32_t acc = 0;
#pragma unroll
      for (int term = 0; term < 3; ++term) {
        const uint32_t a = term == 1 ? a1_lo : aA1, w = term == 2 ? w1_lo : aW1;
#pragma unroll
        for (int k0 = 0; k0 < K1; k0 += 16) {
          umma_bf16(tmem_d1, chunk_desc(a, 128, k0),
                    smem_desc(w + (uint32_t)(k0 >> 3) * hid * 16 + (uint32_t)p * 128 * 16, (uint32_t)hid * 16, 128), idesc,
                    acc);
          acc = 1;
        }
      }
      umma_commit(bars.d1);
    }
    mbar_wait(bars.d1, bars.ph_d1);
    bars.ph_d1 ^= 1;
    tc_fence_after();
#pragma unroll 1
    for (int j = 0; j < 4; ++j, ++cg) {
      const int buf = cg & 1;
      uint32_t r[32];
      tmem_ld32(tmem_d1 + ((uint32_t)(warp * 32) << 16) + 32 * j, r);
      tmem_ld_wait();
      if (cg >= 2) {      // the UMMAs that read this A2 buffer two chunks ago must be done
        mbar_wait(&bars.a2[buf], bars.ph_a2[buf]);
        bars.ph_a2[buf] ^= 1;
      }
      uint8_t* dst = sA2 + buf * kTileBytes + tid * 16;
      const int hcol0 = p * 128 + 32 * j;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        uint32_t h[4], l[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          float f0 = __uint_as_float(r[8 * q + 2 * e]), f1 = __uint_as_float(r[8 * q + 2 * e + 1]);
          if (kBias1) {
            f0 += sB1[hcol0 + 8 * q + 2 * e];
            f1 += sB1[hcol0 + 8 * q + 2 * e + 1];
          }
          split2(gelu_fast(f0), gelu_fast(f1), h[e], l[e]);
        }
        *reinterpret_cast<uint4*>(dst + q * 2048) = make_uint4(h[0], h[1], h[2], h[3]);
        *reinterpret_cast<uint4*>(dst + kImg + q * 2048) = make_uint4(l[0], l[1], l[2], l[3]);
      }
      fence_async_smem();
      tc_fence_before();
      __syncthreads();
      if (tid == 0) {
        tc_fence_after();
        const uint32_t idesc2 = idesc_bf16(128, N2);
        const uint32_t a2 = aA2 + buf * kTileBytes;
#pragma unroll
        for (int term = 0; term < 3; ++term) {
          const uint32_t a = term == 1 ? a2 + kImg : a2, w = term == 2 ? w2_lo : aW2;
#pragma unroll
          for (int kk = 0; kk < 32; kk += 16)
            umma_bf16(tmem_d2, chunk_desc(a, 128, kk), chunk_desc(w, N2, hcol0 + kk), idesc2, (cg | term | kk) != 0);
        }
        umma_commit(&bars.a2[buf]);
      }
    }
  }
  // drain: both buffers' last commits (the second one covers every UMMA of this tile)
  const int last = cg - 1;
  mbar_wait(&bars.a2[(last - 1) & 1], bars.ph_a2[(last - 1) & 1]);
  bars.ph_a2[(last - 1) & 1] ^= 1;
  mbar_wait(&bars.a2[last & 1], bars.ph_a2[last & 1]);
  bars.ph_a2[last & 1] ^= 1;
  tc_fence_after();
}

}  // namespace

// =============================================================================================
// fno_proj_kernel: out = last + W2 * GELU(W1 * x + b1) + b2, tile image in, NCHW frame out
// =============================================================================================
namespace {
struct ProjSmem {   // for hid = 256
  static constexpr int X = 0, W1 = X + kTileBytes, W2 = W1 + 256 * 32 * 4, A2 = W2 + 16 * 256 * 4, B1 = A2 + 2 * kTileBytes,
                       B2 = B1 + 1024, Total = B2 + 64;
};
}  // namespace

__global__ void __launch_bounds__(128) fno_proj_kernel(const TcProjArgs a) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ uint64_t bar_const, bar_x, bar_d1, bar_a2[2];
  __shared__ uint32_t tmem_s;
  uint8_t *sX = smem + ProjSmem::X, *sW1 = smem + ProjSmem::W1, *sW2 = smem + ProjSmem::W2, *sA2 = smem + ProjSmem::A2;
  float* sB1 = reinterpret_cast<float*>(smem + ProjSmem::B1);
  float* sB2 = reinterpret_cast<float*>(smem + ProjSmem::B2);
  const int tid = threadIdx.x, warp = tid >> 5;
  const TcGeom g = a.g;
  const int hid = a.hid;

  if (tid == 0) {
    mbar_init(&bar_const, 1);
    mbar_init(&bar_x, 1);
    mbar_init(&bar_d1, 1);
    mbar_init(&bar_a2[0], 1);
    mbar_init(&bar_a2[1], 1);
    fence_mbar_init();
  }
  if (warp == 0) {
    tmem_alloc(&tmem_s, 256);
    tmem_relinquish();
  }
  for (int i = tid; i < hid; i += 128) sB1[i] = a.b1[i];
  if (tid < 16) sB2[tid] = tid < a.Cg ? a.b2[tid] : 0.f;
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_d1 = tmem_s, tmem_d2 = tmem_s + 128;
  const uint32_t w1_bytes = (uint32_t)hid * 32 * 4, w2_bytes = 16u * hid * 4;
  if (tid == 0) {
    mbar_arrive_expect_tx(&bar_const, w1_bytes + w2_bytes);
    bulk_g2s(sW1, a.W1img, w1_bytes, &bar_const);
    bulk_g2s(sW2, a.W2img, w2_bytes, &bar_const);
  }
  mbar_wait(&bar_const, 0);
  uint32_t ph_x = 0;
  MlpBars bars{&bar_d1, bar_a2, 0u, {0u, 0u}};

  for (int tile = blockIdx.x; tile < g.ntiles; tile += gridDim.x) {
    if (tid == 0) {
      mbar_arrive_expect_tx(&bar_x, kTileBytes);
      bulk_g2s(sX, a.x_img + (size_t)tile * kTileBytes, kTileBytes, &bar_x);
    }
    mbar_wait(&bar_x, ph_x);
    ph_x ^= 1;
    mlp_tile<32, 16, true>(tid, smem_u32(sX), smem_u32(sW1), hid, sB1, smem_u32(sW2), sA2, tmem_d1, tmem_d2, bars);

    uint32_t r[16];
    tmem_ld16(tmem_d2 + ((uint32_t)(warp * 32) << 16), r);
    tmem_ld_wait();
    const long long p = (long long)tile * kTcTile + tid;
    const long long b = p / g.HW, pix = p - b * g.HW;
    float* o = a.out + b * a.out_bstride + pix;
    const float* lastp = a.last ? a.last + b * a.last_bstride + pix : nullptr;
#pragma unroll
    for (int c = 0; c < 16; ++c)
      if (c < a.Cg) {
        float val = __uint_as_float(r[c]) + sB2[c];
        if (lastp) val += lastp[(long long)c * g.HW];
        o[(long long)c * g.HW] = val;
      }
    tc_fence_before();
    __syncthreads();
  }
  if (warp == 0) tmem_dealloc(tmem_s, 256);
}

int launch_tc_proj(const TcProjArgs& a, cudaStream_t st) {
  static int sms = 0;
  if (!sms) {
    int dev = 0;
    DLWP_CUDA(cudaGetDevice(&dev));
    DLWP_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    DLWP_CUDA(cudaFuncSetAttribute(fno_proj_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, ProjSmem::Total));
  }
  DLWP_REQUIRE(a.hid == 128 || a.hid == 256, "tensor-core projection needs width 128 or 256, got %d", a.hid);
  const int grid = a.g.ntiles < 2 * sms ? a.g.ntiles : 2 * sms;
  fno_proj_kernel<<<grid, 128, ProjSmem::Total, st>>>(a);
  DLWP_LAUNCHED("fno_proj_kernel");
  return DLWP_OK;
}

// =============================================================================================
// fno_lift_kernel: x_t gathered from the NCHW sources -> lifting MLP -> tile image + W-DFT
// =============================================================================================
namespace {
struct LiftSmem {   // for hid = 256; XT / Stage reuse the A2 buffers once the MLP of the tile is done
  static constexpr int A1 = 0, W1 = A1 + 128 * 16 * 4, W2 = W1 + 256 * 16 * 4, A2 = W2 + 32 * 256 * 4, XT = A2,
                       Stage = XT + 2 * kXTBytes, F = A2 + 2 * kTileBytes, B2 = F + 64 * 128 * 2, Total = B2 + 128;
  static_assert(Stage + kStageBytes <= F, "transposed tile + stage must fit in the A2 buffers");
};
}  // namespace

__global__ void __launch_bounds__(128) fno_lift_kernel(const TcLiftArgs a) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ uint64_t bar_const, bar_d1, bar_a2[2], bar_d3;
  __shared__ uint32_t tmem_s;
  uint8_t *sA1 = smem + LiftSmem::A1, *sW1 = smem + LiftSmem::W1, *sW2 = smem + LiftSmem::W2, *sA2 = smem + LiftSmem::A2,
          *sXT = smem + LiftSmem::XT, *sF = smem + LiftSmem::F;
  float* stage = reinterpret_cast<float*>(smem + LiftSmem::Stage);
  float* sB2 = reinterpret_cast<float*>(smem + LiftSmem::B2);
  const int tid = threadIdx.x, warp = tid >> 5;
  const TcGeom g = a.g;
  const int hid = a.hid;

  if (tid == 0) {
    mbar_init(&bar_const, 1);
    mbar_init(&bar_d1, 1);
    mbar_init(&bar_a2[0], 1);
    mbar_init(&bar_a2[1], 1);
    mbar_init(&bar_d3, 1);
    fence_mbar_init();
  }
  if (warp == 0) {
    tmem_alloc(&tmem_s, 256);
    tmem_relinquish();
  }
  if (tid < 32) sB2[tid] = a.b2[tid];
  {
    // constant columns of A1: column Cin is 1 (it multiplies the bias row of W1), the rest 0
    uint4* z = reinterpret_cast<uint4*>(sA1);
    for (int i = tid; i < 128 * 16 * 4 / 16; i += 128) z[i] = make_uint4(0, 0, 0, 0);
  }
  __syncthreads();
  *reinterpret_cast<uint16_t*>(sA1 + (a.Cin >> 3) * 2048 + tid * 16 + (a.Cin & 7) * 2) = 0x3F80;   // bf16(1.0), hi part
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_d1 = tmem_s, tmem_d2 = tmem_s + 128, tmem_d3 = tmem_s + 160;
  const uint32_t w1_bytes = (uint32_t)hid * 16 * 4, w2_bytes = 32u * hid * 4;
  if (tid == 0) {
    mbar_arrive_expect_tx(&bar_const, w1_bytes + w2_bytes + 64 * 128 * 2);
    bulk_g2s(sW1, a.W1img, w1_bytes, &bar_const);
    bulk_g2s(sW2, a.W2img, w2_bytes, &bar_const);
    bulk_g2s(sF, a.Fimg, 64 * 128 * 2, &bar_const);
  }
  mbar_wait(&bar_const, 0);
  uint32_t ph_d3 = 0;
  MlpBars bars{&bar_d1, bar_a2, 0u, {0u, 0u}};

  for (int tile = blockIdx.x; tile < g.ntiles; tile += gridDim.x) {
    // gather: channel order constants | prescribed (t, c) | prognostic (t, c)  (fno.py:58-62)
    {
      const long long p = (long long)tile * kTcTile + tid;
      const long long b = p / g.HW, pix = p - b * g.HW;
      int k = 0;
      for (int gi = 0; gi < a.ngrp; ++gi) {
        const float* src = a.grp[gi].base + b * a.grp[gi].bstride + pix;
        for (int ch = 0; ch < a.grp[gi].nch; ++ch, ++k) {
          const float val = src[(long long)ch * g.HW];
          const __nv_bfloat16 h = __float2bfloat16_rn(val);
          const __nv_bfloat16 l = __float2bfloat16_rn(val - __bfloat162float(h));
          uint8_t* d = sA1 + (k >> 3) * 2048 + tid * 16 + (k & 7) * 2;
          *reinterpret_cast<__nv_bfloat16*>(d) = h;
          *reinterpret_cast<__nv_bfloat16*>(d + 4096) = l;
        }
      }
    }
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    mlp_tile<16, 32, false>(tid, smem_u32(sA1), smem_u32(sW1), hid, nullptr, smem_u32(sW2), sA2, tmem_d1, tmem_d2, bars);

    float v[32];
    {
      uint32_t r[32];
      tmem_ld32(tmem_d2 + ((uint32_t)(warp * 32) << 16), r);
      tmem_ld_wait();
#pragma unroll
      for (int c = 0; c < 32; ++c) v[c] = __uint_as_float(r[c]) + sB2[c];
    }
    finish_tile(v, tid, tile, g, a.y_img, a.Aout, sF, sXT, stage, tmem_d3, &bar_d3, ph_d3);
    tc_fence_before();
    __syncthreads();
  }
  if (warp == 0) tmem_dealloc(tmem_s, 256);
}

int launch_tc_lift(const TcLiftArgs& a, cudaStream_t st) {
  static int sms = 0;
  if (!sms) {
    int dev = 0;
    DLWP_CUDA(cudaGetDevice(&dev));
    DLWP_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    DLWP_CUDA(cudaFuncSetAttribute(fno_lift_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, LiftSmem::Total));
  }
  DLWP_REQUIRE(a.hid == 128 || a.hid == 256, "tensor-core lifting needs width 128 or 256, got %d", a.hid);
  DLWP_REQUIRE(a.Cin >= 1 && a.Cin <= 15, "tensor-core lifting needs 1..15 input channels, got %d", a.Cin);
  const int grid = a.g.ntiles < 2 * sms ? a.g.ntiles : 2 * sms;
  fno_lift_kernel<<<grid, 128, LiftSmem::Total, st>>>(a);
  DLWP_LAUNCHED("fno_lift_kernel");
  return DLWP_OK;
}

// =============================================================================================
// host-side packing of constant operands into chunk-major bf16 images
// =============================================================================================
namespace {
inline uint16_t f2bf(float f) {
  uint32_t u;
  std::memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return 0x7fc0;
  u += 0x7fffu + ((u >> 16) & 1u);
  return (uint16_t)(u >> 16);
}
inline float bf2f(uint16_t h) {
  const uint32_t u = (uint32_t)h << 16;
  float f;
  std::memcpy(&f, &u, 4);
  return f;
}
inline size_t img_off(int r, int k, int R) { return (size_t)(k >> 3) * R * 16 + (size_t)r * 16 + (size_t)(k & 7) * 2; }
inline void put(uint8_t* img, int r, int k, int R, uint16_t v) { std::memcpy(img + img_off(r, k, R), &v, 2); }
}  // namespace

void tc_pack_images_f64(const double* m, int rows, int cols, int ld, int R, int K, uint8_t* out) {
  const size_t part = (size_t)R * K * 2;
  std::memset(out, 0, 2 * part);
  for (int r = 0; r < rows; ++r)
    for (int k = 0; k < cols; ++k) {
      const double v = m[(size_t)r * ld + k];
      const uint16_t h = f2bf((float)v);
      const uint16_t l = f2bf((float)(v - (double)bf2f(h)));
      put(out, r, k, R, h);
      put(out + part, r, k, R, l);
    }
}

void tc_pack_images(const float* m, int rows, int cols, int ld, int R, int K, uint8_t* out) {
  const size_t part = (size_t)R * K * 2;
  std::memset(out, 0, 2 * part);
  for (int r = 0; r < rows; ++r)
    for (int k = 0; k < cols; ++k) {
      const float v = m[(size_t)r * ld + k];
      const uint16_t h = f2bf(v);
      const uint16_t l = f2bf(v - bf2f(h));
      put(out, r, k, R, h);
      put(out + part, r, k, R, l);
    }
}

void tc_pack_stacked_f64(const double* m, int rows, int cols, int ld, int R, int K, uint8_t* out) {
  std::memset(out, 0, (size_t)R * K * 2);
  const int half = R / 2;
  for (int r = 0; r < rows && r < half; ++r)
    for (int k = 0; k < cols; ++k) {
      const double v = m[(size_t)r * ld + k];
      const uint16_t h = f2bf((float)v);
      const uint16_t l = f2bf((float)(v - (double)bf2f(h)));
      put(out, r, k, R, h);
      put(out, half + r, k, R, l);
    }
}

}  // namespace dlwp
