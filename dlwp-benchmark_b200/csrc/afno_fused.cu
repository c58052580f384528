// Fused fp32 kernels of the AFNO2D token mixer at embed_dim 64 (fourcastnet.py:78-127,176-187):
// three launches per block instead of LayerNorm + four transforms + the block MLP.
//
//   afno_fwd_kernel   y = LayerNorm(x) per token, then the truncated W-DFT of y           one CTA per grid row
//                     A[b][kx][h][c] = sum_w Fw[kx, w] y[b, h, w, c]                      (:84-85, rfft2 W half)
//   afno_mid_kernel   per (sample, retained column): H-DFT, the block-diagonal complex 2-layer MLP with
//                     ReLU / softshrink on re and im separately, inverse H-DFT            (:85-124)
//   afno_inv_kernel   t2 = Re(Gw Z) + y + x: C2R inverse W-DFT, AFNO2D's "+ bias" (its own input, :127)
//                     and Block's double skip (:185-187); y is recomputed from x          one CTA per grid row
// Layouts: A [B][Kc][H][64] and Z [B][H][Kc][64] complex, so that every producer writes and every
// consumer reads 512-byte channel rows; the spectrum of one (sample, column) stays in shared memory
// through the whole mid kernel.  All three start their successor early (programmatic dependent
// launch) and load their constant operands before they wait for their predecessor.
#include "tc_afno.cuh"
#include "tc_common.cuh"

namespace dlwp {

namespace {
constexpr int kC = kAfnoE;   // 64 channels: one thread per channel and quarter of the outputs

// LayerNorm of the W tokens of one grid row held in shared memory [W][64], in place (warp per token, two
// passes like torch); `plus_x` leaves LayerNorm(x) + x.  g / be: this lane's two channels of the affine.
__device__ __forceinline__ void row_layernorm(float* sy, int W, float2 g, float2 be, float eps, bool plus_x) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int w = warp; w < W; w += 8) {
    float2* p = reinterpret_cast<float2*>(sy + (size_t)w * kC) + lane;
    const float2 v = *p;
    float s = v.x + v.y;
#pragma unroll
    for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    const float mean = s * (1.0f / kC);
    const float d0 = v.x - mean, d1 = v.y - mean;
    float q = d0 * d0 + d1 * d1;
#pragma unroll
    for (int o = 16; o; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
    const float rstd = rsqrtf(q * (1.0f / kC) + eps);
    float2 y = make_float2(d0 * rstd * g.x + be.x, d1 * rstd * g.y + be.y);
    if (plus_x) {
      y.x += v.x;
      y.y += v.y;
    }
    *p = y;
  }
}

// two mbarriers per CTA: [0] constant operands (requested before the predecessor kernel has finished),
// [1] what the predecessor wrote.  Every operand arrives by bulk async copy: one request, no per-thread
// load latency chains.
__device__ __forceinline__ void init_bars(uint64_t* bar) {
  if (threadIdx.x == 0) {
    tc::mbar_init(&bar[0], 1);
    tc::mbar_init(&bar[1], 1);
    tc::fence_mbar_init();
  }
}
}  // namespace

// ---------------------------------------------------------------------------------------------
template <int NK, bool EMBED>   // NK retained columns per thread: kx = kg + 4 j
__global__ void __launch_bounds__(256) afno_fwd_kernel(const AfnoFwdArgs a) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  __shared__ uint64_t bar[2];
  tc::pdl_launch_dependents();
  float* sy = reinterpret_cast<float*>(smem_raw);                       // [W][64]
  float2* sF = reinterpret_cast<float2*>(smem_raw + (size_t)a.W * kC * 4);   // [4 * NK][W]
  float* sIn = reinterpret_cast<float*>(sF + 4 * NK * a.W);              // EMBED: [Cin][W] input row, then [Cin][64] weights
  float* sWe = sIn + 16 * 64;
  const int tid = threadIdx.x, c = tid & 63, kg = tid >> 6, lane = tid & 31;
  init_bars(bar);
  for (int i = a.Kc * a.W + tid; i < 4 * NK * a.W; i += 256) sF[i] = make_float2(0.f, 0.f);
  const float2 g = reinterpret_cast<const float2*>(a.gamma)[lane], be = reinterpret_cast<const float2*>(a.beta)[lane];
  const int bh = blockIdx.x, b = bh / a.H, h = bh - b * a.H;
  float xe[16];     // EMBED: bias + position embedding of this thread's (token, channel) outputs, w = kg + 4 j
  if constexpr (EMBED) {
    for (int i = tid; i < a.Cin * kC; i += 256) {     // transposed: one channel row per input channel
      const int k = i >> 6, cc = i & 63;
      sWe[i] = a.We[cc * a.Cin + k];
    }
    const float bias = a.be[c];
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      const int w = kg + 4 * j;
      xe[j] = bias + ((a.pos != nullptr && w < a.W) ? a.pos[((size_t)h * a.W + w) * kC + c] : 0.f);
    }
  }
  __syncthreads();
  if (tid == 0) {
    tc::mbar_arrive_expect_tx(&bar[0], (uint32_t)(a.Kc * a.W * 8));
    tc::bulk_g2s(sF, a.Fw, (uint32_t)(a.Kc * a.W * 8), &bar[0]);
  }
  tc::pdl_wait();
  if constexpr (EMBED) {
    // gather the row of x_t from the NCHW sources (fourcastnet.py:294-307): coalesced along w
    const int HWp = a.H * a.W;
    for (int i = tid; i < a.Cin * a.W; i += 256) {
      const int k = i / a.W, w = i - k * a.W;
      int ch = k;
      float v = 0.f;
#pragma unroll
      for (int gi = 0; gi < 4; ++gi) {
        if (gi < a.ngrp) {
          if (ch >= 0 && ch < a.grp[gi].nch) v = a.grp[gi].base[(size_t)b * a.grp[gi].bstride + (size_t)ch * HWp + (size_t)h * a.W + w];
          ch -= a.grp[gi].nch;
        }
      }
      sIn[k * 64 + w] = v;
    }
    __syncthreads();
    float* xrow = a.xout + (size_t)bh * a.W * kC;
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      const int w = kg + 4 * j;
      if (w < a.W) {
        float acc = xe[j];
        for (int k = 0; k < a.Cin; ++k) acc = fmaf(sIn[k * 64 + w], sWe[k * kC + c], acc);
        sy[w * kC + c] = acc;
        xrow[w * kC + c] = acc;
      }
    }
    __syncthreads();
  } else {
    if (tid == 0) {
      tc::mbar_arrive_expect_tx(&bar[1], (uint32_t)(a.W * kC * 4));
      tc::bulk_g2s(sy, a.x + (size_t)bh * a.W * kC, (uint32_t)(a.W * kC * 4), &bar[1]);
    }
    tc::mbar_wait(&bar[1], 0);
  }
  row_layernorm(sy, a.W, g, be, a.eps, false);
  tc::mbar_wait(&bar[0], 0);
  __syncthreads();
  float2 acc[NK];
#pragma unroll
  for (int j = 0; j < NK; ++j) acc[j] = make_float2(0.f, 0.f);
#pragma unroll 4
  for (int w = 0; w < a.W; ++w) {
    const float y = sy[w * kC + c];
#pragma unroll
    for (int j = 0; j < NK; ++j) {
      const float2 f = sF[(kg + 4 * j) * a.W + w];
      acc[j].x = fmaf(f.x, y, acc[j].x);
      acc[j].y = fmaf(f.y, y, acc[j].y);
    }
  }
#pragma unroll
  for (int j = 0; j < NK; ++j) {
    const int kx = kg + 4 * j;
    if (kx < a.Kc) a.A[(((size_t)b * a.Kc + kx) * a.H + h) * kC + c] = acc[j];
  }
}

template <int NK>
static int fwd_go(const AfnoFwdArgs& a, cudaStream_t st) {
  const dim3 grid((unsigned)(a.B * a.H)), block(256);
  const size_t smem = (size_t)a.W * kC * 4 + (size_t)4 * NK * a.W * 8 + (a.ngrp > 0 ? (size_t)2 * 16 * 64 * 4 : 0);
  if (a.ngrp > 0) DLWP_CUDA(launch_pdl(afno_fwd_kernel<NK, true>, grid, block, smem, st, a));
  else DLWP_CUDA(launch_pdl(afno_fwd_kernel<NK, false>, grid, block, smem, st, a));
  return DLWP_OK;
}

int launch_afno_fwd(const AfnoFwdArgs& a, cudaStream_t st) {
  DLWP_REQUIRE(a.B > 0 && a.H > 0 && a.W > 0 && a.W <= 64 && a.Kc > 0 && a.Kc <= 36, "afno_fwd_kernel: unsupported shape");
  DLWP_REQUIRE(a.ngrp == 0 || (a.Cin >= 1 && a.Cin <= 16 && a.We && a.be && a.xout), "afno_fwd_kernel: bad embedding arguments");
  const int nk = ceil_div(a.Kc, 4);
  if (nk <= 3) DLWP_TRY(fwd_go<3>(a, st));
  else if (nk <= 5) DLWP_TRY(fwd_go<5>(a, st));
  else DLWP_TRY(fwd_go<9>(a, st));
  DLWP_LAUNCHED("afno_fwd_kernel");
  return DLWP_OK;
}

// ---------------------------------------------------------------------------------------------
namespace {
__device__ __forceinline__ void cfma(float2& acc, float2 f, float2 v) {   // acc += f * v (complex)
  acc.x = fmaf(f.x, v.x, acc.x);
  acc.x = fmaf(-f.y, v.y, acc.x);
  acc.y = fmaf(f.x, v.y, acc.y);
  acc.y = fmaf(f.y, v.x, acc.y);
}

// one layer of the block-diagonal complex MLP on the modes r = rg + 4 m of this CTA's column:
// out[r][k * BS + j] = act(sum_i in[r][k * BS + i] * w[k][i][j] + b[k][j]); the thread's BS complex
// weights move from shared memory to registers once.  w [2, nb, BS, BS], b [2, nb, BS] (planes: real, imaginary)
template <int BS, bool RELU>
__device__ __forceinline__ void block_layer(const float2* in, float2* out, const float* w, const float* bias,
                                            int R, float lambda) {
  constexpr int nb = kC / BS;
  const int kj = threadIdx.x & 63, rg = threadIdx.x >> 6, k = kj / BS, j = kj - k * BS;
  float2 wv[BS];
#pragma unroll
  for (int i = 0; i < BS; ++i)
    wv[i] = make_float2(w[(k * BS + i) * BS + j], w[((nb + k) * BS + i) * BS + j]);
  const float2 bv = make_float2(bias[k * BS + j], bias[(nb + k) * BS + j]);
  float2 acc[8];
#pragma unroll
  for (int m = 0; m < 8; ++m) acc[m] = bv;
#pragma unroll
  for (int i = 0; i < BS; ++i)
#pragma unroll
    for (int m = 0; m < 8; ++m) cfma(acc[m], wv[i], in[(rg + 4 * m) * kC + k * BS + i]);
#pragma unroll
  for (int m = 0; m < 8; ++m) {
    const int r = rg + 4 * m;
    if (r < R)
      out[r * kC + kj] = RELU ? make_float2(fmaxf(acc[m].x, 0.f), fmaxf(acc[m].y, 0.f))
                              : make_float2(softshrink(acc[m].x, lambda), softshrink(acc[m].y, lambda));
  }
}
// two [32][64] complex buffers + Fh + Gh + the filter weights and biases
constexpr int mid_smem(int bs) { return 2 * 32 * kC * 8 + 2 * 32 * 32 * 8 + 2 * (2 * kC * bs * 4) + 2 * (2 * kC * 4); }
}  // namespace

template <int BS>
__global__ void __launch_bounds__(256) afno_mid_kernel(const AfnoMidArgs a) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  __shared__ uint64_t bar[2];
  tc::pdl_launch_dependents();
  float2* bufA = reinterpret_cast<float2*>(smem_raw);     // [32][64]
  float2* bufB = bufA + 32 * kC;
  float2* sFh = bufB + 32 * kC;                           // [32][H] (rows >= R are zero)
  float2* sGh = sFh + 32 * 32;                            // [32][R] (rows >= H are zero)
  float* sW1 = reinterpret_cast<float*>(sGh + 32 * 32);   // [2, nb, BS, BS]
  float* sW2 = sW1 + 2 * kC * BS;
  float* sB1 = sW2 + 2 * kC * BS;                         // [2, nb, BS]
  float* sB2 = sB1 + 2 * kC;
  const int tid = threadIdx.x, c = tid & 63, g4 = tid >> 6;
  const int H = a.H, R = a.R;
  init_bars(bar);
  for (int i = R * H + tid; i < 32 * H; i += 256) sFh[i] = make_float2(0.f, 0.f);
  for (int i = H * R + tid; i < 32 * R; i += 256) sGh[i] = make_float2(0.f, 0.f);
  __syncthreads();
  const int bk = blockIdx.x, b = bk / a.Kc, kx = bk - b * a.Kc;
  if (tid == 0) {
    const uint32_t fb = (uint32_t)(R * H * 8), wb = 2u * kC * BS * 4u, bb = 2u * kC * 4u;
    tc::mbar_arrive_expect_tx(&bar[0], 2 * fb + 2 * wb + 2 * bb);
    tc::bulk_g2s(sFh, a.Fh, fb, &bar[0]);
    tc::bulk_g2s(sGh, a.Gh, fb, &bar[0]);
    tc::bulk_g2s(sW1, a.w1, wb, &bar[0]);
    tc::bulk_g2s(sW2, a.w2, wb, &bar[0]);
    tc::bulk_g2s(sB1, a.b1, bb, &bar[0]);
    tc::bulk_g2s(sB2, a.b2, bb, &bar[0]);
  }
  tc::pdl_wait();
  if (tid == 0) {   // the column's spectrum after the W-DFT: H x 64 complex, contiguous
    tc::mbar_arrive_expect_tx(&bar[1], (uint32_t)(H * kC * 8));
    tc::bulk_g2s(bufA, a.A + (size_t)bk * H * kC, (uint32_t)(H * kC * 8), &bar[1]);
  }
  tc::mbar_wait(&bar[0], 0);
  tc::mbar_wait(&bar[1], 0);
  {   // H-DFT: X[r][c] = sum_h Fh[r][h] A[h][c]
    float2 acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = make_float2(0.f, 0.f);
#pragma unroll 2
    const float* pl = reinterpret_cast<const float*>(bufA);
    for (int h = 0; h < H; ++h) {
      const float2 v = a.planarA ? make_float2(pl[(2 * h) * kC + c], pl[(2 * h + 1) * kC + c]) : bufA[h * kC + c];
#pragma unroll
      for (int j = 0; j < 8; ++j) cfma(acc[j], sFh[(g4 + 4 * j) * H + h], v);
    }
#pragma unroll
    for (int j = 0; j < 8; ++j)
      if (g4 + 4 * j < R) bufB[(g4 + 4 * j) * kC + c] = acc[j];
  }
  __syncthreads();
  block_layer<BS, true>(bufB, bufA, sW1, sB1, R, 0.f);
  __syncthreads();
  block_layer<BS, false>(bufA, bufB, sW2, sB2, R, a.lambda);
  __syncthreads();
  {   // inverse H-DFT: Z[h][c] = sum_r Gh[h][r] Y[r][c]
    float2 acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = make_float2(0.f, 0.f);
#pragma unroll 2
    for (int r = 0; r < R; ++r) {
      const float2 v = bufB[r * kC + c];
#pragma unroll
      for (int j = 0; j < 8; ++j) cfma(acc[j], sGh[(g4 + 4 * j) * R + r], v);
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int h = g4 + 4 * j;
      if (h < H) a.Z[(((size_t)b * H + h) * a.Kc + kx) * kC + c] = acc[j];
    }
  }
}

int launch_afno_mid(const AfnoMidArgs& a, cudaStream_t st) {
  DLWP_REQUIRE(a.B > 0 && a.H > 0 && a.H <= 32 && a.R > 0 && a.R <= 32 && a.Kc > 0, "afno_mid_kernel: unsupported shape");
  DLWP_REQUIRE(a.bs == 8 || a.bs == 16, "afno_mid_kernel: block size %d", a.bs);
  DLWP_TRY(ensure_dynamic_smem((const void*)afno_mid_kernel<8>, mid_smem(8)));
  DLWP_TRY(ensure_dynamic_smem((const void*)afno_mid_kernel<16>, mid_smem(16)));
  const dim3 grid((unsigned)(a.B * a.Kc)), block(256);
  if (a.bs == 8) DLWP_CUDA(launch_pdl(afno_mid_kernel<8>, grid, block, (size_t)mid_smem(8), st, a));
  else DLWP_CUDA(launch_pdl(afno_mid_kernel<16>, grid, block, (size_t)mid_smem(16), st, a));
  DLWP_LAUNCHED("afno_mid_kernel");
  return DLWP_OK;
}

// ---------------------------------------------------------------------------------------------
template <int NW>   // pixels per thread: w = wg + 4 j
__global__ void __launch_bounds__(256) afno_inv_kernel(const AfnoInvArgs a) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  __shared__ uint64_t bar[2];
  tc::pdl_launch_dependents();
  float* sy = reinterpret_cast<float*>(smem_raw);                            // [4 * NW][64]: LayerNorm(x) + x
  float2* sG = reinterpret_cast<float2*>(smem_raw + (size_t)4 * NW * kC * 4);    // [4 * NW][Kc]
  float2* sZ = sG + 4 * NW * a.Kc;                                            // [Kc][64]
  const int tid = threadIdx.x, c = tid & 63, wg = tid >> 6, lane = tid & 31;
  init_bars(bar);
  for (int i = a.W * a.Kc + tid; i < 4 * NW * a.Kc; i += 256) sG[i] = make_float2(0.f, 0.f);
  const float2 g = reinterpret_cast<const float2*>(a.gamma)[lane], be = reinterpret_cast<const float2*>(a.beta)[lane];
  __syncthreads();
  const int bh = blockIdx.x;
  if (tid == 0) {
    tc::mbar_arrive_expect_tx(&bar[0], (uint32_t)(a.W * a.Kc * 8));
    tc::bulk_g2s(sG, a.Gw, (uint32_t)(a.W * a.Kc * 8), &bar[0]);
  }
  tc::pdl_wait();
  if (tid == 0) {
    tc::mbar_arrive_expect_tx(&bar[1], (uint32_t)(a.Kc * kC * 8 + a.W * kC * 4));
    tc::bulk_g2s(sZ, a.Z + (size_t)bh * a.Kc * kC, (uint32_t)(a.Kc * kC * 8), &bar[1]);
    tc::bulk_g2s(sy, a.x + (size_t)bh * a.W * kC, (uint32_t)(a.W * kC * 4), &bar[1]);
  }
  tc::mbar_wait(&bar[1], 0);
  row_layernorm(sy, a.W, g, be, a.eps, true);
  tc::mbar_wait(&bar[0], 0);
  __syncthreads();
  float acc[NW];
#pragma unroll
  for (int j = 0; j < NW; ++j) acc[j] = 0.f;
  for (int kx = 0; kx < a.Kc; ++kx) {
    const float2 z = sZ[kx * kC + c];
#pragma unroll
    for (int j = 0; j < NW; ++j) {
      const float2 g = sG[(wg + 4 * j) * a.Kc + kx];
      acc[j] = fmaf(g.x, z.x, acc[j]);
      acc[j] = fmaf(-g.y, z.y, acc[j]);
    }
  }
  float* orow = a.out + (size_t)bh * a.W * kC;
#pragma unroll
  for (int j = 0; j < NW; ++j) {
    const int w = wg + 4 * j;
    if (w < a.W) orow[w * kC + c] = acc[j] + sy[w * kC + c];
  }
}

int launch_afno_inv(const AfnoInvArgs& a, cudaStream_t st) {
  DLWP_REQUIRE(a.B > 0 && a.H > 0 && a.W > 0 && a.W <= 64 && a.Kc > 0 && a.Kc <= 36, "afno_inv_kernel: unsupported shape");
  const dim3 grid((unsigned)(a.B * a.H)), block(256);
  const int nw = a.W <= 32 ? 8 : 16;
  const size_t smem = (size_t)4 * nw * kC * 4 + (size_t)4 * nw * a.Kc * 8 + (size_t)a.Kc * kC * 8;
  DLWP_TRY(ensure_dynamic_smem((const void*)afno_inv_kernel<8>, 64 * 1024));
  DLWP_TRY(ensure_dynamic_smem((const void*)afno_inv_kernel<16>, 64 * 1024));
  DLWP_REQUIRE(smem <= 64 * 1024, "afno_inv_kernel: shared memory");
  if (nw == 8) DLWP_CUDA(launch_pdl(afno_inv_kernel<8>, grid, block, smem, st, a));
  else DLWP_CUDA(launch_pdl(afno_inv_kernel<16>, grid, block, smem, st, a));
  DLWP_LAUNCHED("afno_inv_kernel");
  return DLWP_OK;
}

}  // namespace dlwp
