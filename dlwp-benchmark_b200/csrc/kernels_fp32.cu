// fp32 CUDA-core kernel set (precision mode DLWP_PREC_FP32): the parity path.
//
// Data layout: activations are channels-last ("NHWC") fp32, [B, H, W, C]; spectra are
// complex float2.  The rfft2 / irfft2 of the reference (torch.fft.rfft2 at
// fourcastnet.py:85,124; unet.py:49,68; neuralop rfftn/irfftn) are evaluated ONLY on the
// retained bins as products with explicit truncated-DFT matrices (built in fp64 on the
// host, plan.cu), W axis first then H, and back H then W.
#include "kernels.cuh"

#include <cuda_bf16.h>
#include <cstdint>

namespace dlwp {

// =====================================================================================
// tdft_kernel: out[bt, k, c] = epi(sum_n T[k, n] * in[bt, n, c])
//   block 256 = 8 k-groups (warps) x 32 channels (lanes); each thread owns KPT outputs
//   k = k0 + g + 8 j.  T tile reads are warp-broadcast, in-tile reads are conflict-free.
// =====================================================================================
constexpr int TD_NT = 32;  // contraction chunk held in shared memory

template <bool IN_CPLX, bool OUT_CPLX, int KPT>
__global__ void __launch_bounds__(256) tdft_kernel(TdftArgs a) {
  using in_t = typename std::conditional<IN_CPLX, float2, float>::type;
  __shared__ in_t s_in[TD_NT][32];
  __shared__ float2 s_T[8 * KPT][TD_NT];

  const int tid = threadIdx.x, lane = tid & 31, g = tid >> 5;
  const int bt = blockIdx.x;
  const int b0 = bt / a.nb1, b1 = bt - b0 * a.nb1;
  const int c0 = blockIdx.y * 32;
  const int k0 = blockIdx.z * (8 * KPT);
  const int c = c0 + lane;

  const in_t* __restrict__ in = reinterpret_cast<const in_t*>(a.in) + b0 * a.in_sb0 + b1 * a.in_sb1;
  const float2* __restrict__ T = a.T;

  float2 acc[KPT];
#pragma unroll
  for (int j = 0; j < KPT; ++j) acc[j] = make_float2(0.f, 0.f);

  for (int n0 = 0; n0 < a.N; n0 += TD_NT) {
    // input tile [TD_NT][32]
#pragma unroll
    for (int it = 0; it < (TD_NT * 32) / 256; ++it) {
      const int idx = tid + it * 256;
      const int nn = idx >> 5, cc = idx & 31;
      const int n = n0 + nn;
      in_t v;
      if constexpr (IN_CPLX) v = make_float2(0.f, 0.f); else v = 0.f;
      if (n < a.N && c0 + cc < a.C) v = in[(long long)n * a.in_sn + c0 + cc];
      s_in[nn][cc] = v;
    }
    // T tile [8*KPT][TD_NT]
    for (int idx = tid; idx < 8 * KPT * TD_NT; idx += 256) {
      const int kk = idx / TD_NT, nn = idx - kk * TD_NT;
      const int k = k0 + kk, n = n0 + nn;
      s_T[kk][nn] = (k < a.K && n < a.N) ? T[(long long)k * a.N + n] : make_float2(0.f, 0.f);
    }
    __syncthreads();
#pragma unroll 8
    for (int nn = 0; nn < TD_NT; ++nn) {
      const in_t v = s_in[nn][lane];
#pragma unroll
      for (int j = 0; j < KPT; ++j) {
        const float2 t = s_T[g + 8 * j][nn];
        if constexpr (IN_CPLX) {
          acc[j].x = fmaf(t.x, v.x, acc[j].x);
          acc[j].x = fmaf(-t.y, v.y, acc[j].x);
          if constexpr (OUT_CPLX) {
            acc[j].y = fmaf(t.x, v.y, acc[j].y);
            acc[j].y = fmaf(t.y, v.x, acc[j].y);
          }
        } else {
          acc[j].x = fmaf(t.x, v, acc[j].x);
          acc[j].y = fmaf(t.y, v, acc[j].y);
        }
      }
    }
    __syncthreads();
  }

  if (c >= a.C) return;
  const long long obase = b0 * a.out_sb0 + b1 * a.out_sb1 + c;
#pragma unroll
  for (int j = 0; j < KPT; ++j) {
    const int k = k0 + g + 8 * j;
    if (k >= a.K) continue;
    const long long o = obase + (long long)k * a.out_sk;
    if constexpr (OUT_CPLX) {
      reinterpret_cast<float2*>(a.out)[o] = acc[j];
    } else {
      float v = acc[j].x;
      if (a.add1) v += a.add1[o];
      if (a.add2) v += a.add2[o];
      if (a.bias) v += a.bias[c];
      if (a.act == 1) v = gelu_erf(v);
      reinterpret_cast<float*>(a.out)[o] = v;
    }
  }
}

template <bool IC, bool OC, int KPT>
static int tdft_go(const TdftArgs& a, dim3 grid, cudaStream_t st) {
  tdft_kernel<IC, OC, KPT><<<grid, 256, 0, st>>>(a);
  DLWP_LAUNCHED("tdft_kernel");
  return DLWP_OK;
}

template <bool IC, bool OC>
static int tdft_pick(const TdftArgs& a, cudaStream_t st) {
  int kpt = ceil_div(a.K, 8);
  if (kpt > 8) kpt = 8;
  // few (batch, channel) blocks: split the outputs over more blocks so that the grid covers the machine
  const long long bc = (long long)a.nb0 * a.nb1 * ceil_div(a.C, 32);
  while (kpt > 2 && bc * ceil_div(a.K, 8 * kpt) < 296) kpt = (kpt + 1) / 2;
  dim3 grid((unsigned)(a.nb0 * a.nb1), (unsigned)ceil_div(a.C, 32), (unsigned)ceil_div(a.K, 8 * kpt));
  switch (kpt) {
    case 1: return tdft_go<IC, OC, 1>(a, grid, st);
    case 2: return tdft_go<IC, OC, 2>(a, grid, st);
    case 3: return tdft_go<IC, OC, 3>(a, grid, st);
    case 4: return tdft_go<IC, OC, 4>(a, grid, st);
    case 5: return tdft_go<IC, OC, 5>(a, grid, st);
    case 6: return tdft_go<IC, OC, 6>(a, grid, st);
    case 7: return tdft_go<IC, OC, 7>(a, grid, st);
    default: return tdft_go<IC, OC, 8>(a, grid, st);
  }
}

int launch_tdft(const TdftArgs& a, cudaStream_t st) {
  DLWP_REQUIRE(a.N > 0 && a.K > 0 && a.C > 0 && a.nb0 > 0 && a.nb1 > 0, "tdft: empty problem");
  DLWP_REQUIRE((long long)a.nb0 * a.nb1 < (1ll << 31) && ceil_div(a.C, 32) < 65536, "tdft: grid too large");
  if (!a.in_complex && a.out_complex) return tdft_pick<false, true>(a, st);
  if (a.in_complex && a.out_complex) return tdft_pick<true, true>(a, st);
  if (a.in_complex && !a.out_complex) return tdft_pick<true, false>(a, st);
  return fail(DLWP_ERR_INVALID, "tdft: real->real transform is not used on this path");
}

// =====================================================================================
// pw_linear_kernel: out[p, n] = epi(sum_k in[p, k] * Wt[n, k])      (64 x 64 x 16 tiles)
//   covers: lifting / projection 1x1 convs (neuralop MLP), the FNO skip conv, PatchEmbed
//   (gather with p1 x p2 patches, fourcastnet.py:537-542), Mlp fc1/fc2 (:51-57) and the
//   head + un-patchify + residual on the last context frame (:349-358, fno.py:103).
// =====================================================================================
constexpr int PW_TM = 64, PW_TN = 64, PW_TK = 16;

__device__ __forceinline__ float pw_gather(const PwArgs& a, int p, int k) {
  const int pp = a.p1 * a.p2;
  const int ch = k / pp, ij = k - ch * pp;
  const int i = ij / a.p2, j = ij - i * a.p2;
  const int b = p / a.tokens, tok = p - b * a.tokens;
  const int wp = a.W / a.p2;
  const int hp = tok / wp, wq = tok - hp * wp;
  const long long pix = (long long)(hp * a.p1 + i) * a.W + (wq * a.p2 + j);
  int c = ch;
#pragma unroll
  for (int gi = 0; gi < 4; ++gi) {
    if (gi < a.ngrp) {
      if (c < a.grp[gi].nch)
        return a.grp[gi].base[b * a.grp[gi].bstride + (long long)c * a.H * a.W + pix];
      c -= a.grp[gi].nch;
    }
  }
  return 0.f;
}

template <bool GATHER>
__global__ void __launch_bounds__(256) pw_linear_kernel(PwArgs a) {
  __shared__ float As[PW_TK][PW_TM + 4];
  __shared__ float Bs[PW_TK][PW_TN + 4];
  const int tid = threadIdx.x;
  const int p0 = blockIdx.x * PW_TM, n0 = blockIdx.y * PW_TN;
  const int tm = tid & 15, tn = tid >> 4;

  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  for (int k0 = 0; k0 < a.K; k0 += PW_TK) {
    if constexpr (GATHER) {
      const int m = tid & 63;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int kk = (tid >> 6) + 4 * i;
        const int p = p0 + m, k = k0 + kk;
        As[kk][m] = (p < a.P && k < a.K) ? pw_gather(a, p, k) : 0.f;
      }
    } else {
      const int m = tid >> 2, kq = (tid & 3) * 4;
      const int p = p0 + m;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int k = k0 + kq + i;
        As[kq + i][m] = (p < a.P && k < a.K) ? a.in[(long long)p * a.K + k] : 0.f;
      }
    }
    {
      const int nn = tid >> 2, kq = (tid & 3) * 4;
      const int n = n0 + nn;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int k = k0 + kq + i;
        Bs[kq + i][nn] = (n < a.N && k < a.K) ? a.Wt[(long long)n * a.K + k] : 0.f;
      }
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < PW_TK; ++kk) {
      float av[4], bv[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) av[i] = As[kk][tm * 4 + i];
#pragma unroll
      for (int j = 0; j < 4; ++j) bv[j] = Bs[kk][tn * 4 + j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    __syncthreads();
  }

#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int n = n0 + tn * 4 + j;
    if (n >= a.N) continue;
    const float bias = a.bias ? a.bias[n] : 0.f;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int p = p0 + tm * 4 + i;
      if (p >= a.P) continue;
      float v = acc[i][j] + bias;
      if (a.pre) v += a.pre[(long long)p * a.N + n];
      if (a.preact) a.preact[(long long)p * a.N + n] = v;
      if (a.act == 1) v = gelu_erf(v);
      if (a.pos) v += a.pos[(long long)(p % a.tokens) * a.N + n];
      if (a.resid) v += a.resid[(long long)p * a.N + n];
      if (a.resid2) v += a.resid2[(long long)p * a.N + n];
      if (!a.unpatch) {
        a.out[(long long)p * a.N + n] = v;
      } else {
        // n = (i, j, co)  <->  "b h w (p1 p2 c_out) -> b c_out (h p1) (w p2)"
        const int b = p / a.tokens, tok = p - b * a.tokens;
        const int wp = a.W / a.p2;
        const int hp = tok / wp, wq = tok - hp * wp;
        const int co = n % a.Cg, ij = n / a.Cg;
        const int pi = ij / a.p2, pj = ij - pi * a.p2;
        const long long off = (long long)co * a.H * a.W + (long long)(hp * a.p1 + pi) * a.W + (wq * a.p2 + pj);
        const float r = a.last ? a.last[b * a.last_bstride + off] : 0.f;
        a.out[b * a.out_bstride + off] = v + r;
      }
    }
  }
}

int launch_pw_linear(const PwArgs& a, cudaStream_t st) {
  DLWP_REQUIRE(a.P > 0 && a.K > 0 && a.N > 0, "pw_linear: empty problem");
  DLWP_REQUIRE(a.tokens > 0, "pw_linear: tokens per sample must be set");
  dim3 grid((unsigned)ceil_div(a.P, PW_TM), (unsigned)ceil_div(a.N, PW_TN));
  if (a.ngrp > 0) {
    DLWP_REQUIRE(a.p1 > 0 && a.p2 > 0 && a.H % a.p1 == 0 && a.W % a.p2 == 0, "pw_linear: bad patch size");
    pw_linear_kernel<true><<<grid, 256, 0, st>>>(a);
  } else {
    DLWP_REQUIRE(a.in != nullptr, "pw_linear: null input");
    pw_linear_kernel<false><<<grid, 256, 0, st>>>(a);
  }
  DLWP_LAUNCHED("pw_linear_kernel");
  return DLWP_OK;
}

// =====================================================================================
// LayerNorm over C for each NHWC token; one warp per token, two passes (mean, then
// centred variance) like torch's CPU/GPU kernels.  fourcastnet.py:172,175,182,189.
// =====================================================================================
__global__ void __launch_bounds__(256) layernorm_kernel(const float* __restrict__ in, float* __restrict__ out,
                                                        const float* __restrict__ gamma,
                                                        const float* __restrict__ beta, int P, int C, float eps) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= P) return;
  const float* x = in + (long long)warp * C;
  float s = 0.f;
  for (int c = lane; c < C; c += 32) s += x[c];
#pragma unroll
  for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  const float mean = s / (float)C;
  float q = 0.f;
  for (int c = lane; c < C; c += 32) {
    const float d = x[c] - mean;
    q = fmaf(d, d, q);
  }
#pragma unroll
  for (int o = 16; o; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
  const float rstd = rsqrtf(q / (float)C + eps);
  float* y = out + (long long)warp * C;
  for (int c = lane; c < C; c += 32) y[c] = (x[c] - mean) * rstd * gamma[c] + beta[c];
}

int launch_layernorm(const float* in, float* out, const float* gamma, const float* beta, int P, int C,
                     float eps, cudaStream_t st) {
  DLWP_REQUIRE(P > 0 && C > 0, "layernorm: empty problem");
  layernorm_kernel<<<ceil_div(P, 8), 256, 0, st>>>(in, out, gamma, beta, P, C, eps);
  DLWP_LAUNCHED("layernorm_kernel");
  return DLWP_OK;
}

// bf16 image output for the tensor-core inverse H-DFT (tc_wide.cuh): per (sample, retained column)
// a [128 rows = (retained row, re | im)] x [128 channels] chunk-major tile; mode m = row * Kc + column
__device__ __forceinline__ void mix_store_image(const MixWideOpts& im, int m, int b, int o, float2 v) {
  const int r = m / im.Kc, kx = m - r * im.Kc;
  uint8_t* p = im.img + ((size_t)b * im.Kc + kx) * 32768 + (size_t)(o >> 3) * 2048 + (size_t)(2 * r) * 16 + (o & 7) * 2;
  *reinterpret_cast<__nv_bfloat16*>(p) = __float2bfloat16(v.x);
  *reinterpret_cast<__nv_bfloat16*>(p + 16) = __float2bfloat16(v.y);
}

// =====================================================================================
// mix_dense_kernel: Y[m, b, o] = sum_i X[m, b, i] * Wm[m, i, o]   (complex)
//   = einsum("bixy,ioxy->boxy") of unet.py:15-17 / neuralop's dense contraction, one
//   retained mode m = (row, col) per blockIdx.x, 8 samples per blockIdx.y.
// =====================================================================================
__global__ void __launch_bounds__(256) mix_dense_kernel(const float2* __restrict__ X,
                                                        const float2* __restrict__ Wm,
                                                        float2* __restrict__ Y, int B, int Ci, int Co,
                                                        MixWideOpts im) {
  extern __shared__ float2 s_x[];  // [8][Ci]
  const int m = blockIdx.x;
  const int b0 = blockIdx.y * 8;
  const int lane = threadIdx.x, ty = threadIdx.y;
  const int tid = ty * 32 + lane;
  for (int idx = tid; idx < 8 * Ci; idx += 256) {
    const int bb = idx / Ci, i = idx - bb * Ci;
    s_x[idx] = (b0 + bb < B) ? X[((long long)m * B + b0 + bb) * Ci + i] : make_float2(0.f, 0.f);
  }
  __syncthreads();
  const int b = b0 + ty;
  if (b >= B) return;
  const float2* __restrict__ w = Wm + (long long)m * Ci * Co;
  const float2* xs = s_x + ty * Ci;
  for (int o = lane; o < Co; o += 32) {
    float ar = 0.f, ai = 0.f;
    for (int i = 0; i < Ci; ++i) {
      const float2 xv = xs[i];
      const float2 wv = w[(long long)i * Co + o];
      ar = fmaf(xv.x, wv.x, ar);
      ar = fmaf(-xv.y, wv.y, ar);
      ai = fmaf(xv.x, wv.y, ai);
      ai = fmaf(xv.y, wv.x, ai);
    }
    if (im.img) mix_store_image(im, m, b, o, make_float2(ar, ai));
    else Y[((long long)m * B + b) * Co + o] = make_float2(ar, ai);
  }
}

// Few samples (B <= 4, e.g. one 721 x 1440 field per GPU): the launch is bound by reading every
// mode's Ci x Co weights once.  Eight warps per mode each take an eighth of the input channels
// with 16-byte weight loads (two outputs per lane), then the partial sums meet in shared memory.
__global__ void __launch_bounds__(256) mix_dense_fewb_kernel(const float2* __restrict__ X,
                                                             const float2* __restrict__ Wm,
                                                             float2* __restrict__ Y, int B, int Ci, int Co,
                                                             MixWideOpts im) {
  extern __shared__ float2 s_mix[];   // x [4][Ci], then partial sums [8][4][Co]
  float2* s_x = s_mix;
  float2* s_p = s_mix + 4 * Ci;
  const int m = blockIdx.x, lane = threadIdx.x, ty = threadIdx.y, tid = ty * 32 + lane;
  for (int idx = tid; idx < 4 * Ci; idx += 256) {
    const int bb = idx / Ci, i = idx - bb * Ci;
    s_x[idx] = bb < B ? X[((long long)m * B + bb) * Ci + i] : make_float2(0.f, 0.f);
  }
  __syncthreads();
  const int half = Co >> 1;
  const float4* __restrict__ w4 = reinterpret_cast<const float4*>(Wm + (long long)m * Ci * Co);
  const int i0 = ty * Ci / 8, i1 = (ty + 1) * Ci / 8;
  for (int op = lane; op < half; op += 32) {
    float2 a[4][2];
#pragma unroll
    for (int bb = 0; bb < 4; ++bb) a[bb][0] = a[bb][1] = make_float2(0.f, 0.f);
#pragma unroll 4
    for (int i = i0; i < i1; ++i) {
      const float4 wv = w4[(long long)i * half + op];
#pragma unroll
      for (int bb = 0; bb < 4; ++bb) {
        const float2 xv = s_x[bb * Ci + i];
        a[bb][0].x = fmaf(xv.x, wv.x, a[bb][0].x);
        a[bb][0].x = fmaf(-xv.y, wv.y, a[bb][0].x);
        a[bb][0].y = fmaf(xv.x, wv.y, a[bb][0].y);
        a[bb][0].y = fmaf(xv.y, wv.x, a[bb][0].y);
        a[bb][1].x = fmaf(xv.x, wv.z, a[bb][1].x);
        a[bb][1].x = fmaf(-xv.y, wv.w, a[bb][1].x);
        a[bb][1].y = fmaf(xv.x, wv.w, a[bb][1].y);
        a[bb][1].y = fmaf(xv.y, wv.z, a[bb][1].y);
      }
    }
#pragma unroll
    for (int bb = 0; bb < 4; ++bb)
      *reinterpret_cast<float4*>(&s_p[(ty * 4 + bb) * Co + 2 * op]) = make_float4(a[bb][0].x, a[bb][0].y, a[bb][1].x, a[bb][1].y);
  }
  __syncthreads();
  for (int idx = tid; idx < B * Co; idx += 256) {
    const int bb = idx / Co, o = idx - bb * Co;
    float2 acc = make_float2(0.f, 0.f);
#pragma unroll
    for (int w = 0; w < 8; ++w) {
      const float2 v = s_p[(w * 4 + bb) * Co + o];
      acc.x += v.x;
      acc.y += v.y;
    }
    if (im.img) mix_store_image(im, m, bb, o, acc);
    else Y[((long long)m * B + bb) * Co + o] = acc;
  }
}

// The same with the mode weights stored as bf16 (re, im) pairs (the bf16 kernel set): half the bytes.
// A lane takes four outputs per 16-byte load, so that a warp keeps as many bytes in flight as above.
__global__ void __launch_bounds__(256) mix_dense_fewb16_kernel(const float2* __restrict__ X, float2* __restrict__ Y, int B,
                                                               int Ci, int Co, MixWideOpts im) {
  extern __shared__ float2 s_mix[];   // x [4][Ci], then partial sums [8][4][Co]
  float2* s_x = s_mix;
  float2* s_p = s_mix + 4 * Ci;
  const int m = blockIdx.x, lane = threadIdx.x, ty = threadIdx.y, tid = ty * 32 + lane;
  for (int idx = tid; idx < 4 * Ci; idx += 256) {
    const int bb = idx / Ci, i = idx - bb * Ci;
    s_x[idx] = bb < B ? X[((long long)m * B + bb) * Ci + i] : make_float2(0.f, 0.f);
  }
  __syncthreads();
  const int quarter = Co >> 2;
  const uint4* __restrict__ w4 = reinterpret_cast<const uint4*>(im.w16 + (long long)m * Ci * Co);
  const int i0 = ty * Ci / 8, i1 = (ty + 1) * Ci / 8;
  for (int oq = lane; oq < quarter; oq += 32) {
    float2 a[4][4];
#pragma unroll
    for (int bb = 0; bb < 4; ++bb)
#pragma unroll
      for (int j = 0; j < 4; ++j) a[bb][j] = make_float2(0.f, 0.f);
#pragma unroll 8
    for (int i = i0; i < i1; ++i) {
      const uint4 u = w4[(long long)i * quarter + oq];
      const uint32_t uw[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float wr = __uint_as_float(uw[j] << 16), wi = __uint_as_float(uw[j] & 0xffff0000u);
#pragma unroll
        for (int bb = 0; bb < 4; ++bb) {
          const float2 xv = s_x[bb * Ci + i];
          a[bb][j].x = fmaf(xv.x, wr, a[bb][j].x);
          a[bb][j].x = fmaf(-xv.y, wi, a[bb][j].x);
          a[bb][j].y = fmaf(xv.x, wi, a[bb][j].y);
          a[bb][j].y = fmaf(xv.y, wr, a[bb][j].y);
        }
      }
    }
#pragma unroll
    for (int bb = 0; bb < 4; ++bb)
#pragma unroll
      for (int j = 0; j < 2; ++j)
        *reinterpret_cast<float4*>(&s_p[(ty * 4 + bb) * Co + 4 * oq + 2 * j]) =
            make_float4(a[bb][2 * j].x, a[bb][2 * j].y, a[bb][2 * j + 1].x, a[bb][2 * j + 1].y);
  }
  __syncthreads();
  for (int idx = tid; idx < B * Co; idx += 256) {
    const int bb = idx / Co, o = idx - bb * Co;
    float2 acc = make_float2(0.f, 0.f);
#pragma unroll
    for (int w = 0; w < 8; ++w) {
      const float2 v = s_p[(w * 4 + bb) * Co + o];
      acc.x += v.x;
      acc.y += v.y;
    }
    if (im.img) mix_store_image(im, m, bb, o, acc);
    else Y[((long long)m * B + bb) * Co + o] = acc;
  }
}

__global__ void narrow_bf16_kernel(const float2* __restrict__ in, uint32_t* __restrict__ out, long long n) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    const __nv_bfloat162 v = __floats2bfloat162_rn(in[i].x, in[i].y);
    out[i] = *reinterpret_cast<const uint32_t*>(&v);
  }
}
int launch_narrow_bf16(const float2* in, uint32_t* out, long long n, cudaStream_t st) {
  narrow_bf16_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(in, out, n);
  DLWP_LAUNCHED("narrow_bf16_kernel");
  return DLWP_OK;
}

int launch_mix_dense(const float2* X, const float2* Wm, float2* Y, int modes, int B, int Ci, int Co,
                     cudaStream_t st, MixWideOpts im) {
  DLWP_REQUIRE(modes > 0 && B > 0 && Ci > 0 && Co > 0, "mix_dense: empty problem");
  DLWP_REQUIRE(!im.img || (Co == 128 && im.Kc > 0 && modes % im.Kc == 0 && modes / im.Kc <= 64),
               "mix_dense: the image output is laid out for 128 channels and at most 64 retained rows");
  const size_t smem_few = (size_t)(4 * Ci + 32 * Co) * sizeof(float2);
  if (B <= 4 && Co % 2 == 0 && Ci >= 8 && smem_few <= 48 * 1024) {
    if (im.w16 && Co % 4 == 0) mix_dense_fewb16_kernel<<<dim3((unsigned)modes), dim3(32, 8), smem_few, st>>>(X, Y, B, Ci, Co, im);
    else mix_dense_fewb_kernel<<<dim3((unsigned)modes), dim3(32, 8), smem_few, st>>>(X, Wm, Y, B, Ci, Co, im);
    DLWP_LAUNCHED("mix_dense_kernel");
    return DLWP_OK;
  }
  const size_t smem = (size_t)8 * Ci * sizeof(float2);
  DLWP_REQUIRE(smem <= 48 * 1024, "mix_dense: Ci=%d too large for the fp32 path", Ci);
  DLWP_REQUIRE(ceil_div(B, 8) < 65536, "mix_dense: batch too large");
  dim3 grid((unsigned)modes, (unsigned)ceil_div(B, 8));
  mix_dense_kernel<<<grid, dim3(32, 8), smem, st>>>(X, Wm, Y, B, Ci, Co, im);
  DLWP_LAUNCHED("mix_dense_kernel");
  return DLWP_OK;
}

// =====================================================================================
// mix_afno_kernel: the AFNO2D block-diagonal complex MLP on one (mode, sample):
//   o1 = relu_re_im(x W1 + b1);  o2 = o1 W2 + b2;  y = softshrink_re_im(o2)
//   fourcastnet.py:96-121.  Channel c = k * bs + o with k the block index.
// =====================================================================================
__global__ void __launch_bounds__(256) mix_afno_kernel(const float2* __restrict__ X, float2* __restrict__ Y,
                                                       const float* __restrict__ w1, const float* __restrict__ b1,
                                                       const float* __restrict__ w2, const float* __restrict__ b2,
                                                       int nb, int bs, int f, float lambda) {
  extern __shared__ float2 s_buf[];  // x [C] then o1 [C * f]
  const int C = nb * bs, Ch = C * f, bsf = bs * f;
  float2* s_xv = s_buf;
  float2* s_o1 = s_buf + C;
  const long long row = blockIdx.x;  // (mode, sample) flattened
  const float2* __restrict__ x = X + row * C;
  for (int c = threadIdx.x; c < C; c += blockDim.x) s_xv[c] = x[c];
  __syncthreads();
  // layer 1: w1 [2, nb, bs, bs*f], b1 [2, nb, bs*f]
  const long long w1_im = (long long)nb * bs * bsf, b1_im = (long long)nb * bsf;
  for (int hc = threadIdx.x; hc < Ch; hc += blockDim.x) {
    const int k = hc / bsf, j = hc - k * bsf;
    float ar = b1[(long long)k * bsf + j], ai = b1[b1_im + (long long)k * bsf + j];
    const float* wr = w1 + ((long long)k * bs) * bsf + j;
    const float* wi = wr + w1_im;
    const float2* xv = s_xv + k * bs;
    for (int i = 0; i < bs; ++i) {
      const float a = wr[(long long)i * bsf], bq = wi[(long long)i * bsf];
      ar = fmaf(xv[i].x, a, ar);
      ar = fmaf(-xv[i].y, bq, ar);
      ai = fmaf(xv[i].y, a, ai);
      ai = fmaf(xv[i].x, bq, ai);
    }
    s_o1[hc] = make_float2(fmaxf(ar, 0.f), fmaxf(ai, 0.f));
  }
  __syncthreads();
  // layer 2: w2 [2, nb, bs*f, bs], b2 [2, nb, bs]
  const long long w2_im = (long long)nb * bsf * bs, b2_im = (long long)nb * bs;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    const int k = c / bs, o = c - k * bs;
    float ar = b2[(long long)k * bs + o], ai = b2[b2_im + (long long)k * bs + o];
    const float* wr = w2 + ((long long)k * bsf) * bs + o;
    const float* wi = wr + w2_im;
    const float2* hv = s_o1 + k * bsf;
    for (int j = 0; j < bsf; ++j) {
      const float a = wr[(long long)j * bs], bq = wi[(long long)j * bs];
      ar = fmaf(hv[j].x, a, ar);
      ar = fmaf(-hv[j].y, bq, ar);
      ai = fmaf(hv[j].y, a, ai);
      ai = fmaf(hv[j].x, bq, ai);
    }
    Y[row * C + c] = make_float2(softshrink(ar, lambda), softshrink(ai, lambda));
  }
}

int launch_mix_afno(const float2* X, float2* Y, const float* w1, const float* b1, const float* w2,
                    const float* b2, int modes, int B, int nb, int bs, int f, float lambda, cudaStream_t st) {
  DLWP_REQUIRE(modes > 0 && B > 0 && nb > 0 && bs > 0 && f > 0, "mix_afno: empty problem");
  const int C = nb * bs;
  const size_t smem = (size_t)C * (1 + f) * sizeof(float2);
  DLWP_REQUIRE(smem <= 48 * 1024, "mix_afno: embed_dim=%d too large for the fp32 path", C);
  const long long rows = (long long)modes * B;
  DLWP_REQUIRE(rows < (1ll << 31), "mix_afno: too many (mode, sample) rows");
  int threads = ((C < 256 ? C : 256) + 31) / 32 * 32;
  mix_afno_kernel<<<(unsigned)rows, threads, smem, st>>>(X, Y, w1, b1, w2, b2, nb, bs, f, lambda);
  DLWP_LAUNCHED("mix_afno_kernel");
  return DLWP_OK;
}

// =====================================================================================
// Weight packing.  Reference layout [Ci, Co, R, Kc] (unet.py:38-44; neuralop/tltorch dense
// tensor) -> mode-major [R, Kc, Ci, Co] so one mode's Ci x Co matrix is contiguous.
// =====================================================================================
__global__ void pack_modes_kernel(const float2* __restrict__ w, float2* __restrict__ wm, int Ci, int Co,
                                  int R, int Kc) {
  const long long total = (long long)Ci * Co * R * Kc;
  for (long long d = blockIdx.x * (long long)blockDim.x + threadIdx.x; d < total;
       d += (long long)gridDim.x * blockDim.x) {
    // d indexes the destination [r][k][i][o]
    long long t = d;
    const int o = (int)(t % Co); t /= Co;
    const int i = (int)(t % Ci); t /= Ci;
    const int k = (int)(t % Kc); t /= Kc;
    const int r = (int)t;
    wm[d] = w[(((long long)i * Co + o) * R + r) * Kc + k];
  }
}

int launch_pack_modes(const float2* w, float2* wm, int Ci, int Co, int R, int Kc, cudaStream_t st) {
  const long long total = (long long)Ci * Co * R * Kc;
  DLWP_REQUIRE(total > 0, "pack_modes: empty weight");
  const int blocks = (int)(ceil_div64(total, 256) < 148 * 16 ? ceil_div64(total, 256) : 148 * 16);
  pack_modes_kernel<<<blocks, 256, 0, st>>>(w, wm, Ci, Co, R, Kc);
  DLWP_LAUNCHED("pack_modes_kernel");
  return DLWP_OK;
}

// Tucker -> dense (tltorch ComplexTuckerTensor.to_tensor): four successive mode products
//   out[pre, n, post] = sum_r U[n, r] * in[pre, r, post]        (complex, fp64)
__global__ void mode_product_kernel(const double2* __restrict__ in, const float2* __restrict__ U,
                                    double2* __restrict__ out, long long pre, int n_out, int r_in,
                                    long long post) {
  const long long total = pre * n_out * post;
  for (long long d = blockIdx.x * (long long)blockDim.x + threadIdx.x; d < total;
       d += (long long)gridDim.x * blockDim.x) {
    const long long q = d % post;
    const long long t = d / post;
    const int n = (int)(t % n_out);
    const long long p = t / n_out;
    double ar = 0.0, ai = 0.0;
    for (int r = 0; r < r_in; ++r) {
      const float2 u = U[(long long)n * r_in + r];
      const double2 v = in[(p * r_in + r) * post + q];
      ar += (double)u.x * v.x - (double)u.y * v.y;
      ai += (double)u.x * v.y + (double)u.y * v.x;
    }
    out[d] = make_double2(ar, ai);
  }
}

__global__ void widen_kernel(const float2* __restrict__ in, double2* __restrict__ out, long long n) {
  for (long long d = blockIdx.x * (long long)blockDim.x + threadIdx.x; d < n;
       d += (long long)gridDim.x * blockDim.x)
    out[d] = make_double2((double)in[d].x, (double)in[d].y);
}

__global__ void narrow_kernel(const double2* __restrict__ in, float2* __restrict__ out, long long n) {
  for (long long d = blockIdx.x * (long long)blockDim.x + threadIdx.x; d < n;
       d += (long long)gridDim.x * blockDim.x)
    out[d] = make_float2((float)in[d].x, (float)in[d].y);
}

size_t tucker_scratch_elems(const int dims[4], const int ranks[4]) {
  // intermediates: [r0 r1 r2 r3] -> [d0 r1 r2 r3] -> [d0 d1 r2 r3] -> [d0 d1 d2 r3] -> dense
  size_t best = 0, cur;
  int shape[4] = {ranks[0], ranks[1], ranks[2], ranks[3]};
  for (int s = 0; s <= 4; ++s) {
    cur = (size_t)shape[0] * shape[1] * shape[2] * shape[3];
    if (cur > best) best = cur;
    if (s < 4) shape[s] = dims[s];
  }
  return best;
}

static inline int grid_for(long long total) {
  const long long b = ceil_div64(total, 256);
  return (int)(b < 148 * 16 ? (b > 0 ? b : 1) : 148 * 16);
}

int launch_tucker_to_dense(const float2* core, const float2* const U[4], const int dims[4],
                           const int ranks[4], double2* scratch, float2* dense, cudaStream_t st) {
  for (int k = 0; k < 4; ++k) DLWP_REQUIRE(dims[k] > 0 && ranks[k] > 0, "tucker: bad dims");
  const size_t half = tucker_scratch_elems(dims, ranks);
  double2* buf[2] = {scratch, scratch + half};
  long long n = (long long)ranks[0] * ranks[1] * ranks[2] * ranks[3];
  widen_kernel<<<grid_for(n), 256, 0, st>>>(core, buf[0], n);
  DLWP_LAUNCHED("widen_kernel");
  int shape[4] = {ranks[0], ranks[1], ranks[2], ranks[3]};
  int cur = 0;
  for (int k = 0; k < 4; ++k) {
    long long pre = 1, post = 1;
    for (int q = 0; q < k; ++q) pre *= shape[q];
    for (int q = k + 1; q < 4; ++q) post *= shape[q];
    const long long total = pre * dims[k] * post;
    mode_product_kernel<<<grid_for(total), 256, 0, st>>>(buf[cur], U[k], buf[cur ^ 1], pre, dims[k], shape[k],
                                                          post);
    DLWP_LAUNCHED("mode_product_kernel");
    shape[k] = dims[k];
    cur ^= 1;
  }
  n = (long long)dims[0] * dims[1] * dims[2] * dims[3];
  narrow_kernel<<<grid_for(n), 256, 0, st>>>(buf[cur], dense, n);
  DLWP_LAUNCHED("narrow_kernel");
  return DLWP_OK;
}

// =====================================================================================
// NCHW <-> NHWC (only the stand-alone dlwp_spectral_conv2d and dlwp_step with a
// pre-concatenated input use these; the rollout gathers NCHW planes inside pw_linear).
// =====================================================================================
__global__ void __launch_bounds__(256) transpose_kernel(const float* __restrict__ in, float* __restrict__ out,
                                                        int rows, int cols) {
  // per batch (blockIdx.z): in [rows, cols] -> out [cols, rows]
  __shared__ float tile[32][33];
  const long long base = (long long)blockIdx.z * rows * cols;
  const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
  for (int j = threadIdx.y; j < 32; j += 8) {
    const int r = r0 + j, c = c0 + threadIdx.x;
    if (r < rows && c < cols) tile[j][threadIdx.x] = in[base + (long long)r * cols + c];
  }
  __syncthreads();
  for (int j = threadIdx.y; j < 32; j += 8) {
    const int c = c0 + j, r = r0 + threadIdx.x;
    if (r < rows && c < cols) out[base + (long long)c * rows + r] = tile[threadIdx.x][j];
  }
}

static int launch_transpose(const float* in, float* out, int B, int rows, int cols, cudaStream_t st) {
  DLWP_REQUIRE(B > 0 && rows > 0 && cols > 0, "transpose: empty problem");
  DLWP_REQUIRE(B < 65536 && ceil_div(rows, 32) < 65536, "transpose: grid too large");
  dim3 grid((unsigned)ceil_div(cols, 32), (unsigned)ceil_div(rows, 32), (unsigned)B);
  transpose_kernel<<<grid, dim3(32, 8), 0, st>>>(in, out, rows, cols);
  DLWP_LAUNCHED("transpose_kernel");
  return DLWP_OK;
}

int launch_nchw_to_nhwc(const float* in, float* out, int B, int C, int HW, cudaStream_t st) {
  return launch_transpose(in, out, B, C, HW, st);
}
int launch_nhwc_to_nchw(const float* in, float* out, int B, int C, int HW, cudaStream_t st) {
  return launch_transpose(in, out, B, HW, C, st);
}

// =====================================================================================
// Latitude-weighted error sums, scripts/evaluate.py:790-820.  One block per (t, c, chunk of
// samples, slice of the plane); fp64 accumulation; one atomicAdd per block and statistic.
// =====================================================================================
// VEC = 4: W % 4 == 0 and 16-byte aligned planes, one float4 per thread and plane position, the loads of four
// samples in flight together (PU = false) or, with few samples per CTA, of four plane positions (PU = true);
// VEC = 1: any shape.
template <int VEC, bool PU = false>
__global__ void __launch_bounds__(256) metrics_kernel(const float* __restrict__ o, const float* __restrict__ t,
                                                      const float* __restrict__ clim,
                                                      const float* __restrict__ latw, int B, int T, int C, int H,
                                                      int W, double* __restrict__ sums) {
  const int tc = blockIdx.x;  // t * C + c
  const int ti = tc / C, c = tc - ti * C;
  const int HW = H * W;
  int slice = (HW + gridDim.z - 1) / gridDim.z;
  slice = (slice + VEC - 1) / VEC * VEC;
  const int p_lo = blockIdx.z * slice, p_hi = p_lo + slice < HW ? p_lo + slice : HW;
  const long long bstride = (long long)T * C * HW;
  const long long plane = ((long long)ti * C + c) * HW;
  double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
  auto acc = [&](float w, float ov, float tv, float cv) {
    const double d = (double)ov - (double)tv, oa = (double)ov - (double)cv, ta = (double)tv - (double)cv;
    s0 += (double)w * d * d;
    s1 += (double)w * oa * ta;
    s2 += (double)w * oa * oa;
    s3 += (double)w * ta * ta;
  };
  if constexpr (VEC == 4 && PU) {
    constexpr int U = 4;
    const int pstep = blockDim.x * 4;
    for (int b = blockIdx.y; b < B; b += gridDim.y) {
      const long long base = b * bstride + plane;
      for (int p = p_lo + threadIdx.x * 4; p < p_hi; p += pstep * U) {
        float4 ov[U], tv[U], cv[U];
        float w[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int pp = p + u * pstep;
          cv[u] = make_float4(0.f, 0.f, 0.f, 0.f);
          if (pp < p_hi) {
            ov[u] = __ldcs(reinterpret_cast<const float4*>(o + base + pp));
            tv[u] = __ldcs(reinterpret_cast<const float4*>(t + base + pp));
            if (clim) cv[u] = *reinterpret_cast<const float4*>(clim + (long long)c * HW + pp);
            w[u] = latw[pp / W];
          }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          if (p + u * pstep < p_hi) {
            acc(w[u], ov[u].x, tv[u].x, cv[u].x);
            acc(w[u], ov[u].y, tv[u].y, cv[u].y);
            acc(w[u], ov[u].z, tv[u].z, cv[u].z);
            acc(w[u], ov[u].w, tv[u].w, cv[u].w);
          }
        }
      }
    }
  } else if constexpr (VEC == 4) {
    constexpr int U = 4;
    for (int p = p_lo + threadIdx.x * 4; p < p_hi; p += blockDim.x * 4) {
      const float w = latw[p / W];
      float4 cv = make_float4(0.f, 0.f, 0.f, 0.f);
      if (clim) cv = *reinterpret_cast<const float4*>(clim + (long long)c * HW + p);
      for (int b = blockIdx.y; b < B; b += gridDim.y * U) {
        float4 ov[U], tv[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int bb = b + u * gridDim.y;
          if (bb < B) {
            ov[u] = __ldcs(reinterpret_cast<const float4*>(o + bb * bstride + plane + p));
            tv[u] = __ldcs(reinterpret_cast<const float4*>(t + bb * bstride + plane + p));
          }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          if (b + u * gridDim.y < B) {
            acc(w, ov[u].x, tv[u].x, cv.x);
            acc(w, ov[u].y, tv[u].y, cv.y);
            acc(w, ov[u].z, tv[u].z, cv.z);
            acc(w, ov[u].w, tv[u].w, cv.w);
          }
        }
      }
    }
  } else {
    for (int b = blockIdx.y; b < B; b += gridDim.y) {
      const long long base = b * bstride + plane;
      for (int p = p_lo + threadIdx.x; p < p_hi; p += blockDim.x) {
        const float cv = clim ? clim[(long long)c * HW + p] : 0.f;
        acc(latw[p / W], o[base + p], t[base + p], cv);
      }
    }
  }
  __shared__ double red[4][8];
  double v[4] = {s0, s1, s2, s3};
#pragma unroll
  for (int k = 0; k < 4; ++k) {
#pragma unroll
    for (int off = 16; off; off >>= 1) v[k] += __shfl_xor_sync(0xffffffffu, v[k], off);
    if ((threadIdx.x & 31) == 0) red[k][threadIdx.x >> 5] = v[k];
  }
  __syncthreads();
  if (threadIdx.x < 4) {
    double a = 0.0;
    for (int wdx = 0; wdx < 8; ++wdx) a += red[threadIdx.x][wdx];
    atomicAdd(&sums[(long long)tc * 4 + threadIdx.x], a);
  }
}

int launch_metrics(const float* o, const float* t, const float* clim, const float* latw, int B, int T, int C,
                   int H, int W, double* sums, cudaStream_t st) {
  DLWP_REQUIRE(B > 0 && T > 0 && C > 0 && H > 0 && W > 0, "metrics: empty problem");
  DLWP_REQUIRE((long long)T * C < (1ll << 31), "metrics: too many (lead time, variable) pairs");
  int by = B < 64 ? B : 64;
  // many samples: fewer, longer CTAs (a CTA's fixed part -- start, block reduction, four fp64 atomics -- is paid per 64 KB otherwise)
  while (by > 16 && (long long)T * C * (by / 2) >= 2 * 1184) by /= 2;
  // large planes with few samples: slice the plane so that the grid still covers the machine
  long long bz = (1184 + (long long)T * C * by - 1) / ((long long)T * C * by);
  const long long bz_max = ((long long)H * W + 4095) / 4096;
  if (bz > bz_max) bz = bz_max;
  if (bz < 1) bz = 1;
  if (bz > 65535) bz = 65535;
  dim3 grid((unsigned)(T * C), (unsigned)by, (unsigned)bz);
  const bool vec = W % 4 == 0 && (((uintptr_t)o | (uintptr_t)t | (uintptr_t)clim) & 15) == 0;
  if (vec && B < 4 * by)      // fewer than four samples per CTA: keep four plane positions in flight instead
    metrics_kernel<4, true><<<grid, 256, 0, st>>>(o, t, clim, latw, B, T, C, H, W, sums);
  else if (vec)
    metrics_kernel<4><<<grid, 256, 0, st>>>(o, t, clim, latw, B, T, C, H, W, sums);
  else
    metrics_kernel<1><<<grid, 256, 0, st>>>(o, t, clim, latw, B, T, C, H, W, sums);
  DLWP_LAUNCHED("metrics_kernel");
  return DLWP_OK;
}

}  // namespace dlwp
