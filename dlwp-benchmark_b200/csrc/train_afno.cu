// Training primitives of the FourCastNet (AFNO) model (continued from train_ops.cu): what
// scripts/train.py:263-271 back-propagates through beyond the Linear layers:
//   dlwp_layernorm_forward / _backward       nn.LayerNorm(dim, eps=1e-6) of Block (fourcastnet.py:170,174)
//   dlwp_afno_filter_forward / _backward     AFNO2D (fourcastnet.py:60-127): truncated rfft2 (ortho), block-diagonal complex
//                                            two-layer MLP with ReLU, softshrink, irfft2, + input; backward = adjoint chain
//                                            (train_ops.cu) around the backward pass of the block MLP
// fp32 kernel set on channels-last tokens; the spectrum and the hidden layer are recomputed in the backward pass.
#include <cmath>

#include "kernels.cuh"
#include "spectral_chain.cuh"

namespace dlwp {
namespace {

// ---- LayerNorm backward: one warp per token, lane l owns channels l, l + 32, ... ----------------------------------------
// NJ = channels per lane held in registers for the dgamma / dbeta partial sums (C <= 32 NJ)
template <int NJ>
__global__ void __launch_bounds__(256) layernorm_bwd_kernel(const float* __restrict__ dy, const float* __restrict__ x,
                                                            const float* __restrict__ gamma, float* __restrict__ dx,
                                                            float* __restrict__ dgamma, float* __restrict__ dbeta, int P, int C, float eps) {
  __shared__ float s_red[8][32];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int warp0 = blockIdx.x * 8 + wib, nwarps = gridDim.x * 8;
  float ag[NJ], ab[NJ], gm[NJ];
#pragma unroll
  for (int j = 0; j < NJ; ++j) {
    ag[j] = ab[j] = 0.f;
    gm[j] = (j * 32 + lane < C) ? gamma[j * 32 + lane] : 0.f;
  }
  const float invC = 1.0f / (float)C;
  for (int p = warp0; p < P; p += nwarps) {
    const float* xr = x + (long long)p * C;
    const float* gr = dy + (long long)p * C;
    float xv[NJ], gv[NJ];
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < NJ; ++j) {
      const bool in = j * 32 + lane < C;
      xv[j] = in ? xr[j * 32 + lane] : 0.f;
      gv[j] = in ? gr[j * 32 + lane] : 0.f;
      s += xv[j];
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    const float mean = s * invC;
    float q = 0.f;
#pragma unroll
    for (int j = 0; j < NJ; ++j) {
      const float d = (j * 32 + lane < C) ? xv[j] - mean : 0.f;
      xv[j] = d;
      q = fmaf(d, d, q);
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
    const float rstd = rsqrtf(q * invC + eps);
    float s1 = 0.f, s2 = 0.f;      // sum of g, sum of g * xhat with g = dy * gamma
#pragma unroll
    for (int j = 0; j < NJ; ++j) {
      xv[j] *= rstd;               // xhat
      ag[j] = fmaf(gv[j], xv[j], ag[j]);
      ab[j] += gv[j];
      gv[j] *= gm[j];
      s1 += gv[j];
      s2 = fmaf(gv[j], xv[j], s2);
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) {
      s1 += __shfl_xor_sync(0xffffffffu, s1, o);
      s2 += __shfl_xor_sync(0xffffffffu, s2, o);
    }
    s1 *= invC;
    s2 *= invC;
    if (dx != nullptr) {
      float* dr = dx + (long long)p * C;
#pragma unroll
      for (int j = 0; j < NJ; ++j)
        if (j * 32 + lane < C) dr[j * 32 + lane] = rstd * (gv[j] - s1 - xv[j] * s2);
    }
  }
  // the block's eight partial sums per channel, then one atomic per channel and block
#pragma unroll
  for (int j = 0; j < NJ; ++j) {
    for (int pass = 0; pass < 2; ++pass) {
      __syncthreads();
      s_red[wib][lane] = pass == 0 ? ag[j] : ab[j];
      __syncthreads();
      if (wib == 0 && j * 32 + lane < C) {
        float t = 0.f;
#pragma unroll
        for (int w8 = 0; w8 < 8; ++w8) t += s_red[w8][lane];
        atomicAdd((pass == 0 ? dgamma : dbeta) + j * 32 + lane, t);
      }
    }
  }
}

// ---- backward of the AFNO2D block MLP on one (mode, sample) row -----------------------------------------------------------
//   forward (recomputed): z1 = x w1 + b1, a = relu_re_im(z1), z2 = a w2 + b2, y = softshrink_re_im(z2)
//   backward: g2 = dy * [|z2| > lambda], ga = g2 conj(w2)^T, g1 = ga * [z1 > 0], dx = g1 conj(w1)^T    (re and im separately)
// writes a (A1), g1 (G1), g2 (in place over dY) for the weight-gradient kernel and dX
__global__ void __launch_bounds__(256) afno_mlp_bwd_kernel(const float2* __restrict__ X, float2* __restrict__ dY, float2* __restrict__ A1,
                                                           float2* __restrict__ G1, float2* __restrict__ dX, const float* __restrict__ w1,
                                                           const float* __restrict__ b1, const float* __restrict__ w2,
                                                           const float* __restrict__ b2, int nb, int bs, int f, float lambda) {
  extern __shared__ float2 s_buf[];      // x [C], a [Ch], g2 [C], g1 [Ch]
  const int C = nb * bs, Ch = C * f, bsf = bs * f;
  float2* s_x = s_buf;
  float2* s_a = s_x + C;
  float2* s_g2 = s_a + Ch;
  float2* s_g1 = s_g2 + C;
  const long long row = blockIdx.x;
  for (int c = threadIdx.x; c < C; c += blockDim.x) s_x[c] = X[row * C + c];
  __syncthreads();
  const long long w1_im = (long long)nb * bs * bsf, b1_im = (long long)nb * bsf;
  for (int hc = threadIdx.x; hc < Ch; hc += blockDim.x) {
    const int k = hc / bsf, j = hc - k * bsf;
    float ar = b1[(long long)k * bsf + j], ai = b1[b1_im + (long long)k * bsf + j];
    const float* wr = w1 + ((long long)k * bs) * bsf + j;
    const float* wi = wr + w1_im;
    const float2* xv = s_x + k * bs;
    for (int i = 0; i < bs; ++i) {
      const float a = wr[(long long)i * bsf], bq = wi[(long long)i * bsf];
      ar = fmaf(xv[i].x, a, ar);
      ar = fmaf(-xv[i].y, bq, ar);
      ai = fmaf(xv[i].y, a, ai);
      ai = fmaf(xv[i].x, bq, ai);
    }
    const float2 av = make_float2(fmaxf(ar, 0.f), fmaxf(ai, 0.f));
    s_a[hc] = av;
    A1[row * Ch + hc] = av;
  }
  __syncthreads();
  const long long w2_im = (long long)nb * bsf * bs, b2_im = (long long)nb * bs;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    const int k = c / bs, o = c - k * bs;
    float ar = b2[(long long)k * bs + o], ai = b2[b2_im + (long long)k * bs + o];
    const float* wr = w2 + ((long long)k * bsf) * bs + o;
    const float* wi = wr + w2_im;
    const float2* hv = s_a + k * bsf;
    for (int j = 0; j < bsf; ++j) {
      const float a = wr[(long long)j * bs], bq = wi[(long long)j * bs];
      ar = fmaf(hv[j].x, a, ar);
      ar = fmaf(-hv[j].y, bq, ar);
      ai = fmaf(hv[j].y, a, ai);
      ai = fmaf(hv[j].x, bq, ai);
    }
    const float2 g = dY[row * C + c];
    const float2 g2 = make_float2(fabsf(ar) > lambda ? g.x : 0.f, fabsf(ai) > lambda ? g.y : 0.f);
    s_g2[c] = g2;
    dY[row * C + c] = g2;
  }
  __syncthreads();
  // ga[k, j] = sum_o g2[k, o] conj(w2[k, j, o]), masked by the ReLU
  for (int hc = threadIdx.x; hc < Ch; hc += blockDim.x) {
    const int k = hc / bsf, j = hc - k * bsf;
    const float* wr = w2 + ((long long)k * bsf + j) * bs;
    const float* wi = wr + w2_im;
    const float2* gv = s_g2 + k * bs;
    float ar = 0.f, ai = 0.f;
    for (int o = 0; o < bs; ++o) {
      const float a = wr[o], bq = wi[o];
      ar = fmaf(gv[o].x, a, ar);
      ar = fmaf(gv[o].y, bq, ar);
      ai = fmaf(gv[o].y, a, ai);
      ai = fmaf(-gv[o].x, bq, ai);
    }
    const float2 av = s_a[hc];
    const float2 g1 = make_float2(av.x > 0.f ? ar : 0.f, av.y > 0.f ? ai : 0.f);
    s_g1[hc] = g1;
    G1[row * Ch + hc] = g1;
  }
  __syncthreads();
  // dx[k, i] = sum_j g1[k, j] conj(w1[k, i, j])
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    const int k = c / bs, i = c - k * bs;
    const float* wr = w1 + ((long long)k * bs + i) * bsf;
    const float* wi = wr + w1_im;
    const float2* gv = s_g1 + k * bsf;
    float ar = 0.f, ai = 0.f;
    for (int j = 0; j < bsf; ++j) {
      const float a = wr[j], bq = wi[j];
      ar = fmaf(gv[j].x, a, ar);
      ar = fmaf(gv[j].y, bq, ar);
      ai = fmaf(gv[j].y, a, ai);
      ai = fmaf(-gv[j].x, bq, ai);
    }
    dX[row * C + c] = make_float2(ar, ai);
  }
}

// ---- weight gradients of one layer of the block MLP --------------------------------------------------------------------
//   dW[k, i, j] = sum_rows conj(P[row, k ni + i]) Q[row, k nj + j]  (real plane, then imaginary plane, as the parameters are stored)
//   db[k, j]    = sum_rows Q[row, k nj + j]
// grid (nb, row splits); a block stages 32 rows of its block's columns, a thread owns the entries e = tid + 256 q; partial sums by atomics
constexpr int AW_ROWS = 32;
template <int NQ>
__global__ void __launch_bounds__(256) afno_wgrad_kernel(const float2* __restrict__ Pm, const float2* __restrict__ Qm, float* __restrict__ dW,
                                                         float* __restrict__ db, long long rows, int per, int nb, int ni, int nj) {
  extern __shared__ float2 s_buf[];      // P tile [32][ni], Q tile [32][nj]
  float2* s_p = s_buf;
  float2* s_q = s_buf + AW_ROWS * ni;
  const int k = blockIdx.x, ne = ni * nj;
  const long long r_lo = (long long)blockIdx.y * per, r_hi = r_lo + per < rows ? r_lo + per : rows;
  const long long ldp = (long long)nb * ni, ldq = (long long)nb * nj;
  float accr[NQ], acci[NQ];
  int ei[NQ], ej[NQ];
#pragma unroll
  for (int q = 0; q < NQ; ++q) {
    accr[q] = acci[q] = 0.f;
    const int e = threadIdx.x + 256 * q;
    ei[q] = e < ne ? e / nj : 0;
    ej[q] = e < ne ? e - ei[q] * nj : 0;
  }
  float br = 0.f, bi = 0.f;
  for (long long r0 = r_lo; r0 < r_hi; r0 += AW_ROWS) {
    __syncthreads();
    for (int t = threadIdx.x; t < AW_ROWS * ni; t += 256) {
      const int rr = t / ni, i = t - rr * ni;
      s_p[t] = r0 + rr < r_hi ? Pm[(r0 + rr) * ldp + (long long)k * ni + i] : make_float2(0.f, 0.f);
    }
    for (int t = threadIdx.x; t < AW_ROWS * nj; t += 256) {
      const int rr = t / nj, j = t - rr * nj;
      s_q[t] = r0 + rr < r_hi ? Qm[(r0 + rr) * ldq + (long long)k * nj + j] : make_float2(0.f, 0.f);
    }
    __syncthreads();
#pragma unroll 4
    for (int rr = 0; rr < AW_ROWS; ++rr) {
#pragma unroll
      for (int q = 0; q < NQ; ++q) {
        const float2 p = s_p[rr * ni + ei[q]], g = s_q[rr * nj + ej[q]];
        accr[q] = fmaf(p.x, g.x, accr[q]);
        accr[q] = fmaf(p.y, g.y, accr[q]);
        acci[q] = fmaf(p.x, g.y, acci[q]);
        acci[q] = fmaf(-p.y, g.x, acci[q]);
      }
    }
    if (threadIdx.x < nj)
      for (int rr = 0; rr < AW_ROWS; ++rr) {
        br += s_q[rr * nj + threadIdx.x].x;
        bi += s_q[rr * nj + threadIdx.x].y;
      }
  }
  const long long plane = (long long)nb * ne;
#pragma unroll
  for (int q = 0; q < NQ; ++q) {
    const int e = threadIdx.x + 256 * q;
    if (e < ne) {
      atomicAdd(&dW[(long long)k * ne + e], accr[q]);
      atomicAdd(&dW[plane + (long long)k * ne + e], acci[q]);
    }
  }
  if (threadIdx.x < nj) {
    atomicAdd(&db[(long long)k * nj + threadIdx.x], br);
    atomicAdd(&db[(long long)nb * nj + (long long)k * nj + threadIdx.x], bi);
  }
}

}  // namespace
}  // namespace dlwp

using namespace dlwp;

namespace dlwp {
namespace {
inline size_t up256a(size_t v) { return (v + 255) & ~(size_t)255; }

// workspace of the AFNO2D training entries: the eight bases (forward / inverse and their adjoints), the four spectra of the
// truncated-DFT chain and the two hidden-layer buffers of the backward pass
struct AfnoTrain {
  SpecChain c;
  float2 *GwH = nullptr, *GhH = nullptr, *FhH = nullptr, *FwH = nullptr;
  float2 *A1 = nullptr, *G1 = nullptr;      // relu(z1), its gradient: [R Kc B, C f]
  size_t total = 0;
  int nb = 0, bs = 0, f = 0;
  int layout(int batch, int C, int h, int w, int R, int Kc, const int32_t* rows, int num_blocks, int factor, unsigned char* base,
             const char* who) {
    c.B = batch; c.Ci = c.Co = C; c.H = h; c.W = w; c.R = R; c.Kc = Kc;
    DLWP_TRY(c.check(rows, rows, DLWP_FFT_NORM_ORTHO, who));
    DLWP_REQUIRE(factor > 0, "%s: hidden_factor %d", who, factor);
    f = factor;
    if (num_blocks != 0) {      // 0: only the sizes are wanted
      DLWP_REQUIRE(num_blocks > 0 && C % num_blocks == 0, "%s: hidden size %d not divisible by num_blocks %d", who, C, num_blocks);
      nb = num_blocks; bs = C / nb;
      DLWP_REQUIRE((size_t)C * (2 + 2 * f) * sizeof(float2) <= 48 * 1024, "%s: hidden size %d too large for the fp32 block-MLP kernels", who, C);
      DLWP_REQUIRE(bs * f <= 256 && bs * bs * f <= 4096, "%s: block size %d x %d outside the weight-gradient kernel's range", who, bs, bs * f);
    }
    const size_t Bn = (size_t)batch, Rn = (size_t)R, K = (size_t)Kc;
    const size_t spec_bh = Bn * h * K * C * sizeof(float2), spec_r = Rn * K * Bn * C * sizeof(float2);
    const size_t sizes[14] = {K * w * 8, Rn * h * 8, (size_t)h * Rn * 8, (size_t)w * K * 8,      // Fw, Fh, Gh, Gw
                              K * w * 8, Rn * h * 8, (size_t)h * Rn * 8, (size_t)w * K * 8,      // GwH, GhH, FhH, FwH
                              spec_bh, spec_r, spec_r, spec_bh,                                   // A, X, Y, Z
                              spec_r * f, spec_r * f};                                            // A1, G1
    float2** dst[14] = {&c.Fw, &c.Fh, &c.Gh, &c.Gw, &GwH, &GhH, &FhH, &FwH, &c.A, &c.X, &c.Y, &c.Z, &A1, &G1};
    size_t off = 0;
    for (int i = 0; i < 14; ++i) {
      *dst[i] = reinterpret_cast<float2*>(base ? base + off : nullptr);
      off += up256a(sizes[i]);
    }
    total = off;
    return DLWP_OK;
  }
  // x [B, h, w, C] -> X [R, Kc, B, C]
  int forward_spectrum(const float* x, cudaStream_t st) {
    const int B = c.B, H = c.H, W = c.W, C = c.Ci, R = c.R;
    const long long K = c.Kc;
    TdftArgs a{};
    a.in = x; a.out = c.A; a.T = c.Fw; a.N = W; a.K = c.Kc; a.C = C; a.nb0 = B; a.nb1 = H;
    a.in_sn = C; a.in_sb0 = (long long)H * W * C; a.in_sb1 = (long long)W * C;
    a.out_sk = C; a.out_sb0 = (long long)H * K * C; a.out_sb1 = K * C;
    a.in_complex = false; a.out_complex = true;
    DLWP_TRY(launch_tdft(a, st));
    a = TdftArgs{};
    a.in = c.A; a.out = c.X; a.T = c.Fh; a.N = H; a.K = R; a.C = C; a.nb0 = B; a.nb1 = c.Kc;
    a.in_sn = K * C; a.in_sb0 = (long long)H * K * C; a.in_sb1 = C;
    a.out_sk = K * B * C; a.out_sb0 = C; a.out_sb1 = (long long)B * C;
    a.in_complex = true; a.out_complex = true;
    return launch_tdft(a, st);
  }
};

template <int NQ>
int launch_wgrad(const float2* Pm, const float2* Qm, float* dW, float* db, long long rows, int nb, int ni, int nj, cudaStream_t st) {
  int nsplit = ceil_div(148 * 4, nb);
  const long long max_split = (rows + 4 * AW_ROWS - 1) / (4 * AW_ROWS);
  if (nsplit > max_split) nsplit = (int)max_split;
  if (nsplit < 1) nsplit = 1;
  long long per = (rows + nsplit - 1) / nsplit;
  per = (per + AW_ROWS - 1) / AW_ROWS * AW_ROWS;
  nsplit = (int)((rows + per - 1) / per);
  const size_t smem = (size_t)AW_ROWS * (ni + nj) * sizeof(float2);
  afno_wgrad_kernel<NQ><<<dim3((unsigned)nb, (unsigned)nsplit), 256, smem, st>>>(Pm, Qm, dW, db, rows, (int)per, nb, ni, nj);
  DLWP_LAUNCHED("afno_wgrad_kernel");
  return DLWP_OK;
}
int wgrad(const float2* Pm, const float2* Qm, float* dW, float* db, long long rows, int nb, int ni, int nj, cudaStream_t st) {
  const int ne = ni * nj;
  if (ne <= 256) return launch_wgrad<1>(Pm, Qm, dW, db, rows, nb, ni, nj, st);
  if (ne <= 1024) return launch_wgrad<4>(Pm, Qm, dW, db, rows, nb, ni, nj, st);
  return launch_wgrad<16>(Pm, Qm, dW, db, rows, nb, ni, nj, st);
}
}  // namespace
}  // namespace dlwp

extern "C" int dlwp_layernorm_forward(const float* x, const float* gamma, const float* beta, float* y, int P, int C, float eps, void* stream) {
  if (!x || !gamma || !beta || !y) return fail(DLWP_ERR_INVALID, "dlwp_layernorm_forward: null argument");
  return launch_layernorm(x, y, gamma, beta, P, C, eps, (cudaStream_t)stream);
}

extern "C" int dlwp_layernorm_backward(const float* dy, const float* x, const float* gamma, float* dx, float* dgamma, float* dbeta, int P,
                                       int C, float eps, void* stream) {
  if (!dy || !x || !gamma || !dgamma || !dbeta) return fail(DLWP_ERR_INVALID, "dlwp_layernorm_backward: null argument");
  DLWP_REQUIRE(P > 0 && C > 0 && C <= 1024, "dlwp_layernorm_backward: C=%d outside [1, 1024]", C);
  cudaStream_t st = (cudaStream_t)stream;
  DLWP_CUDA(cudaMemsetAsync(dgamma, 0, (size_t)C * sizeof(float), st));
  DLWP_CUDA(cudaMemsetAsync(dbeta, 0, (size_t)C * sizeof(float), st));
  int blocks = ceil_div(P, 8 * 4);      // at least four tokens per warp: one atomic per channel and block
  if (blocks > 148 * 4) blocks = 148 * 4;
  if (blocks < 1) blocks = 1;
  const int nj = ceil_div(C, 32);
#define DLWP_LNB(NJ) layernorm_bwd_kernel<NJ><<<blocks, 256, 0, st>>>(dy, x, gamma, dx, dgamma, dbeta, P, C, eps)
  if (nj <= 2) DLWP_LNB(2);
  else if (nj <= 4) DLWP_LNB(4);
  else if (nj <= 8) DLWP_LNB(8);
  else if (nj <= 16) DLWP_LNB(16);
  else DLWP_LNB(32);
#undef DLWP_LNB
  DLWP_LAUNCHED("layernorm_bwd_kernel");
  return DLWP_OK;
}

extern "C" int dlwp_afno_filter_workspace(int batch, int C, int h, int w, int n_rows, int n_cols, int hidden_factor, size_t* bytes) {
  if (!bytes) return fail(DLWP_ERR_INVALID, "dlwp_afno_filter_workspace: null argument");
  DLWP_REQUIRE(n_rows >= 1 && n_rows <= kSpecMaxRows, "dlwp_afno_filter_workspace: n_rows=%d outside [1, %d]", n_rows, kSpecMaxRows);
  AfnoTrain t;
  int32_t rows[kSpecMaxRows] = {0};
  DLWP_TRY(t.layout(batch, C, h, w, n_rows, n_cols, rows, 0, hidden_factor, nullptr, "dlwp_afno_filter_workspace"));
  *bytes = t.total;
  return DLWP_OK;
}

extern "C" int dlwp_afno_filter_forward(const float* x, const float* extra, const float* w1, const float* b1, const float* w2,
                                        const float* b2, float* y, int batch, int C, int h, int w, int n_rows, int n_cols,
                                        const int32_t* rows, int num_blocks, int hidden_factor, float lambda, void* workspace,
                                        size_t workspace_bytes, void* stream) {
  if (!x || !w1 || !b1 || !w2 || !b2 || !y || !rows || !workspace) return fail(DLWP_ERR_INVALID, "dlwp_afno_filter_forward: null argument");
  DLWP_REQUIRE(n_rows >= 1 && n_rows <= kSpecMaxRows, "dlwp_afno_filter_forward: n_rows=%d outside [1, %d]", n_rows, kSpecMaxRows);
  AfnoTrain t;
  DLWP_TRY(t.layout(batch, C, h, w, n_rows, n_cols, rows, num_blocks, hidden_factor, static_cast<unsigned char*>(workspace), "dlwp_afno_filter_forward"));
  DLWP_REQUIRE(workspace_bytes >= t.total, "dlwp_afno_filter_forward: workspace of %zu bytes, %zu needed", workspace_bytes, t.total);
  cudaStream_t st = (cudaStream_t)stream;
  SpecChain& c = t.c;
  const long long K = n_cols;
  DLWP_TRY(c.bases(st));
  DLWP_TRY(t.forward_spectrum(x, st));
  DLWP_TRY(launch_mix_afno(c.X, c.Y, w1, b1, w2, b2, n_rows * n_cols, batch, t.nb, t.bs, t.f, lambda, st));
  TdftArgs a{};
  a.in = c.Y; a.out = c.Z; a.T = c.Gh; a.N = n_rows; a.K = h; a.C = C; a.nb0 = batch; a.nb1 = n_cols;
  a.in_sn = K * batch * C; a.in_sb0 = C; a.in_sb1 = (long long)batch * C;
  a.out_sk = K * C; a.out_sb0 = (long long)h * K * C; a.out_sb1 = C;
  a.in_complex = true; a.out_complex = true;
  DLWP_TRY(launch_tdft(a, st));
  a = TdftArgs{};
  a.in = c.Z; a.out = y; a.T = c.Gw; a.N = n_cols; a.K = w; a.C = C; a.nb0 = batch; a.nb1 = h;
  a.in_sn = C; a.in_sb0 = (long long)h * K * C; a.in_sb1 = K * C;
  a.out_sk = C; a.out_sb0 = (long long)h * w * C; a.out_sb1 = (long long)w * C;
  a.in_complex = true; a.out_complex = false;
  a.add1 = x; a.add2 = extra;      // AFNO2D's own "+ bias" (:127) and Block's second skip (:185-187)
  return launch_tdft(a, st);
}

extern "C" int dlwp_afno_filter_backward(const float* x, const float* w1, const float* b1, const float* w2, const float* b2,
                                         const float* dy, float* dx, float* dw1, float* db1, float* dw2, float* db2, int batch, int C,
                                         int h, int w, int n_rows, int n_cols, const int32_t* rows, int num_blocks, int hidden_factor,
                                         float lambda, void* workspace, size_t workspace_bytes, void* stream) {
  if (!x || !w1 || !b1 || !w2 || !b2 || !dy || !dw1 || !db1 || !dw2 || !db2 || !rows || !workspace)
    return fail(DLWP_ERR_INVALID, "dlwp_afno_filter_backward: null argument");
  DLWP_REQUIRE(n_rows >= 1 && n_rows <= kSpecMaxRows, "dlwp_afno_filter_backward: n_rows=%d outside [1, %d]", n_rows, kSpecMaxRows);
  AfnoTrain t;
  DLWP_TRY(t.layout(batch, C, h, w, n_rows, n_cols, rows, num_blocks, hidden_factor, static_cast<unsigned char*>(workspace), "dlwp_afno_filter_backward"));
  DLWP_REQUIRE(workspace_bytes >= t.total, "dlwp_afno_filter_backward: workspace of %zu bytes, %zu needed", workspace_bytes, t.total);
  cudaStream_t st = (cudaStream_t)stream;
  SpecChain& c = t.c;
  const int B = batch, R = n_rows, Kc = n_cols, nb = t.nb, bs = t.bs, f = t.f, bsf = bs * f;
  const long long K = Kc, HW = (long long)h * w, rows_n = (long long)R * Kc * B;
  DLWP_REQUIRE(rows_n < (1ll << 31), "dlwp_afno_filter_backward: too many (mode, sample) rows");
  DLWP_TRY(c.bases(st));
  DLWP_TRY(launch_adjoint_bases(c, t.GwH, t.GhH, t.FhH, t.FwH, st));
  DLWP_TRY(t.forward_spectrum(x, st));      // X, recomputed
  // adjoint of y = Re(Gw Z), Z = Gh Y:  dZ = Gw^H dy, dY = Gh^H dZ
  TdftArgs a{};
  a.in = dy; a.out = c.Z; a.T = t.GwH; a.N = w; a.K = Kc; a.C = C; a.nb0 = B; a.nb1 = h;
  a.in_sn = C; a.in_sb0 = HW * C; a.in_sb1 = (long long)w * C;
  a.out_sk = C; a.out_sb0 = (long long)h * K * C; a.out_sb1 = K * C;
  a.in_complex = false; a.out_complex = true;
  DLWP_TRY(launch_tdft(a, st));
  a = TdftArgs{};
  a.in = c.Z; a.out = c.Y; a.T = t.GhH; a.N = h; a.K = R; a.C = C; a.nb0 = B; a.nb1 = Kc;
  a.in_sn = K * C; a.in_sb0 = (long long)h * K * C; a.in_sb1 = C;
  a.out_sk = K * B * C; a.out_sb0 = C; a.out_sb1 = (long long)B * C;
  a.in_complex = true; a.out_complex = true;
  DLWP_TRY(launch_tdft(a, st));
  // block MLP: dY -> g2 in place, A1, G1, and dX into the Z buffer (free from here on)
  const size_t smem = (size_t)C * (2 + 2 * f) * sizeof(float2);
  const int threads = ((C * f < 256 ? C * f : 256) + 31) / 32 * 32;
  afno_mlp_bwd_kernel<<<(unsigned)rows_n, threads, smem, st>>>(c.X, c.Y, t.A1, t.G1, c.Z, w1, b1, w2, b2, nb, bs, f, lambda);
  DLWP_LAUNCHED("afno_mlp_bwd_kernel");
  DLWP_CUDA(cudaMemsetAsync(dw1, 0, (size_t)2 * nb * bs * bsf * sizeof(float), st));
  DLWP_CUDA(cudaMemsetAsync(db1, 0, (size_t)2 * nb * bsf * sizeof(float), st));
  DLWP_CUDA(cudaMemsetAsync(dw2, 0, (size_t)2 * nb * bsf * bs * sizeof(float), st));
  DLWP_CUDA(cudaMemsetAsync(db2, 0, (size_t)2 * nb * bs * sizeof(float), st));
  DLWP_TRY(wgrad(c.X, t.G1, dw1, db1, rows_n, nb, bs, bsf, st));
  DLWP_TRY(wgrad(t.A1, c.Y, dw2, db2, rows_n, nb, bsf, bs, st));
  if (dx == nullptr) return DLWP_OK;
  // adjoint of X = Fh A, A = Fw x:  dA = Fh^H dX, dx = Re(Fw^H dA) + dy (the "+ x" of the forward pass)
  a = TdftArgs{};
  a.in = c.Z; a.out = c.A; a.T = t.FhH; a.N = R; a.K = h; a.C = C; a.nb0 = B; a.nb1 = Kc;
  a.in_sn = K * B * C; a.in_sb0 = C; a.in_sb1 = (long long)B * C;
  a.out_sk = K * C; a.out_sb0 = (long long)h * K * C; a.out_sb1 = C;
  a.in_complex = true; a.out_complex = true;
  DLWP_TRY(launch_tdft(a, st));
  a = TdftArgs{};
  a.in = c.A; a.out = dx; a.T = t.FwH; a.N = Kc; a.K = w; a.C = C; a.nb0 = B; a.nb1 = h;
  a.in_sn = C; a.in_sb0 = (long long)h * K * C; a.in_sb1 = K * C;
  a.out_sk = C; a.out_sb0 = HW * C; a.out_sb1 = (long long)w * C;
  a.in_complex = true; a.out_complex = false;
  a.add1 = dy;
  return launch_tdft(a, st);
}
