// Stand-alone spectral operators of the in-repo UNet family on the fp32 kernel set:
//   dlwp_spectral_conv2d          SpectralConv2d.forward              (models/unet/unet.py:19-69), any Ci / Co
//   dlwp_fourier_residual_block   FourierResidualBlock.forward        (models/unet/unet.py:780-836)
// Both run on channels-last copies inside ONE workspace block (the caller's, or one cudaMalloc per call when none is
// given) through the truncated-DFT chain that the rollout plans use (plan.cu:spectral_layer): W-DFT, H-DFT, per-mode
// channel mixing, inverse H-DFT, inverse W-DFT with the additions / bias / GELU fused into its epilogue.
#include <cmath>

#include "kernels.cuh"
#include "spectral_chain.cuh"

namespace dlwp {

namespace {

__device__ __forceinline__ float2 twiddle(long long k, long long n, long long N, double sign, double scale) {
  long long r = (k * n) % N;
  if (r < 0) r += N;
  double s, c;
  sincospi(2.0 * (double)r / (double)N, &s, &c);
  return make_float2((float)(scale * c), (float)(scale * sign * s));
}

// The truncated-DFT bases of plan.cu:build_bases, generated on the device in fp64 (no host staging copy):
//   Fw [Kc, W] = s exp(-2 pi i k w / W)          Fh [R, H] = exp(-2 pi i rows_in[j] h / H)
//   Gh [H, R]  = exp(+2 pi i rows_out[j] h / H)  Gw [W, Kc] = t c_k exp(+2 pi i k w / W), c = 1 for the DC / Nyquist column, else 2
__global__ void dft_bases_kernel(float2* Fw, float2* Fh, float2* Gh, float2* Gw, int H, int W, int R, int Kc, double s, double t,
                                 SpecRows rows) {
  const int n_fw = Kc * W, n_fh = R * H, total = 2 * n_fw + 2 * n_fh;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    if (i < n_fw) {
      const int k = i / W, w = i - k * W;
      Fw[i] = twiddle(k, w, W, -1.0, s);
    } else if (i < n_fw + n_fh) {
      const int j = (i - n_fw) / H, h = (i - n_fw) - j * H;
      Fh[i - n_fw] = twiddle(rows.in[j], h, H, -1.0, 1.0);
    } else if (i < n_fw + 2 * n_fh) {
      const int e = i - n_fw - n_fh, h = e / R, j = e - h * R;
      Gh[e] = twiddle(rows.out[j], h, H, +1.0, 1.0);
    } else {
      const int e = i - n_fw - 2 * n_fh, w = e / Kc, k = e - w * Kc;
      const bool self_conj = (k == 0) || (W % 2 == 0 && k == W / 2);
      Gw[e] = twiddle(k, w, W, +1.0, t * (self_conj ? 1.0 : 2.0));
    }
  }
}

inline size_t up256(size_t v) { return (v + 255) & ~(size_t)255; }

}  // namespace

size_t SpecChain::carve(unsigned char* base) {
  const size_t Bn = (size_t)B, Rn = (size_t)R, K = (size_t)Kc;
  const size_t sizes[9] = {K * W * sizeof(float2),      Rn * H * sizeof(float2),      (size_t)H * Rn * sizeof(float2),
                           (size_t)W * K * sizeof(float2), Bn * H * K * Ci * sizeof(float2), Rn * K * Bn * Ci * sizeof(float2),
                           Rn * K * Bn * Co * sizeof(float2), Bn * H * K * Co * sizeof(float2), (size_t)Ci * Co * Rn * K * sizeof(float2)};
  float2** dst[9] = {&Fw, &Fh, &Gh, &Gw, &A, &X, &Y, &Z, &wm};
  size_t off = 0;
  for (int i = 0; i < 9; ++i) {
    *dst[i] = reinterpret_cast<float2*>(base ? base + off : nullptr);
    off += up256(sizes[i]);
  }
  return off;
}

int SpecChain::check(const int32_t* rows_in, const int32_t* rows_out, int fft_norm, const char* who) {
  DLWP_REQUIRE(B > 0 && Ci > 0 && Co > 0 && H > 0 && W > 0, "%s: empty problem", who);
  DLWP_REQUIRE(R >= 1 && R <= kSpecMaxRows && R <= H, "%s: n_rows=%d outside [1, min(%d, H=%d)]", who, R, kSpecMaxRows, H);
  DLWP_REQUIRE(Kc >= 1 && Kc <= W / 2 + 1, "%s: n_cols=%d outside [1, W/2+1=%d]", who, Kc, W / 2 + 1);
  for (int j = 0; j < R; ++j) {
    DLWP_REQUIRE(rows_in[j] >= 0 && rows_in[j] < H && rows_out[j] >= 0 && rows_out[j] < H, "%s: row bin out of range", who);
    rows.in[j] = rows_in[j];
    rows.out[j] = rows_out[j];
  }
  s = t = 1.0;
  const double hw = (double)H * (double)W;
  switch (fft_norm) {
    case DLWP_FFT_NORM_BACKWARD: t = 1.0 / hw; break;
    case DLWP_FFT_NORM_FORWARD: s = 1.0 / hw; break;
    case DLWP_FFT_NORM_ORTHO: s = t = 1.0 / std::sqrt(hw); break;
    default: return fail(DLWP_ERR_INVALID, "%s: unknown fft_norm %d", who, fft_norm);
  }
  return DLWP_OK;
}

int SpecChain::bases(cudaStream_t st) {
  const int total = 2 * Kc * W + 2 * R * H;
  dft_bases_kernel<<<ceil_div(total, 256), 256, 0, st>>>(Fw, Fh, Gh, Gw, H, W, R, Kc, s, t, rows);
  DLWP_LAUNCHED("dft_bases_kernel");
  return DLWP_OK;
}

// yl = act( irfft2( mix( rfft2(xl)[rows_in, :Kc] ) -> [rows_out, :Kc] ) + add1 + add2 + bias ), channels-last; wm = packed weights
int SpecChain::run(const float* xl, float* yl, const float* add1, const float* add2, const float* bias, int act, cudaStream_t st) {
  const long long HW = (long long)H * W, K = Kc;
  TdftArgs a{};
  a.in = xl; a.out = A; a.T = Fw; a.N = W; a.K = Kc; a.C = Ci; a.nb0 = B; a.nb1 = H;      // W axis: x [B,H,W,Ci] -> A [B,H,Kc,Ci]
  a.in_sn = Ci; a.in_sb0 = HW * Ci; a.in_sb1 = (long long)W * Ci;
  a.out_sk = Ci; a.out_sb0 = (long long)H * K * Ci; a.out_sb1 = K * Ci;
  a.in_complex = false; a.out_complex = true;
  DLWP_TRY(launch_tdft(a, st));
  a = TdftArgs{};
  a.in = A; a.out = X; a.T = Fh; a.N = H; a.K = R; a.C = Ci; a.nb0 = B; a.nb1 = Kc;         // H axis: A -> X [R,Kc,B,Ci]
  a.in_sn = K * Ci; a.in_sb0 = (long long)H * K * Ci; a.in_sb1 = Ci;
  a.out_sk = K * B * Ci; a.out_sb0 = Ci; a.out_sb1 = (long long)B * Ci;
  a.in_complex = true; a.out_complex = true;
  DLWP_TRY(launch_tdft(a, st));
  DLWP_TRY(launch_mix_dense(X, wm, Y, R * Kc, B, Ci, Co, st));                               // einsum("bixy,ioxy->boxy"), unet.py:15-17
  a = TdftArgs{};
  a.in = Y; a.out = Z; a.T = Gh; a.N = R; a.K = H; a.C = Co; a.nb0 = B; a.nb1 = Kc;         // inverse H: Y -> Z [B,H,Kc,Co]
  a.in_sn = K * B * Co; a.in_sb0 = Co; a.in_sb1 = (long long)B * Co;
  a.out_sk = K * Co; a.out_sb0 = (long long)H * K * Co; a.out_sb1 = Co;
  a.in_complex = true; a.out_complex = true;
  DLWP_TRY(launch_tdft(a, st));
  a = TdftArgs{};
  a.in = Z; a.out = yl; a.T = Gw; a.N = Kc; a.K = W; a.C = Co; a.nb0 = B; a.nb1 = H;        // inverse W (C2R): Z -> y [B,H,W,Co]
  a.in_sn = Co; a.in_sb0 = (long long)H * K * Co; a.in_sb1 = K * Co;
  a.out_sk = Co; a.out_sb0 = HW * Co; a.out_sb1 = (long long)W * Co;
  a.in_complex = true; a.out_complex = false;
  a.add1 = add1; a.add2 = add2; a.bias = bias; a.act = act;
  DLWP_TRY(launch_tdft(a, st));
  return DLWP_OK;
}

}  // namespace dlwp

// =============================================================================================
// public entries
// =============================================================================================
namespace dlwp {
namespace {

// out = gelu(x) elementwise (activation of unet.py:826,830 without normalisation)
__global__ void gelu_kernel(const float* __restrict__ x, float* __restrict__ y, long long n) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) y[i] = gelu_erf(x[i]);
}

// out = gelu(GroupNorm(x)) on channels-last [B, HW, C]: one block per (sample, group); statistics over HW * C / G elements
// (torch.nn.GroupNorm: biased variance, eps inside the square root, per-channel affine)
__global__ void __launch_bounds__(256) groupnorm_gelu_kernel(const float* __restrict__ x, float* __restrict__ y, const float* __restrict__ gamma,
                                                             const float* __restrict__ beta, int HW, int C, int G, float eps) {
  __shared__ float red[8];
  __shared__ float stat[2];
  const int b = blockIdx.x / G, g = blockIdx.x - b * G, cg = C / G, n = HW * cg;
  const float* xb = x + (size_t)b * HW * C + (size_t)g * cg;
  float* yb = y + (size_t)b * HW * C + (size_t)g * cg;
  auto block_sum = [&](float v) {
#pragma unroll
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    float s = 0.f;
    for (int w = 0; w < 8; ++w) s += red[w];
    return s;
  };
  float s = 0.f;
  for (int i = threadIdx.x; i < n; i += 256) s += xb[(size_t)(i / cg) * C + (i % cg)];
  const float mean = block_sum(s) / (float)n;
  float q = 0.f;
  for (int i = threadIdx.x; i < n; i += 256) {
    const float d = xb[(size_t)(i / cg) * C + (i % cg)] - mean;
    q = fmaf(d, d, q);
  }
  const float rstd = rsqrtf(block_sum(q) / (float)n + eps);
  (void)stat;
  for (int i = threadIdx.x; i < n; i += 256) {
    const int c = i % cg;
    const size_t o = (size_t)(i / cg) * C + c;
    yb[o] = gelu_erf((xb[o] - mean) * rstd * gamma[g * cg + c] + beta[g * cg + c]);
  }
}

int act_norm(const float* in, float* out, const float* gamma, const float* beta, int B, int HW, int C, int G, float eps, cudaStream_t st) {
  if (gamma == nullptr) {
    const long long n = (long long)B * HW * C;
    gelu_kernel<<<(unsigned)((n + 1023) / 1024 < 148 * 8 ? (n + 1023) / 1024 : 148 * 8), 256, 0, st>>>(in, out, n);
    DLWP_LAUNCHED("gelu_kernel");
    return DLWP_OK;
  }
  DLWP_REQUIRE(G >= 1 && C % G == 0 && beta != nullptr, "GroupNorm: %d channels do not split into %d groups (or the bias is missing)", C, G);
  groupnorm_gelu_kernel<<<B * G, 256, 0, st>>>(in, out, gamma, beta, HW, C, G, eps);
  DLWP_LAUNCHED("groupnorm_gelu_kernel");
  return DLWP_OK;
}

PwArgs pw(const float* in, const float* Wt, const float* bias, float* out, int P, int K, int N, int tokens) {
  PwArgs a{};
  a.in = in; a.Wt = Wt; a.bias = bias; a.out = out; a.P = P; a.K = K; a.N = N; a.tokens = tokens; a.p1 = a.p2 = 1;
  return a;
}

inline size_t up256b(size_t v) { return (v + 255) & ~(size_t)255; }

}  // namespace
}  // namespace dlwp

using namespace dlwp;

extern "C" int dlwp_spectral_conv2d(const float* x, const float* weight, const float* bias, float* y, int batch, int Ci, int Co, int H,
                                    int W, int n_rows, int n_cols, const int32_t* rows_in, const int32_t* rows_out, int fft_norm,
                                    void* stream) {
  if (!x || !weight || !y || !rows_in || !rows_out) return fail(DLWP_ERR_INVALID, "dlwp_spectral_conv2d: null argument");
  SpecChain c;
  c.B = batch; c.Ci = Ci; c.Co = Co; c.H = H; c.W = W; c.R = n_rows; c.Kc = n_cols;
  DLWP_TRY(c.check(rows_in, rows_out, fft_norm, "dlwp_spectral_conv2d"));
  const size_t P = (size_t)batch * H * W, chain = c.carve(nullptr);
  const size_t total = chain + up256b(P * Ci * 4) + up256b(P * Co * 4);
  unsigned char* ws = nullptr;      // a convenience entry: one allocation per call (dlwp_spectral_conv2d_static allocates nothing)
  DLWP_CUDA(cudaMalloc((void**)&ws, total));
  c.carve(ws);
  float *xl = (float*)(ws + chain), *yl = (float*)(ws + chain + up256b(P * Ci * 4));
  cudaStream_t st = (cudaStream_t)stream;
  int rc = c.bases(st);
  if (rc == DLWP_OK) rc = launch_pack_modes(reinterpret_cast<const float2*>(weight), c.wm, Ci, Co, n_rows, n_cols, st);
  if (rc == DLWP_OK) rc = launch_nchw_to_nhwc(x, xl, batch, Ci, H * W, st);
  if (rc == DLWP_OK) rc = c.run(xl, yl, nullptr, nullptr, bias, 0, st);
  if (rc == DLWP_OK) rc = launch_nhwc_to_nchw(yl, y, batch, Co, H * W, st);
  cudaStreamSynchronize(st);
  cudaFree(ws);
  return rc;
}

namespace dlwp {
namespace {
struct FrbPlan {
  SpecChain c1, c2;
  size_t total = 0, o_wm2 = 0, o_act[7] = {};
};
// workspace layout: chain buffers (sized for the wider of the two layers), second weight block, seven channels-last tensors
int frb_layout(FrbPlan& f, int batch, int Ci, int Co, int H, int W, int m1, int m2, unsigned char* base) {
  DLWP_REQUIRE(m1 >= 1 && 2 * m1 <= H && m2 >= 1 && m2 <= W / 2 + 1, "dlwp_fourier_residual_block: modes %dx%d do not fit the %dx%d grid", m1,
               m2, H, W);
  int32_t rows[kSpecMaxRows];
  DLWP_REQUIRE(2 * m1 <= kSpecMaxRows, "dlwp_fourier_residual_block: at most %d retained rows", kSpecMaxRows);
  for (int j = 0; j < m1; ++j) {      // weights1: rows [:m1], weights2: rows [-m1:]  (unet.py:60-65)
    rows[j] = j;
    rows[m1 + j] = H - m1 + j;
  }
  SpecChain big;
  big.B = batch; big.Ci = Ci > Co ? Ci : Co; big.Co = Co; big.H = H; big.W = W; big.R = 2 * m1; big.Kc = m2;
  DLWP_TRY(big.check(rows, rows, DLWP_FFT_NORM_BACKWARD, "dlwp_fourier_residual_block"));
  size_t off = big.carve(base);
  f.c1 = big; f.c1.Ci = Ci;
  f.c2 = big; f.c2.Ci = Co;
  f.o_wm2 = off;
  f.c2.wm = reinterpret_cast<float2*>(base ? base + off : nullptr);
  off += up256b((size_t)big.Ci * Co * big.R * m2 * sizeof(float2));
  const size_t P = (size_t)batch * H * W;
  const int width[7] = {Ci, Ci, Co, Co, Co, Co, Co};      // xl, h, t1, o1, t2, t3, yl
  for (int i = 0; i < 7; ++i) {
    f.o_act[i] = off;
    off += up256b(P * width[i] * 4);
  }
  f.total = off;
  return DLWP_OK;
}
}  // namespace
}  // namespace dlwp

extern "C" int dlwp_fourier_residual_block_workspace(int batch, int Ci, int Co, int H, int W, int modes1, int modes2, size_t* bytes) {
  if (!bytes) return fail(DLWP_ERR_INVALID, "dlwp_fourier_residual_block_workspace: null argument");
  FrbPlan f;
  DLWP_TRY(frb_layout(f, batch, Ci, Co, H, W, modes1, modes2, nullptr));
  *bytes = f.total;
  return DLWP_OK;
}

extern "C" int dlwp_fourier_residual_block(const float* x, const dlwp_frb_weights* w, float* y, int batch, int Ci, int Co, int H, int W,
                                           int modes1, int modes2, void* workspace, size_t workspace_bytes, void* stream) {
  if (!x || !w || !y || !workspace) return fail(DLWP_ERR_INVALID, "dlwp_fourier_residual_block: null argument");
  DLWP_REQUIRE(w->f1_w1 && w->f1_w2 && w->f2_w1 && w->f2_w2 && w->conv1_w && w->conv1_b && w->conv2_w && w->conv2_b,
               "dlwp_fourier_residual_block: a weight pointer is null");
  DLWP_REQUIRE((w->short_w != nullptr) || Ci == Co, "dlwp_fourier_residual_block: in_channels != out_channels needs the shortcut convolution");
  FrbPlan f;
  unsigned char* ws = static_cast<unsigned char*>(workspace);
  DLWP_TRY(frb_layout(f, batch, Ci, Co, H, W, modes1, modes2, ws));
  DLWP_REQUIRE(workspace_bytes >= f.total, "dlwp_fourier_residual_block: workspace of %zu bytes, %zu needed", workspace_bytes, f.total);
  cudaStream_t st = (cudaStream_t)stream;
  const int HW = H * W, P = batch * HW, R = 2 * modes1, m1 = modes1, m2 = modes2;
  float* act[7];
  for (int i = 0; i < 7; ++i) act[i] = reinterpret_cast<float*>(ws + f.o_act[i]);
  float *xl = act[0], *h = act[1], *t1 = act[2], *o1 = act[3], *t2 = act[4], *t3 = act[5], *yl = act[6];
  DLWP_TRY(f.c1.bases(st));
  // the two corner weight tensors of a SpectralConv2d fill the two halves of the mode-major weight
  const size_t half1 = (size_t)m1 * m2 * Ci * Co, half2 = (size_t)m1 * m2 * Co * Co;
  DLWP_TRY(launch_pack_modes(reinterpret_cast<const float2*>(w->f1_w1), f.c1.wm, Ci, Co, m1, m2, st));
  DLWP_TRY(launch_pack_modes(reinterpret_cast<const float2*>(w->f1_w2), f.c1.wm + half1, Ci, Co, m1, m2, st));
  DLWP_TRY(launch_pack_modes(reinterpret_cast<const float2*>(w->f2_w1), f.c2.wm, Co, Co, m1, m2, st));
  DLWP_TRY(launch_pack_modes(reinterpret_cast<const float2*>(w->f2_w2), f.c2.wm + half2, Co, Co, m1, m2, st));
  (void)R;
  DLWP_TRY(launch_nchw_to_nhwc(x, xl, batch, Ci, HW, st));
  // h = act(norm1(x));  out = fourier1(h) + conv1(h)                                   (unet.py:826-829)
  DLWP_TRY(act_norm(xl, h, w->norm1_w, w->norm1_b, batch, HW, Ci, w->n_groups, w->eps, st));
  DLWP_TRY(launch_pw_linear(pw(h, w->conv1_w, w->conv1_b, t1, P, Ci, Co, HW), st));
  const bool norm2 = w->norm2_w != nullptr;      // without normalisation the GELU of :830 runs in the inverse transform's epilogue
  DLWP_TRY(f.c1.run(h, norm2 ? t2 : o1, t1, nullptr, nullptr, norm2 ? 0 : 1, st));
  if (norm2) DLWP_TRY(act_norm(t2, o1, w->norm2_w, w->norm2_b, batch, HW, Co, w->n_groups, w->eps, st));
  // out = fourier2(out) + conv2(out) + shortcut(x)                                     (unet.py:831-835)
  DLWP_TRY(launch_pw_linear(pw(o1, w->conv2_w, w->conv2_b, t2, P, Co, Co, HW), st));
  const float* sc = xl;
  if (w->short_w) {
    DLWP_TRY(launch_pw_linear(pw(xl, w->short_w, w->short_b, t3, P, Ci, Co, HW), st));
    sc = t3;
  }
  DLWP_TRY(f.c2.run(o1, yl, t2, sc, nullptr, 0, st));
  DLWP_TRY(launch_nhwc_to_nchw(yl, y, batch, Co, HW, st));
  return DLWP_OK;
}
