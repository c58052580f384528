// Launcher declarations for the fp32 CUDA-core kernel set (kernels_fp32.cu).
// All launchers enqueue on `st`, never synchronise, and return a dlwp_status.
#pragma once
#include "common.cuh"

namespace dlwp {

// ---- batched truncated-DFT product --------------------------------------------------
//   out[bt, k, c] = epi( sum_n T[k, n] * in[bt, n, c] ),   bt = b0 * nb1 + b1
// `T` is a complex K x N matrix (row-major float2).  in/out element strides are in units of
// the element type (float or float2); the channel stride is 1 on both sides.
struct TdftArgs {
  const void* in;
  void* out;
  const float2* T;
  int N, K, C;
  int nb0, nb1;
  long long in_sn, in_sb0, in_sb1;
  long long out_sk, out_sb0, out_sb1;
  bool in_complex, out_complex;
  // real-output epilogue: out = act(re(acc) + add1 + add2 + bias[c]); add* share out's indexing
  const float* add1;
  const float* add2;
  const float* bias;
  int act;  // 0 none, 1 exact GELU
};
int launch_tdft(const TdftArgs& a, cudaStream_t st);

// ---- pointwise linear (1x1 conv / Linear / patch embed / head) -----------------------
struct SrcGroup {
  const float* base;    // NCHW planes: base[b * bstride + ch * H * W + y * W + x]
  long long bstride;
  int nch;
};
struct PwArgs {
  // input: NHWC rows `in[p * K + k]`, or a gather over up to 4 NCHW channel groups with
  // p1 x p2 patches (k = (ch, i, j)) when ngrp > 0
  const float* in;
  SrcGroup grp[4];
  int ngrp;
  int H, W, p1, p2;
  // GEMM: out[p, n] = sum_k in[p, k] * Wt[n * K + k]
  const float* Wt;
  const float* bias;  // [N] or null
  int P, K, N;
  int tokens;         // tokens per sample = (H / p1) * (W / p2)
  // epilogue
  int act;            // 0 none, 1 exact GELU
  const float* pos;   // [tokens, N] or null (added after act)
  const float* resid; // NHWC [P, N] or null
  const float* resid2; // a second one (the FNO2D token mixer adds its input and the block residual)
  const float* pre;    // NHWC [P, N] or null: added BEFORE the activation (training path: spectral branch of an FNO block)
  float* preact;       // NHWC [P, N] or null: the pre-activation value is stored here too (saved for the backward pass)
  // output: NHWC `out[p * N + n]`, or (unpatch != 0) NCHW frame with n = (i, j, co):
  //   out[b * out_bstride + co*H*W + y*W + x] = v + last[b * last_bstride + same]
  float* out;
  int unpatch;
  long long out_bstride;
  const float* last;
  long long last_bstride;
  int Cg;
};
int launch_pw_linear(const PwArgs& a, cudaStream_t st);

// ---- LayerNorm over the channel dim of NHWC tokens -----------------------------------
int launch_layernorm(const float* in, float* out, const float* gamma, const float* beta, int P, int C,
                     float eps, cudaStream_t st);

// ---- per-mode channel mixing ----------------------------------------------------------
// X, Y: [modes, B, C] complex;  Wm: [modes, Ci, Co] complex
// Options of the bf16 kernel set (tc_wide.cuh).  `img` set: Y is not written; the result goes out as
// bf16 tile images per (sample, retained column) for the tensor-core inverse H-DFT, modes = rows * Kc.
// `w16` set (few-sample kernel only): the mode weights are read from this bf16 (re, im)-pair copy of Wm.
struct MixWideOpts {
  uint8_t* img = nullptr;
  int Kc = 0;
  const uint32_t* w16 = nullptr;
};
// complex fp32 -> bf16 (re, im) pairs
int launch_narrow_bf16(const float2* in, uint32_t* out, long long n, cudaStream_t st);
int launch_mix_dense(const float2* X, const float2* Wm, float2* Y, int modes, int B, int Ci, int Co,
                     cudaStream_t st, MixWideOpts im = MixWideOpts{});
// AFNO block-diagonal 2-layer complex MLP with ReLU (re, im separately) and softshrink.
// w1 [2, nb, bs, bs*f], b1 [2, nb, bs*f], w2 [2, nb, bs*f, bs], b2 [2, nb, bs]
int launch_mix_afno(const float2* X, float2* Y, const float* w1, const float* b1, const float* w2,
                    const float* b2, int modes, int B, int nb, int bs, int f, float lambda,
                    cudaStream_t st);

// ---- weight packing --------------------------------------------------------------------
// dense complex [Ci, Co, R, Kc] (float2) -> mode-major [R, Kc, Ci, Co]
int launch_pack_modes(const float2* w, float2* wm, int Ci, int Co, int R, int Kc, cudaStream_t st);
// Tucker: W[i,o,x,y] = sum core[a,b,c,d] Ui[i,a] Uo[o,b] Ux[x,c] Uy[y,d]  (fp64 accumulate).
// `scratch` holds 2 * max-intermediate double2 elements (see tucker_scratch_elems).
size_t tucker_scratch_elems(const int dims[4], const int ranks[4]);
int launch_tucker_to_dense(const float2* core, const float2* const U[4], const int dims[4],
                           const int ranks[4], double2* scratch, float2* dense, cudaStream_t st);

// ---- layout helpers ----------------------------------------------------------------------
int launch_nchw_to_nhwc(const float* in, float* out, int B, int C, int HW, cudaStream_t st);
int launch_nhwc_to_nchw(const float* in, float* out, int B, int C, int HW, cudaStream_t st);

// ---- metrics -------------------------------------------------------------------------------
int launch_metrics(const float* o, const float* t, const float* clim, const float* latw, int B, int T,
                   int C, int H, int W, double* sums, cudaStream_t st);

}  // namespace dlwp
