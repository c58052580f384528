// Tensor-core (tcgen05) kernels for FourCastNet's AFNO block (fourcastnet.py:156-193) at the
// reference configuration: embed_dim 64, mlp_ratio 4 (configs/model/fourcastnet.yaml:9-12).
//
// Activations stay channels-last fp32 [tokens, 64] between kernels (the residual stream of the
// block); inside a kernel every fp32 product runs as three bf16 tensor-core products with fp32
// accumulation in tensor memory (hi*hi + lo*hi + hi*lo, see tc_fno.cuh).
//
//   afno_mlp_kernel    out = x + fc2(GELU(fc1(LayerNorm(x))))          fourcastnet.py:41-57,189-192
//                      LayerNorm statistics per token in the prologue (gamma / beta are folded into
//                      fc1 at commit), the 256-wide hidden never leaves the SM.
#pragma once
#include "kernels.cuh"

namespace dlwp {

constexpr int kAfnoE = 64;        // embed_dim of the tensor-core AFNO kernels
constexpr int kAfnoHid = 256;     // Mlp hidden width (mlp_ratio 4)

struct AfnoMlpArgs {
  int ntiles;             // tokens / 128
  const float* x;         // [tokens, 64] block input (the residual), channels-last
  float* out;             // [tokens, 64]; may alias x (each thread reads the rows it writes first)
  const uint8_t* W1img;   // fc1 * diag(gamma): [hi | lo] images, R = 256 rows, K = 64
  const uint8_t* W2img;   // fc2 / 2 (the epilogue hands on 2 gelu): [hi | lo] images, R = 64 rows, K = 256
  // the biases travel as kernel parameters: warp-uniform reads from the constant bank instead of 40 global loads
  // per thread and tile through the L1 data pipe (which is what bounds this kernel)
  float b1c[kAfnoHid];    // fc1 bias + fc1 * beta
  float b2c[kAfnoE];      // fc2 bias
  float eps;
  // optional fused head of the network (last block, 1x1 patches; fourcastnet.py:349-358): instead of storing
  // `out`, frame[b, co, pix] = last[b, co, pix] + sum_c out[token, c] * Wh[co, c]
  const float* Wh;        // [Cg, 64] head.weight, or null
  int Cg;                 // <= 4
  int HW;                 // tokens (= pixels) per sample
  float* frame;
  long long frame_bstride;
  const float* last;      // residual frame or null
  long long last_bstride;
};
int launch_afno_mlp(const AfnoMlpArgs& a, cudaStream_t st);

// ---- fused fp32 kernels of the AFNO2D token mixer (afno_fused.cu) -------------------------------
struct AfnoFwdArgs {
  // optional fused patch embedding of the network (first block, 1x1 patches; fourcastnet.py:285-289,537-542):
  // x = conv1x1(gathered NCHW input) + bias + pos_embed is computed here and stored to `xout`
  SrcGroup grp[4];
  int ngrp;               // 0: no embedding, read x
  int Cin;                // <= 16
  const float* We;        // [64, Cin] patch_embed.proj.weight
  const float* be;        // [64]
  const float* pos;       // [H * W, 64] or null
  float* xout;            // [B, H, W, 64]
  const float* x;         // [B, H, W, 64] block input
  float2* A;              // [B, Kc, H, 64] W-DFT of LayerNorm(x) on the retained columns
  const float2* Fw;       // [Kc, W]
  const float *gamma, *beta;
  float eps;
  int B, H, W, Kc;
};
struct AfnoMidArgs {
  const float2* A;        // [B, Kc, H, 64]
  float2* Z;              // [B, H, Kc, 64] after H-DFT, block MLP, inverse H-DFT
  const float2 *Fh, *Gh;  // [R, H], [H, R]
  const float *w1, *b1, *w2, *b2;   // filter weights, reference layout (fourcastnet.py:73-76), hidden_size_factor 1
  float lambda;
  int B, H, R, Kc, bs;
  int planarA;            // A is [B, Kc, H, (re | im), 64] (what afno_fwd_tc_kernel writes) instead of interleaved complex
};
struct AfnoInvArgs {
  const float2* Z;        // [B, H, Kc, 64]
  const float* x;         // block input (the residual; LayerNorm(x) is recomputed)
  float* out;             // [B, H, W, 64] = Re(Gw Z) + LayerNorm(x) + x
  const float2* Gw;       // [W, Kc]
  const float *gamma, *beta;
  float eps;
  int B, H, W, Kc;
};
int launch_afno_fwd(const AfnoFwdArgs& a, cudaStream_t st);

// ---- the W-DFTs of the token mixer on tcgen05 (tc_afno.cu) -------------------------------------------
// Tiles of 128 tokens = RT grid rows (W = 128 / RT in {32, 64, 128}); retained columns RT * Kc * 2 <= 128.
struct AfnoWTcArgs {
  int ntiles, H, W, RT, Kc;
  const float* x;         // [tokens, 64] block input
  const float *gamma, *beta;   // norm1
  float eps;
  // forward: A [B, Kc, H, (re | im), 64] = W-DFT of LayerNorm(x)
  const uint8_t* Fimg;    // [hi | lo] images of F''[(h2, kx, re | im), (h2', w)], R = 128 rows, K = 128
  float* A;
  uint8_t* Aimg;          // when set: instead of A, the split-bf16 operand image afno_mid_tc_kernel reads (AfnoMid2Args::Aimg)
  int ncols;              // B * Kc (part stride of Aimg)
  // forward, first block with 1x1 patches: x is not read but computed from the NCHW sources of the step (patch embedding +
  // position embedding, fourcastnet.py:285-289,537-542) and stored to `xout`
  SrcGroup grp[4];
  int ngrp;               // 0: read x
  int Cin;                // <= 16
  const float* We;        // [64, Cin] patch_embed.proj.weight
  const float* be;        // [64]
  const float* pos;       // [H * W, 64] or null
  float* xout;
  // inverse: out = Re(Gw Z) + LayerNorm(x) + x
  const uint8_t* Gimg;    // [hi | lo] images of G'[(h2, w), (h2', kx, re | im)], R = 128 rows, K = KZ
  int KZ;                 // RT * Kc * 2 rounded up to a multiple of 16
  const float2* Z;        // [B, H, Kc, 64]
  const uint8_t* Zimg;    // when set: instead of Z, ready MN-major operand images per tile (AfnoMid2Args::Zimg)
  float* out;             // [tokens, 64]
};
// ---- H-DFT, block MLP, inverse H-DFT on tcgen05 (tc_afno_mid.cu) ----------------------------------------
struct AfnoMid2Args {
  const uint8_t* Aimg;    // [hi | lo][column = b Kc + kx][8 channel octets][2H K rows k = 2h + (re|im)][8 ch] bf16: afno_fwd_tc_kernel
  uint8_t* Zimg;          // per 128-token tile [hi | lo][8 channel octets][KZ K rows (h2, kx, re|im)][8 ch]: afno_inv_tc_kernel's operand
  const uint8_t *Fimg, *Gimg;      // Fs / Gs [hi | lo] images, 64 rows x K = 64 (afno_mid_tc_pack_bases)
  const uint8_t *W1img, *W2img;    // per channel block [hi | lo] images of the (re|im)-stacked weights (afno_mid_tc_pack_layer)
  const float *b1, *b2;   // [128], column order (block, re|im, o)
  float lambda;           // softshrink threshold
  int ncols;              // B * Kc
  int H, R, Kc, bs, RT, KZ;
};
bool afno_mid_tc_covers(int H, int R, int bs);
int launch_afno_mid_tc(const AfnoMid2Args& a, cudaStream_t st);
void afno_mid_tc_pack_bases(const double* Fr, const double* Fi, const double* Gr, const double* Gi, int H, int R, uint8_t* Fimg, uint8_t* Gimg);
void afno_mid_tc_pack_layer(const float* w, const float* b, int nb, int bs, uint8_t* Wimg, float* bias);

int launch_afno_fwd_tc(const AfnoWTcArgs& a, cudaStream_t st);
int launch_afno_inv_tc(const AfnoWTcArgs& a, cudaStream_t st);
int afno_inv_tc_smem_bytes(int RT, int Kc, int KZ);
int launch_afno_mid(const AfnoMidArgs& a, cudaStream_t st);
int launch_afno_inv(const AfnoInvArgs& a, cudaStream_t st);

}  // namespace dlwp
