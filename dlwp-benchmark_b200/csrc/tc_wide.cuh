// Tensor-core (tcgen05) kernel set for the FNO step at hidden width 128 in bf16
// (precision mode DLWP_PREC_BF16; BASELINE config 5: 721 x 1440 grid, 128 channels, 64 x 33 modes).
//
// Every grid row of W pixels is cut into TW = ceil(W / 128) tiles of 128 pixel slots (slots past
// W are padding: they hold finite values that no result depends on).  Activations live in HBM
// as bf16 "tile images": [128 slots] x [K channels] in the chunk-major UMMA operand layout of
// tc_common.cuh, so a consumer brings a tile in with one bulk async copy and hands it to
// tcgen05.mma untouched -- as the K-major A operand of the pointwise GEMMs, and as the MN-major
// B operand of the forward W-DFT (which contracts over pixels).
//
// One model evaluation (reference: neuralop FNO.forward behind fno.py:103):
//   wide_gather_kernel   x_t from the NCHW sources (fno.py:49-62) -> [128 x 32] input images
//                        (hi parts, a ones column for the bias, lo parts)
//   tcw_lift             x0 = W2 gelu(W1 x + b1) + b2, hidden on chip  -> [128 x 128] activation images
//   per layer:
//     tcw FWD            A[h, kx, c] = sum_w F[kx, w] x[h, w, c]   one item per grid row -> bf16 spectrum images
//                        per (sample, kx, block of 64 grid rows): rows (h, re | im), columns = channels
//     tcw HFWD           X[r, kx, c] = sum_h Fh[r, h] A[h, kx, c]  complex product as one real GEMM with
//                        (re | im)-stacked rows, contracting over the image rows -> fp32 [R, Kc, B, 128]
//     mix_dense          per-mode channel mixing (fp32 FMA, kernels_fp32.cu) -> bf16 images per (sample, kx):
//                        rows (r, re | im), columns = channels
//     tcw HINV           Z[h, kx, c] = sum_r Gh[h, r] Y[r, kx, c]  -> per grid row [KZ rows (kx, re | im)] x
//                        [128 channels] bf16 (the MN-major B operand of the inverse W-DFT)
//     tcw LAYER          x' = act(G Z + Wskip x + bias)            -> [128 x 128] images
//   tcw_proj             out = P2 gelu(P1 x + b1) + b2 + last frame, hidden on chip -> NCHW fp32
// GELU epilogues store 2 * gelu (tc_gelu.cuh); the consumers' weights carry the 1/2.
#pragma once
#include "kernels.cuh"

namespace dlwp {

constexpr int kWideC = 128;     // hidden channels
constexpr int kWideHid = 256;   // lifting / projection width

enum TcwMode { TCW_FWD = 0, TCW_HFWD = 1, TCW_HINV = 2, TCW_LAYER = 3 };

struct TcwArgs {
  int nitems;             // tiles (B * H * TW); grid rows (B * H) for TCW_FWD; B * Kc (* HB) for TCW_HFWD (TCW_HINV)
  int TW, H, W;           // tiles per grid row, grid rows per sample, grid width
  int B, R, HB;           // samples, retained rows (2 R <= 128), blocks of 64 grid rows per sample
  int Kc, KZ;             // retained columns; K extent of the inverse W-DFT operands (2 Kc padded to 16)
  const uint8_t* a_img;   // streamed tile images: input [128 x 32] / hidden [128 x 256] / activation [128 x 128]
  const uint8_t* b_res;   // resident weight image [N rows x K]
  const uint8_t* bias_img;// [N rows x 16]: columns 0 / 1 = hi / lo of the bias (times a ones operand), or null
  const uint8_t* w2_img;  // fused MLP kernels: second weight image [N2 rows x 256] and its bias image
  const uint8_t* bias2_img;
  const uint8_t* g_img;   // TCW_LAYER: TW inverse W-DFT tiles [128 x KZ]; DFT modes: the [128 x 128] DFT tiles
  const uint8_t* z_img;   // TCW_LAYER: per grid row [KZ x 128 channels] (MN-major)
  uint8_t* out_img;       // tile images out (TCW_FWD: spectrum images; TCW_HINV: Z images)
  float2* spec;           // TCW_HFWD: X [R, Kc, B, 128] complex
  float scale;            // TCW_FWD / TCW_HINV: factor on the result (1/2 when the input image holds 2 * gelu)
  int gelu;               // TCW_LAYER: apply GELU
  float* out;             // projection: NCHW frame(s) out[b * out_bstride + (c, y, x)]
  long long out_bstride;
  const float* last;      // residual frame or null
  long long last_bstride;
  int Cg;
};
int launch_tcw(int mode, const TcwArgs& a, cudaStream_t st);
// fused pointwise MLP with the 256-wide hidden on chip: kind 0 = lifting (input images -> activation
// images), kind 1 = projection (activation images -> NCHW frame + residual)
int launch_tcw_mlp(int kind, const TcwArgs& a, cudaStream_t st);

struct WideGatherArgs {
  SrcGroup grp[4];
  int ngrp, Cin;          // Cin <= 14
  int ntiles, TW, H, W;
  uint8_t* in_img;        // [ntiles] x [128 x 32]
};
int launch_wide_gather(const WideGatherArgs& a, cudaStream_t st);

// few-sample mode mixing on tcgen05 (tc_wide_mix.cu): weights as per-mode operand images [modes][64 KB] (launch_mix_few_pack from
// the mode-major fp32 weights [modes][128][128] complex), spectrum X [modes, B, 128] complex -> the bf16 tile images `ys`
bool mix_few_tc_covers(int modes, int B, int C, int Kc);
int launch_mix_few_pack(const float2* wm, uint8_t* Wimg, int modes, cudaStream_t st);
int launch_mix_few_tc(const float2* X, const uint8_t* Wimg, uint8_t* ys, int modes, int B, int Kc, cudaStream_t st);

// host: chunk-major bf16 image of a row-major matrix `m` [rows, cols] padded to R x K with zeros
void tcw_pack_image(const float* m, int rows, int cols, int ld, int R, int K, uint8_t* out);
void tcw_pack_image_f64(const double* m, int rows, int cols, int ld, int R, int K, uint8_t* out);

}  // namespace dlwp
