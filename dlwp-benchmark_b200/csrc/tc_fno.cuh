// Tensor-core (tcgen05) kernel set for the FNO / TFNO step at hidden width 32.
//
// Activations between kernels live in HBM as "tile images": 128 consecutive pixels x 32
// channels, split into bf16 hi and lo parts (v = hi + lo), each part stored in exactly the
// chunk-major shared-memory layout the UMMA descriptors read (tc_common.cuh).  One tile is
// 16 KB (8 KB hi image, then 8 KB lo image) = the bytes of the same tile in fp32; a consumer
// brings it in with one bulk async copy and feeds it to tcgen05.mma without touching it.
// Every fp32 product a*b is evaluated as a_hi*b_hi + a_lo*b_hi + a_hi*b_lo on the bf16 tensor
// cores with fp32 accumulation in tensor memory (error ~2^-17 relative per product).
//
// Per model evaluation (reference: neuralop FNO.forward as called from fno.py:103):
//   fno_lift_kernel    gather x_t from the NCHW sources (fno.py:49-62) -> lifting MLP
//                      (1x1 conv, GELU, 1x1 conv; the hidden layer stays in tensor memory) -> tile image
//                      + forward truncated W-DFT of the result (a third MMA on the same tile)
//   fno_mid_tc_kernel  H-DFT, per-mode channel mixing, inverse H-DFT on the retained columns as three
//                      chained GEMMs (tc_mid2.cu; fno_mid_kernel is the fp32 FMA form) -> split-bf16 operand image of Z
//   fno_layer_kernel   inverse truncated W-DFT GEMM K-concatenated with the 1x1 skip conv,
//                      + bias, GELU -> tile image + forward W-DFT for the next layer
//   fno_proj_kernel    projection MLP + residual on the last context frame -> NCHW output
#pragma once
#include "kernels.cuh"

namespace dlwp {

constexpr int kTcTile = 128;      // pixels per tile (UMMA M)
constexpr int kTcC = 32;          // hidden channels of the tensor-core path
constexpr int kTcImgBytes = kTcTile * kTcC * 2;   // one bf16 image (hi or lo) of a tile: 8 KB

// geometry shared by the three kernels
struct TcGeom {
  int ntiles;          // B * H * W / 128
  int HW;              // pixels per sample
  int H;               // grid rows per sample
  int W;               // grid width (128 % W == 0)
  int RT;              // grid rows per tile = 128 / W
  int KXP;             // padded real columns per grid row = 32 / RT  (2 * Kc <= KXP)
  int Kc;              // retained half-spectrum columns
  int KZ;              // K extent of the inverse W-DFT operands: 8 per retained column, padded to 16
  int abf;             // the forward W-DFT leaves as split-bf16 operand tiles for fno_mid_tc_kernel (tc_mid2.cu) instead of planar fp32
  int Bn;              // samples of this launch (ntiles * 128 / HW)
  int TPS;             // tiles per sample = HW / 128 = H / RT (a tile lies within one sample: 32-bit index arithmetic per tile)
  int l2hint;          // L2 eviction priorities on the bulk copies of the activation images (tc_common.cuh); DLWP_L2_HINTS=0 clears it
  int rev;             // fno_layer_kernel: the tile list is walked from the last tile down (see tc_plan.cu)
};

struct TcLiftArgs {
  TcGeom g;
  SrcGroup grp[4];     // NCHW sources of x_t
  int ngrp;
  int Cin;             // <= 15 (column Cin of the first GEMM carries the bias)
  int hid;             // lifting width: 128 or 256
  int workers4;        // the four-worker kernel (fno_lift4_kernel) when every worker has a tile
  const uint8_t* W1img;   // [hi | lo] images, R = hid rows, K = 16
  const uint8_t* W2img;   // [hi | lo] images, R = 32 rows, K = hid
  const float* b2;        // [32]
  const uint8_t* Fimg;    // forward W-DFT operand [hi | lo] images, R = 32, K = 128
  uint8_t* y_img;         // tile images out
  float* Aout;            // planar spectrum [B][Kc][H][re|im][32]: forward W-DFT of the output
};

struct TcLayerArgs {
  TcGeom g;
  const uint8_t* x_img;
  uint8_t* y_img;
  const uint8_t* Zimg;    // per tile: split-bf16 MN-major operand image of the spectrum after the inverse H-DFT (layout: tc_mid.cu)
  float* Aout;            // forward W-DFT of the output (planar, as above), or null (last layer)
  const uint8_t* Gimg;    // inverse W-DFT operand [hi | lo] images, R = 128, K = KZ
  const uint8_t* Fimg;
  const uint8_t* Wsimg;   // skip conv [hi | lo] images, R = 32, K = 32
  const float* bias;      // [32] spectral-conv bias of this layer
  int gelu;
  int ny;                 // output image buffers per worker (set by the launcher)
};

struct TcProjArgs {
  TcGeom g;
  const uint8_t* x_img;
  int hid;                // projection width: 128 or 256
  int Cg;                 // output channels <= 16
  const uint8_t* W1img;   // [hi | lo] images, R = hid rows, K = 32
  const uint8_t* W1ximg;  // one bf16 image, R = hid rows, K = 16: columns 0 / 1 = hi / lo of the first bias
  const uint8_t* W2img;   // [hi | lo] images, R = 16 rows, K = hid
  const float* b2;        // [Cg]
  float* out;             // NCHW frame(s): out[b * out_bstride + (c, y, x)]
  long long out_bstride;
  const float* last;      // residual frame or null
  long long last_bstride;
  int last_bulk;          // set by the launcher: the residual rows are 16-byte aligned (bulk copies)
};

// H-DFT, per-mode channel mixing and inverse H-DFT on the retained columns (tc_mid.cu)
constexpr int kTcMidSamples = 4;   // samples one CTA pass works on
struct TcMidArgs {
  const float* A;         // planar spectrum [B][Kc][H][re|im][32]: forward W-DFT of the layer input
  const float2* Wm;       // [R, Kc, 32, 32] mode-major spectral weights of the layer
  const float2* Fh;       // [R, H] forward H-DFT rows
  const float2* Gh;       // [H, R] inverse H-DFT columns
  uint8_t* Zimg;          // [B*H/RT] operand images of KZ * 128 bytes
  int B, H, R, Kc, RT, KZ;
};
size_t tc_mid_smem_bytes(int H, int R);
int launch_tc_mid(const TcMidArgs& a, cudaStream_t st);

// The same three products on tcgen05 (tc_mid2.cu): the forward W-DFT of the lift / layer kernels arrives as
// split-bf16 operand tiles, one per (retained column kx, part hi | lo, sample b):
//   A[kx][part][b] : [4 channel groups g][2H rows k = 2 h + (re|im)][8 channels] bf16 = H * 128 bytes,
// i.e. the MN-major B operand of   X[(re|im, r), (b, i)] = Fs[(re|im, r), k] * A[k, (b, i)]   for NB samples at once.
struct TcMid2Args {
  const uint8_t* A;       // [Kc][2][B][H * 128 bytes]
  const uint8_t* Wimg;    // this layer's mode weights as operand images [Kc][groups of 4 modes][hi 16 KB | lo 16 KB]: rows (mode, o), K = (Wr^T | Wi^T)
  const uint8_t* Fimg;    // forward H-DFT, (re|im)-stacked: [hi | lo] images of 64 rows (Re X, Im X, -Im X, Re X) x 2H
  const uint8_t* Gimg;    // inverse H-DFT, (re|im)-stacked: [hi | lo] images of max(2H, 64) rows x 32
  uint8_t* Zimg;          // as TcMidArgs
  int B, H, R, Kc, RT, KZ;
  int l2hint;             // as TcGeom
  int rev;                // the sample groups are taken from the last one down
};
// samples per work item (0 = this geometry is not covered: the fp32 FMA kernel runs instead)
int tc_mid2_samples(int H, int R);
int launch_tc_mid2(const TcMid2Args& a, cudaStream_t st);
// mode-major fp32 weights [R][Kc][32][32] complex -> operand images (device kernel)
int launch_tc_mid2_pack(const float2* wm, uint8_t* Wimg, int R, int Kc, cudaStream_t st);
// host: the stacked DFT operand images; Fh [R][H], Gh [H][R] complex as in plan.cu:build_bases (fp64)
void tc_mid2_pack_bases(const double* Fh_re, const double* Fh_im, const double* Gh_re, const double* Gh_im, int H, int R,
                        uint8_t* Fimg, uint8_t* Gimg);

// last spectral layer + projection MLP in one kernel (tc_fno_lp.cu): the layer's output tile goes accumulator -> bias ->
// split-bf16 pairs written back in place -> A operand of the projection's first GEMM, never to HBM
struct TcLayerProjArgs {
  TcLayerArgs layer;      // y_img, Aout, Fimg unused; gelu as given (0 for neuralop's last block)
  TcProjArgs proj;        // x_img unused
};
int launch_tc_layerproj(const TcLayerProjArgs& a, cudaStream_t st);
bool tc_layerproj_fits(int KZ);
int launch_tc_lift(const TcLiftArgs& a, cudaStream_t st);
int launch_tc_layer(const TcLayerArgs& a, cudaStream_t st);
int launch_tc_proj(const TcProjArgs& a, cudaStream_t st);

// ---- host-side packing (plan_commit) ----------------------------------------------------
// Chunk-major [hi | lo] bf16 images of a row-major fp32 matrix `m` [rows, cols] padded to
// R x K (zeros); returns 2 * R * K * 2 bytes.  `lo` is the bf16 of (v - hi).
void tc_pack_images(const float* m, int rows, int cols, int ld, int R, int K, uint8_t* out);
// same from fp64 (DFT bases)
void tc_pack_images_f64(const double* m, int rows, int cols, int ld, int R, int K, uint8_t* out);
// one bf16 image [R, 16] with columns 0 / 1 = hi / lo parts of bias[r] (multiplied by a ones operand)
void tc_pack_bias_image(const float* bias, int n, int R, uint8_t* out);
// a single bf16 image whose rows [0, R/2) are hi parts and [R/2, R) lo parts of the R/2 x K matrix
void tc_pack_stacked_f64(const double* m, int rows, int cols, int ld, int R, int K, uint8_t* out);

}  // namespace dlwp
