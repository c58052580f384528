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
  const float* b1;        // [256] fc1 bias + fc1 * beta
  const uint8_t* W2img;   // fc2 / 2 (the epilogue hands on 2 gelu): [hi | lo] images, R = 64 rows, K = 256
  const float* b2;        // [64]
  float eps;
};
int launch_afno_mlp(const AfnoMlpArgs& a, cudaStream_t st);

}  // namespace dlwp
