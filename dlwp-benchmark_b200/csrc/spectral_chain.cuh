// One dense Fourier layer on channels-last fp32 activations with caller-provided scratch: shared by the stand-alone
// entries dlwp_spectral_conv2d / dlwp_fourier_residual_block (spectral_ops.cu) and dlwp_spectral_conv2d_static
// (static_scratch.cu).  Same mathematics as plan.cu:spectral_layer with separate input / output widths.
#pragma once
#include "kernels.cuh"

namespace dlwp {

constexpr int kSpecMaxRows = DLWP_STATIC_MAX_ROWS;      // retained rows the by-value row table holds

struct SpecRows {
  int32_t in[kSpecMaxRows], out[kSpecMaxRows];
};

struct SpecChain {
  int B = 0, Ci = 0, Co = 0, H = 0, W = 0, R = 0, Kc = 0;
  double s = 1.0, t = 1.0;      // forward / inverse scale of the fft_norm
  SpecRows rows{};
  float2 *Fw = nullptr, *Fh = nullptr, *Gh = nullptr, *Gw = nullptr;      // truncated-DFT bases
  float2 *A = nullptr, *X = nullptr, *Y = nullptr, *Z = nullptr;          // the four spectra
  float2* wm = nullptr;                                                  // mode-major weights [R, Kc, Ci, Co]

  // validates the geometry, copies the row bins, sets the scales
  int check(const int32_t* rows_in, const int32_t* rows_out, int fft_norm, const char* who);
  // assigns the nine buffers inside `base` (null: only measures); returns the bytes they take
  size_t carve(unsigned char* base);
  // fills the bases (device kernel, fp64 twiddles)
  int bases(cudaStream_t st);
  // yl = act( irfft2( mix( rfft2(xl)[bins] ) ) + add1 + add2 + bias ) on channels-last tensors
  int run(const float* xl, float* yl, const float* add1, const float* add2, const float* bias, int act, cudaStream_t st);
};

// adjoint bases of a chain whose bases are filled (train_ops.cu): GwH [Kc, W] = conj(Gw)^T, GhH [R, H] = conj(Gh)^T,
// FhH [H, R] = conj(Fh)^T, FwH [W, Kc] = conj(Fw)^T
int launch_adjoint_bases(const SpecChain& c, float2* GwH, float2* GhH, float2* FhH, float2* FwH, cudaStream_t st);

}  // namespace dlwp
