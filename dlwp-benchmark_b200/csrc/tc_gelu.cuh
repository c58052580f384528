// Exact-erf GELU on the packed fp32x2 pipe with one special-function op per element; shared by the
// tensor-core kernel sets (tc_fno.cu, tc_wide.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace dlwp {
namespace tc {

// Exact-erf GELU (F.gelu default) with ONE special-function op per element:
//   gelu(x) = x Phi(x) = x/2 + |x| (1/2 - Phi(-|x|)),   Phi(-a) = erfc(a / sqrt 2) / 2 = 2^(-h(a))
// h is smooth; a degree-5 polynomial (tools/fit_gelu.py: minimax on the GELU error over [0, 6],
// max |error| 5.4e-7 * max(1, |x|) in fp32 Horner arithmetic, checked over [-40, 40]) feeds
// ex2.approx.  Its leading coefficient is positive, so beyond the fitted range h keeps growing
// and 2^(-h) underflows to 0: no clamp is needed and the tails are exact to fp32.  Two values at
// once on the packed fp32x2 pipe (FFMA2 / FMUL2): 8 packed ops + 2 MUFU + 2 ALU ops per pair.
__device__ __forceinline__ float2 gelu2(float2 x) {
  const float2 a = make_float2(fabsf(x.x), fabsf(x.y));
  float2 p = __ffma2_rn(make_float2(-4.732980128e-04f, -4.732980128e-04f), a, make_float2(7.084488528e-03f, 7.084488528e-03f));
  p = __ffma2_rn(p, a, make_float2(-5.182723514e-02f, -5.182723514e-02f));
  p = __ffma2_rn(p, a, make_float2(-4.599925752e-01f, -4.599925752e-01f));
  p = __ffma2_rn(p, a, make_float2(-1.150787801e+00f, -1.150787801e+00f));
  p = __ffma2_rn(p, a, make_float2(-1.000037628e+00f, -1.000037628e+00f));      // p = -h(a)
  float2 e;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e.x) : "f"(p.x));
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e.y) : "f"(p.y));
  const float2 t = __ffma2_rn(e, make_float2(-1.0f, -1.0f), make_float2(0.5f, 0.5f));
  const float2 hx = __fmul2_rn(x, make_float2(0.5f, 0.5f));
  return __ffma2_rn(a, t, hx);
}

// The same on eight pairs in lock step (eight independent dependency chains in flight), returning
// TWICE the GELU: 2 gelu(x) = x + |x| (1 - 2 * 2^(-h)); the consumer GEMM's weights carry the 1/2
// (exact), which saves the x/2 multiply.
__device__ __forceinline__ void gelu2x_8(const uint32_t (&r)[16], float2 (&g)[8]) {
  float2 x[8], p[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) x[e] = make_float2(__uint_as_float(r[2 * e]), __uint_as_float(r[2 * e + 1]));
#pragma unroll
  for (int e = 0; e < 8; ++e)
    p[e] = __ffma2_rn(make_float2(-4.732980128e-04f, -4.732980128e-04f), make_float2(fabsf(x[e].x), fabsf(x[e].y)),
                      make_float2(7.084488528e-03f, 7.084488528e-03f));
  const float c[4] = {-5.182723514e-02f, -4.599925752e-01f, -1.150787801e+00f, -1.000037628e+00f};
#pragma unroll
  for (int k = 0; k < 4; ++k)
#pragma unroll
    for (int e = 0; e < 8; ++e) p[e] = __ffma2_rn(p[e], make_float2(fabsf(x[e].x), fabsf(x[e].y)), make_float2(c[k], c[k]));
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(p[e].x) : "f"(p[e].x));
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(p[e].y) : "f"(p[e].y));
  }
#pragma unroll
  for (int e = 0; e < 8; ++e) p[e] = __ffma2_rn(p[e], make_float2(-2.0f, -2.0f), make_float2(1.0f, 1.0f));
#pragma unroll
  for (int e = 0; e < 8; ++e) g[e] = __ffma2_rn(make_float2(fabsf(x[e].x), fabsf(x[e].y)), p[e], x[e]);
}

}  // namespace tc
}  // namespace dlwp
