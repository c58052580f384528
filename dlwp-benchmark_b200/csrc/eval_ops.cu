// Evaluation tail on the device (scripts/evaluate.py:281-296): undo the per-variable normalisation of a forecast and
// split it into one contiguous array per variable, which is what the netCDF writer (build_dataset, :320-416: one
// DataArray per variable, dims sample x time x lat x lon) needs -- one read and one write of the forecast instead of
// the reference's host-side numpy loop over variables followed by per-variable slicing.
#include "kernels.cuh"

namespace dlwp {
namespace {

// one block per (sample, lead time, variable) plane:  out[c][b][t][:] = x[b][t][c][:] * std[c] + mean[c]
template <int VEC>
__global__ void __launch_bounds__(256) denorm_split_kernel(const float* __restrict__ x, const float* __restrict__ mean,
                                                           const float* __restrict__ stdv, float* __restrict__ out, int B, int T, int C,
                                                           long long HW) {
  const long long plane = blockIdx.x;      // (b * T + t) * C + c
  const int c = (int)(plane % C);
  const long long bt = plane / C;
  const float m = mean ? mean[c] : 0.f, s = stdv ? stdv[c] : 1.f;
  const float* src = x + plane * HW;
  float* dst = out + ((long long)c * B * T + bt) * HW;
  if (VEC == 4) {
    const float4* s4 = reinterpret_cast<const float4*>(src);
    float4* d4 = reinterpret_cast<float4*>(dst);
    for (long long i = blockIdx.y * 256 + threadIdx.x; i < HW / 4; i += (long long)gridDim.y * 256) {
      const float4 v = s4[i];
      d4[i] = make_float4(fmaf(v.x, s, m), fmaf(v.y, s, m), fmaf(v.z, s, m), fmaf(v.w, s, m));
    }
  } else {
    for (long long i = blockIdx.y * 256 + threadIdx.x; i < HW; i += (long long)gridDim.y * 256) dst[i] = fmaf(src[i], s, m);
  }
}

}  // namespace
}  // namespace dlwp

using namespace dlwp;

extern "C" int dlwp_denormalize_split(const float* x, const float* mean, const float* stdv, float* out, int B, int T, int C, int H, int W,
                                      void* stream) {
  if (!x || !out) return fail(DLWP_ERR_INVALID, "dlwp_denormalize_split: null argument");
  DLWP_REQUIRE(B > 0 && T > 0 && C > 0 && H > 0 && W > 0, "dlwp_denormalize_split: empty problem");
  const long long HW = (long long)H * W, planes = (long long)B * T * C;
  DLWP_REQUIRE(planes < (1ll << 31), "dlwp_denormalize_split: too many planes");
  cudaStream_t st = (cudaStream_t)stream;
  const bool vec = HW % 4 == 0 && (reinterpret_cast<uintptr_t>(x) & 15) == 0 && (reinterpret_cast<uintptr_t>(out) & 15) == 0;
  long long per = vec ? HW / 4 : HW;
  int gy = (int)((per + 256 * 8 - 1) / (256 * 8));      // eight elements (vectors) per thread
  if (gy < 1) gy = 1;
  if (gy > 64) gy = 64;
  const dim3 grid((unsigned)planes, (unsigned)gy);
  if (vec) denorm_split_kernel<4><<<grid, 256, 0, st>>>(x, mean, stdv, out, B, T, C, HW);
  else denorm_split_kernel<1><<<grid, 256, 0, st>>>(x, mean, stdv, out, B, T, C, HW);
  DLWP_LAUNCHED("denorm_split_kernel");
  return DLWP_OK;
}
