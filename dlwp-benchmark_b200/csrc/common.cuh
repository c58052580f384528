// Shared helpers for the dlwp_b200 native library: error plumbing, launch counting, small
// device math.  Everything here is internal; the public surface is include/dlwp_b200.h.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <string>

#include "../../include/dlwp_b200.h"

namespace dlwp {

// ---- error plumbing ----------------------------------------------------------------
extern thread_local std::string g_last_error;
extern std::atomic<int64_t> g_launches;
// per-kernel CUDA-event profiler (dlwp_profile_begin / _end); off on the hot path
extern bool g_prof_on;
void prof_mark(const char* name, cudaStream_t st);

inline int fail(int code, const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  g_last_error = buf;
  return code;
}

#define DLWP_CUDA(expr)                                                                        \
  do {                                                                                         \
    cudaError_t _e = (expr);                                                                   \
    if (_e != cudaSuccess)                                                                     \
      return ::dlwp::fail(DLWP_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), \
                          __FILE__, __LINE__);                                                 \
  } while (0)

#define DLWP_TRY(expr)            \
  do {                            \
    int _s = (expr);              \
    if (_s != DLWP_OK) return _s; \
  } while (0)

#define DLWP_REQUIRE(cond, ...)                                   \
  do {                                                            \
    if (!(cond)) return ::dlwp::fail(DLWP_ERR_INVALID, __VA_ARGS__); \
  } while (0)

// call after every kernel launch: counts it and surfaces launch-configuration errors
#define DLWP_LAUNCHED(name)                                                                   \
  do {                                                                                        \
    ::dlwp::g_launches.fetch_add(1, std::memory_order_relaxed);                               \
    if (::dlwp::g_prof_on) ::dlwp::prof_mark(name, st);                                       \
    cudaError_t _e = cudaPeekAtLastError();                                                   \
    if (_e != cudaSuccess)                                                                    \
      return ::dlwp::fail(DLWP_ERR_CUDA, "launch of %s failed: %s", name, cudaGetErrorString(_e)); \
  } while (0)

// Per-device launch state (plan.cu).  Function attributes and SM counts belong to a device, and the launchers may be
// reached from several host threads: both are cached per (device, kernel) behind a mutex instead of in function-local
// statics.  `current_sms`: SM count of the current device.  `ensure_dynamic_smem`: raises the kernel's
// MaxDynamicSharedMemorySize on the current device to at least `bytes` (once per device and size).
int current_sms(int* sms);
int ensure_dynamic_smem(const void* kernel, int bytes);

inline int ceil_div(int a, int b) { return (a + b - 1) / b; }
inline int64_t ceil_div64(int64_t a, int64_t b) { return (a + b - 1) / b; }

// ---- device math -------------------------------------------------------------------
__device__ __forceinline__ float gelu_erf(float x) {
  // F.gelu default / nn.GELU(): 0.5 * x * (1 + erf(x / sqrt(2)))
  return 0.5f * x * (1.0f + erff(x * 0.70710678118654752440f));
}

__device__ __forceinline__ float softshrink(float x, float lam) {
  return x > lam ? x - lam : (x < -lam ? x + lam : 0.0f);
}

}  // namespace dlwp
