// dlwp_spectral_conv2d_static: the stand-alone dense Fourier layer of dlwp_spectral_conv2d (unet.py:19-69) for callers
// that cannot allocate -- code under CUDA-graph capture, tight serving loops, harnesses that forbid cudaMalloc.
// No counterpart in the reference.
//
// Every intermediate of the layer (channels-last copies of x and y, the four spectra, the mode-major weight copy and
// the truncated-DFT bases) lives in ONE file-scope device array of kStaticScratchBytes, which every GPU that loads the
// library holds until the process exits.  In host code the array's name is only a stand-in for the device symbol: its
// device address is fetched once per device with cudaGetSymbolAddress (with that device current) and it is that
// pointer the kernels get.  The entry calls no cudaMalloc / cudaFree and no allocating library routine, and it does
// not synchronise: it only enqueues kernels on `stream`.
//
// All calls share the array and it keeps the last call's contents.  Nothing in it has to start at zero: every stage
// below writes the whole region the next stage reads (bases, both layout copies, all four spectra, the packed
// weight), so a call never sees a previous call's data.  Calls that overlap on the GPU -- two streams, two host
// threads, or a captured graph replayed while another call is in flight -- RACE on the array; serialise them.
#include <cmath>
#include <mutex>

#include "kernels.cuh"
#include "spectral_chain.cuh"

namespace dlwp {
namespace {

constexpr size_t kStaticScratchBytes = DLWP_STATIC_SCRATCH_BYTES;

static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

constexpr int kMaxDevices = 64;
std::mutex g_addr_mutex;
void* g_scratch_addr[kMaxDevices] = {};      // device address of g_static_scratch, per device ordinal

// device address of the scratch array on the CURRENT device (looked up once per device)
int scratch_base(unsigned char** out) {
  int dev = 0;
  DLWP_CUDA(cudaGetDevice(&dev));
  DLWP_REQUIRE(dev >= 0 && dev < kMaxDevices, "dlwp_spectral_conv2d_static: device ordinal %d beyond %d", dev, kMaxDevices);
  std::lock_guard<std::mutex> lock(g_addr_mutex);
  if (!g_scratch_addr[dev]) {
    void* p = nullptr;
    DLWP_CUDA(cudaGetSymbolAddress(&p, g_static_scratch));
    DLWP_REQUIRE(p != nullptr, "dlwp_spectral_conv2d_static: cudaGetSymbolAddress returned a null address");
    g_scratch_addr[dev] = p;
  }
  *out = static_cast<unsigned char*>(g_scratch_addr[dev]);
  return DLWP_OK;
}

inline size_t align_up(size_t v) { return (v + 255) & ~(size_t)255; }

}  // namespace
}  // namespace dlwp

using namespace dlwp;

extern "C" int dlwp_spectral_conv2d_static(const float* x, const float* weight, const float* bias, float* y, int batch,
                                           int Ci, int Co, int H, int W, int n_rows, int n_cols, const int32_t* rows_in,
                                           const int32_t* rows_out, int fft_norm, void* stream) {
  if (!x || !weight || !y || !rows_in || !rows_out)
    return fail(DLWP_ERR_INVALID, "dlwp_spectral_conv2d_static: null argument");
  SpecChain c;
  c.B = batch; c.Ci = Ci; c.Co = Co; c.H = H; c.W = W; c.R = n_rows; c.Kc = n_cols;
  DLWP_TRY(c.check(rows_in, rows_out, fft_norm, "dlwp_spectral_conv2d_static"));
  // carve the scratch array: chain buffers (bases, spectra, packed weight), then the two channels-last copies
  const size_t P = (size_t)batch * H * W, chain = c.carve(nullptr);
  const size_t need = chain + align_up(P * Ci * sizeof(float)) + align_up(P * Co * sizeof(float));
  if (need > kStaticScratchBytes)
    return fail(DLWP_ERR_INVALID,
                "dlwp_spectral_conv2d_static: this problem needs %zu bytes of scratch, the static array holds %zu "
                "(DLWP_STATIC_SCRATCH_BYTES); use dlwp_spectral_conv2d or a plan for larger sizes",
                need, kStaticScratchBytes);
  unsigned char* base = nullptr;
  DLWP_TRY(scratch_base(&base));
  c.carve(base);
  float *xl = (float*)(base + chain), *yl = (float*)(base + chain + align_up(P * Ci * sizeof(float)));
  cudaStream_t st = (cudaStream_t)stream;
  DLWP_TRY(c.bases(st));
  DLWP_TRY(launch_pack_modes(reinterpret_cast<const float2*>(weight), c.wm, Ci, Co, n_rows, n_cols, st));
  DLWP_TRY(launch_nchw_to_nhwc(x, xl, batch, Ci, H * W, st));
  DLWP_TRY(c.run(xl, yl, nullptr, nullptr, bias, 0, st));
  DLWP_TRY(launch_nhwc_to_nchw(yl, y, batch, Co, H * W, st));
  return DLWP_OK;
}
