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

namespace dlwp {
namespace {

constexpr size_t kStaticScratchBytes = DLWP_STATIC_SCRATCH_BYTES;
constexpr int kMaxRows = DLWP_STATIC_MAX_ROWS;

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

struct RowBins {
  int32_t in[kMaxRows], out[kMaxRows];
};

// The truncated-DFT bases of plan.cu:build_bases, generated on the device in fp64 (no host staging copy):
//   Fw [Kc, W] = s exp(-2 pi i k w / W)          Fh [R, H] = exp(-2 pi i rows_in[j] h / H)
//   Gh [H, R]  = exp(+2 pi i rows_out[j] h / H)  Gw [W, Kc] = t c_k exp(+2 pi i k w / W), c = 1 for the DC / Nyquist column, else 2
__device__ __forceinline__ float2 twiddle(long long k, long long n, long long N, double sign, double scale) {
  long long r = (k * n) % N;
  if (r < 0) r += N;
  double s, c;
  sincospi(2.0 * (double)r / (double)N, &s, &c);
  return make_float2((float)(scale * c), (float)(scale * sign * s));
}

__global__ void static_bases_kernel(float2* Fw, float2* Fh, float2* Gh, float2* Gw, int H, int W, int R, int Kc, double s,
                                    double t, RowBins rows) {
  const int n_fw = Kc * W, n_fh = R * H, total = 2 * n_fw + 2 * n_fh;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    if (i < n_fw) {
      const int k = i / W, w = i - k * W;
      Fw[i] = twiddle(k, w, W, -1.0, s);
    } else if (i < n_fw + n_fh) {
      const int j = (i - n_fw) / H, h = (i - n_fw) - j * H;
      Fh[i - n_fw] = twiddle(rows.in[j], h, H, -1.0, 1.0);
    } else if (i < n_fw + 2 * n_fh) {
      const int e = i - n_fw - n_fh, h = e / R, j = e - h * R;
      Gh[e] = twiddle(rows.out[j], h, H, +1.0, 1.0);
    } else {
      const int e = i - n_fw - 2 * n_fh, w = e / Kc, k = e - w * Kc;
      const bool self_conj = (k == 0) || (W % 2 == 0 && k == W / 2);
      Gw[e] = twiddle(k, w, W, +1.0, t * (self_conj ? 1.0 : 2.0));
    }
  }
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
  DLWP_REQUIRE(batch > 0 && Ci > 0 && Co > 0 && H > 0 && W > 0, "dlwp_spectral_conv2d_static: empty problem");
  DLWP_REQUIRE(n_rows >= 1 && n_rows <= kMaxRows && n_rows <= H, "dlwp_spectral_conv2d_static: n_rows=%d outside [1, min(%d, H=%d)]",
               n_rows, kMaxRows, H);
  DLWP_REQUIRE(n_cols >= 1 && n_cols <= W / 2 + 1, "dlwp_spectral_conv2d_static: n_cols=%d outside [1, W/2+1=%d]", n_cols, W / 2 + 1);
  RowBins rows{};
  for (int j = 0; j < n_rows; ++j) {
    DLWP_REQUIRE(rows_in[j] >= 0 && rows_in[j] < H && rows_out[j] >= 0 && rows_out[j] < H, "row bin out of range");
    rows.in[j] = rows_in[j];
    rows.out[j] = rows_out[j];
  }
  double s = 1.0, t = 1.0;
  const double hw = (double)H * (double)W;
  switch (fft_norm) {
    case DLWP_FFT_NORM_BACKWARD: t = 1.0 / hw; break;
    case DLWP_FFT_NORM_FORWARD: s = 1.0 / hw; break;
    case DLWP_FFT_NORM_ORTHO: s = t = 1.0 / std::sqrt(hw); break;
    default: return fail(DLWP_ERR_INVALID, "unknown fft_norm %d", fft_norm);
  }

  // carve the scratch array
  const size_t B = (size_t)batch, HW = (size_t)H * W, R = (size_t)n_rows, Kc = (size_t)n_cols;
  const size_t sizes[] = {
      Kc * W * sizeof(float2), R * H * sizeof(float2), H * R * sizeof(float2), W * Kc * sizeof(float2),   // Fw Fh Gh Gw
      B * HW * Ci * sizeof(float), B * HW * Co * sizeof(float),                                         // x, y channels-last
      B * H * Kc * Ci * sizeof(float2), R * Kc * B * Ci * sizeof(float2),                               // A, X
      R * Kc * B * Co * sizeof(float2), B * H * Kc * Co * sizeof(float2),                               // Y, Z
      (size_t)Ci * Co * R * Kc * sizeof(float2)};                                                       // mode-major weight
  size_t off[12] = {0};
  for (int i = 0; i < 11; ++i) off[i + 1] = off[i] + align_up(sizes[i]);
  if (off[11] > kStaticScratchBytes)
    return fail(DLWP_ERR_INVALID,
                "dlwp_spectral_conv2d_static: this problem needs %zu bytes of scratch, the static array holds %zu "
                "(DLWP_STATIC_SCRATCH_BYTES); use dlwp_spectral_conv2d or a plan for larger sizes",
                off[11], kStaticScratchBytes);
  unsigned char* base = nullptr;
  DLWP_TRY(scratch_base(&base));
  float2 *Fw = (float2*)(base + off[0]), *Fh = (float2*)(base + off[1]), *Gh = (float2*)(base + off[2]),
         *Gw = (float2*)(base + off[3]);
  float *xl = (float*)(base + off[4]), *yl = (float*)(base + off[5]);
  float2 *A = (float2*)(base + off[6]), *X = (float2*)(base + off[7]), *Y = (float2*)(base + off[8]),
         *Z = (float2*)(base + off[9]), *wm = (float2*)(base + off[10]);

  cudaStream_t st = (cudaStream_t)stream;
  {
    const int total = 2 * (int)(Kc * W) + 2 * (int)(R * H);
    static_bases_kernel<<<ceil_div(total, 256), 256, 0, st>>>(Fw, Fh, Gh, Gw, H, W, n_rows, n_cols, s, t, rows);
    DLWP_LAUNCHED("static_bases_kernel");
  }
  DLWP_TRY(launch_pack_modes(reinterpret_cast<const float2*>(weight), wm, Ci, Co, n_rows, n_cols, st));
  DLWP_TRY(launch_nchw_to_nhwc(x, xl, batch, Ci, H * W, st));
  // the transform chain of plan.cu:spectral_layer with separate input / output widths
  TdftArgs a{};
  a.in = xl; a.out = A; a.T = Fw; a.N = W; a.K = n_cols; a.C = Ci; a.nb0 = batch; a.nb1 = H;      // W axis: x [B,H,W,Ci] -> A [B,H,Kc,Ci]
  a.in_sn = Ci; a.in_sb0 = (long long)HW * Ci; a.in_sb1 = (long long)W * Ci;
  a.out_sk = Ci; a.out_sb0 = (long long)H * Kc * Ci; a.out_sb1 = (long long)Kc * Ci;
  a.in_complex = false; a.out_complex = true;
  DLWP_TRY(launch_tdft(a, st));
  a = TdftArgs{};
  a.in = A; a.out = X; a.T = Fh; a.N = H; a.K = n_rows; a.C = Ci; a.nb0 = batch; a.nb1 = n_cols;   // H axis: A -> X [R,Kc,B,Ci]
  a.in_sn = (long long)Kc * Ci; a.in_sb0 = (long long)H * Kc * Ci; a.in_sb1 = Ci;
  a.out_sk = (long long)Kc * B * Ci; a.out_sb0 = Ci; a.out_sb1 = (long long)B * Ci;
  a.in_complex = true; a.out_complex = true;
  DLWP_TRY(launch_tdft(a, st));
  DLWP_TRY(launch_mix_dense(X, wm, Y, n_rows * n_cols, batch, Ci, Co, st));                       // einsum("bixy,ioxy->boxy")
  a = TdftArgs{};
  a.in = Y; a.out = Z; a.T = Gh; a.N = n_rows; a.K = H; a.C = Co; a.nb0 = batch; a.nb1 = n_cols;   // inverse H: Y -> Z [B,H,Kc,Co]
  a.in_sn = (long long)Kc * B * Co; a.in_sb0 = Co; a.in_sb1 = (long long)B * Co;
  a.out_sk = (long long)Kc * Co; a.out_sb0 = (long long)H * Kc * Co; a.out_sb1 = Co;
  a.in_complex = true; a.out_complex = true;
  DLWP_TRY(launch_tdft(a, st));
  a = TdftArgs{};
  a.in = Z; a.out = yl; a.T = Gw; a.N = n_cols; a.K = W; a.C = Co; a.nb0 = batch; a.nb1 = H;       // inverse W (C2R): Z -> y [B,H,W,Co]
  a.in_sn = Co; a.in_sb0 = (long long)H * Kc * Co; a.in_sb1 = (long long)Kc * Co;
  a.out_sk = Co; a.out_sb0 = (long long)HW * Co; a.out_sb1 = (long long)W * Co;
  a.in_complex = true; a.out_complex = false;
  a.bias = bias;
  DLWP_TRY(launch_tdft(a, st));
  DLWP_TRY(launch_nhwc_to_nchw(yl, y, batch, Co, H * W, st));
  return DLWP_OK;
}
