/*
 * dlwp_b200.h -- C ABI of the B200-native rollout forward pass for dlwp-benchmark's
 * Fourier-neural-operator backbones (FNO2D / TFNO2D / FourCastNet-AFNO).
 *
 * The reference (AnneLouisedb/dlwp-benchmark) is pure Python and has no FFI of its own:
 * the seam it offers is the nn.Module contract of
 *   src/dlwpbench/models/fno/fno.py:12-146            (FNO2DModule, TFNO2DModule)
 *   src/dlwpbench/models/fourcastnet/fourcastnet.py:214-361 (AFNONet = "FourCastNet")
 * Each entry point below states which reference lines it replaces.  The Python front end
 * (dlwp_benchmark_b200.models) binds these with ctypes; INTEGRATION.md shows the stub.
 *
 * Conventions
 *   - every function returns 0 on success and a negative dlwp_status otherwise; nothing
 *     throws, aborts, exits or prints.  dlwp_last_error() gives the message (thread-local).
 *   - tensors are raw pointers + sizes; fp32 unless stated; no torch types.
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream).
 *   - the caller owns every buffer it passes; the plan owns packed weights, DFT bases and
 *     workspace, all allocated in plan_create / commit_weights.  dlwp_step, dlwp_rollout and
 *     dlwp_metrics_accumulate never allocate, never synchronise and only use `stream`, so
 *     they can be captured in a CUDA graph.  A plan is not thread-safe.
 *   - there is no CPU fallback: without a usable sm_100 device plan_create fails.
 */
#ifndef DLWP_B200_H
#define DLWP_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DLWP_ABI_VERSION 3

typedef enum {
  DLWP_OK = 0,
  DLWP_ERR_INVALID = -1,      /* bad argument / unsupported shape      */
  DLWP_ERR_CUDA = -2,         /* a CUDA runtime call failed            */
  DLWP_ERR_STATE = -3,        /* call order (weights not committed...) */
  DLWP_ERR_MISSING = -4,      /* a required weight tensor is missing   */
  DLWP_ERR_NO_DEVICE = -5     /* no CUDA device / not sm_100           */
} dlwp_status;

typedef enum {
  DLWP_MODEL_FNO = 0,         /* neuralop FNO (dense) and TFNO (Tucker, densified at load) */
  DLWP_MODEL_AFNO = 1         /* FourCastNet AFNONet (token mixer: see dlwp_afno_filter below)       */
} dlwp_model_kind;

typedef enum {
  DLWP_AFNO_FILTER_AFNO2D = 0, /* block-diagonal spectral MLP (fourcastnet.py:60-127)       */
  DLWP_AFNO_FILTER_FNO2D = 1   /* one-layer neuralop FNO as token mixer (fourcastnet.py:130-153):
                                  hidden = lifting = projection = embed_dim; the spectral fields
                                  (fft_norm, rows, n_cols) then describe that FNO's modes    */
} dlwp_afno_filter;

typedef enum {
  DLWP_PREC_FP32 = 0,         /* fp32 storage, fp32 FMA arithmetic (parity mode, <=1e-4)   */
  DLWP_PREC_FP32_TC = 1,      /* fp32 storage, split-bf16 tensor-core GEMMs, fp32 accumulate */
  DLWP_PREC_BF16 = 2          /* bf16 operands on tensor cores, fp32 accumulate            */
} dlwp_precision;

typedef enum {
  DLWP_FFT_NORM_BACKWARD = 0, /* forward unscaled, inverse 1/(H*W)   (unet.py:49,68)       */
  DLWP_FFT_NORM_FORWARD = 1,  /* forward 1/(H*W), inverse unscaled   (neuralop fft_norm)   */
  DLWP_FFT_NORM_ORTHO = 2     /* 1/sqrt(H*W) both ways               (fourcastnet.py:85,124) */
} dlwp_fft_norm;

/* Model description = the reference constructor kwargs after the front end resolved them.
 * FNO fields mirror fno.py:18-47 (and the neuralop defaults they imply); AFNO fields mirror
 * fourcastnet.py:215-268.  Unused fields of the other kind are ignored. */
typedef struct {
  int32_t kind;                 /* dlwp_model_kind */
  int32_t precision;            /* dlwp_precision  */
  int32_t max_batch;            /* workspace is sized for this many samples per call */
  int32_t height, width;        /* grid H, W */
  int32_t constant_channels;    /* Cc */
  int32_t prescribed_channels;  /* Cp */
  int32_t prognostic_channels;  /* Cg */
  int32_t context_size;         /* ctx; in_channels = Cc + (Cp + Cg) * ctx  (fno.py:36) */

  /* spectral transform: which rfft2 bins the layer reads / writes */
  int32_t fft_norm;             /* dlwp_fft_norm */
  int32_t n_rows;               /* number of retained row bins  */
  int32_t n_cols;               /* retained column bins 0..n_cols-1 of the W/2+1 half spectrum */
  const int32_t* rows_in;       /* [n_rows] unshifted FFT row bin read for weight row j  */
  const int32_t* rows_out;      /* [n_rows] unshifted FFT row bin written by weight row j */

  /* FNO / TFNO */
  int32_t hidden_channels, lifting_channels, projection_channels, n_layers;

  /* AFNO */
  int32_t embed_dim, depth, num_blocks, mlp_hidden, patch_h, patch_w, use_pos_embed;
  int32_t hidden_size_factor;   /* AFNO2D hidden_size_factor (1 in the reference) */
  float sparsity_threshold;     /* softshrink lambda */
  float layernorm_eps;          /* 1e-6 (fourcastnet.py:245) */
  int32_t afno_filter;          /* dlwp_afno_filter: the filter key of fourcastnet.yaml, eval'd at fourcastnet.py:259 */
} dlwp_model_desc;

typedef struct dlwp_plan dlwp_plan;

int dlwp_abi_version(void);
const char* dlwp_last_error(void);

/* Replaces module construction (fno.py:38-47,136-146; fourcastnet.py:247-268): validates the
 * description, builds the truncated-DFT bases in fp64, allocates workspace on `device`. */
int dlwp_plan_create(const dlwp_model_desc* desc, int device, dlwp_plan** out);
int dlwp_plan_destroy(dlwp_plan* plan);
int dlwp_plan_workspace_bytes(const dlwp_plan* plan, size_t* bytes);
/* *out = 1 when the plan runs the tcgen05 tensor-core kernel set (DLWP_PREC_FP32_TC on a shape
 * it covers), 0 when it runs the fp32 FMA kernel set. */
int dlwp_plan_uses_tensor_cores(const dlwp_plan* plan, int* out);
/* Diagnostics (no reference counterpart): every plan-owned buffer (workspace, operand images, packed weights) is allocated
 * between two 256-byte guard zones; this synchronises `stream`, reads the zones back and fails with DLWP_ERR_STATE when a
 * kernel wrote outside its buffer.  Used by the GPU tests after rollouts at the multi-tile sizes. */
int dlwp_plan_check_guards(const dlwp_plan* plan, void* stream);
/* Diagnostics: copy the first `max_bytes` bytes of a named internal workspace buffer ("specA", "specZ",
 * "specX", "tcw_in", "tcw_x0", "tcw_x1", "tcw_sa", "tcw_ys", "tcw_z", "tc_img0", "tc_img1") to the device buffer `dst`.
 * The caller must know the buffer's size; used by the stage-by-stage tests.  No reference counterpart. */
int dlwp_plan_debug_read(const dlwp_plan* plan, const char* name, void* dst, size_t max_bytes, void* stream);

/* Replaces load_state_dict (scripts/evaluate.py:148-149).  `name` is the reference
 * state_dict key without the wrapper prefix "fno." (e.g. "lifting.fcs.0.weight",
 * "fno_blocks.convs.weight.0.tensor" | ".core" | ".factors.factor_2", "blocks.1.filter.w1").
 * `data` may be a host or a device pointer (fp32, contiguous); it is copied before return
 * of the stream work, the caller keeps ownership.  Complex tensors are passed as real views
 * with a trailing dimension of 2. */
int dlwp_plan_set_tensor(dlwp_plan* plan, const char* name, const void* data, int ndim,
                         const int64_t* shape, void* stream);
/* Packs what set_tensor collected into kernel layouts (mode-major spectral weights,
 * Tucker -> dense, transposed pointwise weights).  Must be called before step / rollout. */
int dlwp_plan_commit_weights(dlwp_plan* plan, void* stream);

/* One backbone evaluation y = net(x): `self.fno(x_t)` (fno.py:103) or
 * forward_features + head + un-patchify (fourcastnet.py:345-357).
 * x: device [B, Cin, H, W]; y: device [B, Cg, H, W]. */
int dlwp_step(dlwp_plan* plan, const float* x, float* y, int batch, void* stream);

/* The whole autoregressive loop (fno.py:78-106 / fourcastnet.py:323-361, working form
 * :489-527): for t = ctx .. T-1: out_t = prog_t[:, -1] + net(cat(constants, prescribed
 * [t-ctx:t], prog_t)).  Device pointers; constants [B,1,Cc,H,W] or NULL, prescribed
 * [B,T,Cp,H,W] or NULL, prognostic [B,T,Cg,H,W]; outs [B, T-ctx, Cg, H, W] stays on device. */
int dlwp_rollout(dlwp_plan* plan, const float* constants, const float* prescribed,
                 const float* prognostic, int batch, int T, float* outs, void* stream);

/* Same call with HOST buffers (pinned or pageable; what scripts/evaluate.py:214-217,235-243 does
 * around the model): copies to plan-owned staging what the loop reads -- constants, the first `ctx`
 * frames of prescribed and of prognostic (later ground-truth prognostic frames are never read,
 * fno.py:83,91-95) --, runs the rollout while the later prescribed frames arrive on a second
 * stream just ahead of the step that first reads them and every lead time's frame goes back on a
 * third stream under the following steps, and synchronises `stream`.  This is the entry the end-to-end benchmark times.
 * May grow the staging buffers on first use. */
int dlwp_rollout_host(dlwp_plan* plan, const float* constants, const float* prescribed,
                      const float* prognostic, int batch, int T, float* outs, void* stream);

/* Latitude-weighted error sums (scripts/evaluate.py:790-820): adds, per lead time and
 * variable, sum w*(o-t)^2, sum w*(o-c)(t-c), sum w*(o-c)^2, sum w*(t-c)^2 over (sample,
 * lat, lon) into sums[T, C, 4] (device, fp64, caller zeroes).  outputs/targets: device
 * [B, T, C, H, W]; climatology: device [C, H, W] or NULL (= 0); lat_weights: device [H]. */
int dlwp_metrics_accumulate(const float* outputs, const float* targets, const float* climatology,
                            const float* lat_weights, int batch, int T, int C, int H, int W,
                            double* sums, void* stream);

/* Evaluation tail (scripts/evaluate.py:281-296 + the per-variable slicing of build_dataset, :366-383): undoes the
 * normalisation, out[c][b][t][:] = x[b][t][c][:] * std[c] + mean[c], and leaves one contiguous [B, T, H, W] array per
 * variable for the writer.  x: device [B, T, C, H, W]; mean / std: device [C] or NULL (0 / 1); out: device [C, B, T, H, W]. */
int dlwp_denormalize_split(const float* x, const float* mean, const float* std, float* out, int B, int T, int C, int H, int W,
                           void* stream);

/* Stand-alone dense Fourier layer y = irfft2(mix(rfft2(x)[bins])) for NCHW tensors:
 * unet.py:19-69 (SpectralConv2d; Ci != Co as unet.py:26-44 allows) and the neuralop SpectralConv.  weight: device complex
 * (real view) [Ci, Co, n_rows, n_cols, 2]; bias: device [Co] or NULL.  A convenience entry: one cudaMalloc / cudaFree and a
 * stream synchronisation per call (dlwp_spectral_conv2d_static allocates nothing; dlwp_fourier_residual_block takes a workspace). */
int dlwp_spectral_conv2d(const float* x, const float* weight, const float* bias, float* y,
                         int batch, int Ci, int Co, int H, int W, int n_rows, int n_cols,
                         const int32_t* rows_in, const int32_t* rows_out, int fft_norm,
                         void* stream);

/* The same layer for callers that cannot allocate (code under CUDA-graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.  Same arguments
 * and results as dlwp_spectral_conv2d (Ci != Co is accepted), but every intermediate lives in one file-scope device array
 * of DLWP_STATIC_SCRATCH_BYTES that each GPU loading the library holds until exit: the call does no cudaMalloc / cudaFree,
 * calls no allocating library routine and does not synchronise (it only enqueues kernels on `stream`).  Problems that
 * need more scratch, or more than DLWP_STATIC_MAX_ROWS retained rows, are rejected with DLWP_ERR_INVALID and a message
 * that states both sizes.  All calls share the array: calls that overlap on the GPU (other streams, other host threads)
 * race on it -- serialise them.  x, weight, bias, y must belong to the current device. */
#define DLWP_STATIC_SCRATCH_BYTES (16u << 20)
#define DLWP_STATIC_MAX_ROWS 64
int dlwp_spectral_conv2d_static(const float* x, const float* weight, const float* bias, float* y,
                                int batch, int Ci, int Co, int H, int W, int n_rows, int n_cols,
                                const int32_t* rows_in, const int32_t* rows_out, int fft_norm,
                                void* stream);

/* FourierResidualBlock.forward (models/unet/unet.py:780-836), the in-repo statement of the fused spectral pattern:
 *   h = act(norm1(x)); out = fourier1(h) + conv1(h); out = act(norm2(out)); y = fourier2(out) + conv2(out) + shortcut(x)
 * with act = exact GELU (the constructor default, :804), SpectralConv2d corners rows [:modes1] and [-modes1:], columns
 * [:modes2] (:60-65), 1x1 convolutions with bias, GroupNorm when the norm pointers are set (norm=True, :818-820) and the
 * shortcut convolution when in_channels != out_channels (:813-816).  All pointers are device pointers; weights in the
 * reference's state_dict shapes (complex weights as real views [.., 2]).  x [B, Ci, H, W] -> y [B, Co, H, W].
 * The call allocates nothing and does not synchronise: every intermediate lives in `workspace` (device,
 * dlwp_fourier_residual_block_workspace bytes, 256-byte aligned); the additions and the second GELU run in the epilogue of
 * the inverse transforms. */
typedef struct {
  const float *f1_w1, *f1_w2;       /* fourier1.weights1 / .weights2  [Ci, Co, modes1, modes2, 2] */
  const float *conv1_w, *conv1_b;   /* conv1.weight [Co, Ci, 1, 1], conv1.bias [Co] */
  const float *f2_w1, *f2_w2;       /* fourier2.weights1 / .weights2  [Co, Co, modes1, modes2, 2] */
  const float *conv2_w, *conv2_b;   /* conv2.weight [Co, Co, 1, 1], conv2.bias [Co] */
  const float *short_w, *short_b;   /* shortcut.weight [Co, Ci, 1, 1], .bias [Co]; NULL = Identity (Ci == Co) */
  const float *norm1_w, *norm1_b;   /* norm1.weight / .bias [Ci]; NULL = Identity (norm=False) */
  const float *norm2_w, *norm2_b;   /* norm2.weight / .bias [Co]; NULL = Identity */
  int32_t n_groups;                 /* GroupNorm groups (ignored without normalisation) */
  float eps;                        /* GroupNorm eps (torch default 1e-5) */
} dlwp_frb_weights;
int dlwp_fourier_residual_block_workspace(int batch, int Ci, int Co, int H, int W, int modes1, int modes2, size_t* bytes);
int dlwp_fourier_residual_block(const float* x, const dlwp_frb_weights* w, float* y, int batch, int Ci, int Co, int H, int W,
                                int modes1, int modes2, void* workspace, size_t workspace_bytes, void* stream);

/* ---- training primitives (scripts/train.py:263-271 back-propagates through the model) ----------------------
 * fp32 kernel set, channels-last activations [P = B*H*W, C]; dlwp_benchmark_b200/training.py wraps them in
 * torch.autograd.Function objects.  None of them allocates or synchronises.
 *
 * y = act(x W^T + b + pre), act 0 = none, 1 = exact GELU (neuralop MLP 1x1 convolutions, the FNO skip convolution with the
 * spectral branch as `pre`, fno_blocks epilogue).  x [P, K], W [N, K], b [N] | NULL, pre [P, N] | NULL, y [P, N];
 * z [P, N] | NULL receives the pre-activation values (saved for the backward pass). */
int dlwp_linear_forward(const float* x, const float* W, const float* b, const float* pre, float* y, float* z, int P, int K, int N,
                        int act, void* stream);
/* dz = dy * act'(z) (also the gradient of `pre`), dx = dz W (skipped when dx is NULL), dW = dz^T x, db = column sums of dz
 * (NULL: no bias).  wt_scratch: [K, N] floats (needed for dx). */
int dlwp_linear_backward(const float* dy, const float* z, const float* x, const float* W, float* dz, float* dx, float* dW, float* db,
                         float* wt_scratch, int P, int K, int N, int act, void* stream);
/* The dense Fourier layer y = irfft2(mix(rfft2(x)[bins])) (unet.py:15-69 / neuralop SpectralConv, no bias) on channels-last
 * tensors x [B, H, W, Ci] -> y [B, H, W, Co]; weight: complex (real view) [Ci, Co, n_rows, n_cols, 2].  backward: the adjoint of
 * the truncated-DFT chain gives dx (NULL: skipped), dweight[i, o, mode] = sum_b conj(X[mode, b, i]) dY[mode, b, o] in the
 * weight's own layout.  workspace: dlwp_spectral_train_workspace bytes (device, 256-byte aligned). */
int dlwp_spectral_train_workspace(int batch, int Ci, int Co, int H, int W, int n_rows, int n_cols, size_t* bytes);
int dlwp_spectral_forward(const float* x, const float* weight, float* y, int batch, int Ci, int Co, int H, int W, int n_rows,
                          int n_cols, const int32_t* rows_in, const int32_t* rows_out, int fft_norm, void* workspace,
                          size_t workspace_bytes, void* stream);
int dlwp_spectral_backward(const float* x, const float* weight, const float* dy, float* dx, float* dweight, int batch, int Ci, int Co,
                           int H, int W, int n_rows, int n_cols, const int32_t* rows_in, const int32_t* rows_out, int fft_norm,
                           void* workspace, size_t workspace_bytes, void* stream);
/* LayerNorm over the channel axis of channels-last tokens x [P, C] (fourcastnet.py:170,174: nn.LayerNorm(dim, eps=1e-6)),
 * biased variance, eps inside the square root.  backward recomputes the row statistics from x:
 * dx = rstd (g - mean(g) - xhat mean(g xhat)) with g = dy gamma; dgamma = column sums of dy xhat, dbeta = column sums of dy. */
int dlwp_layernorm_forward(const float* x, const float* gamma, const float* beta, float* y, int P, int C, float eps, void* stream);
int dlwp_layernorm_backward(const float* dy, const float* x, const float* gamma, float* dx, float* dgamma, float* dbeta, int P,
                            int C, float eps, void* stream);
/* The AFNO2D token mixer (fourcastnet.py:60-127) on channels-last tokens x [B, h, w, C]:
 *   y = irfft2( softshrink( relu(X w1 + b1) w2 + b2 ) ) + x (+ extra),  X = rfft2(x, norm="ortho")[rows, :n_cols]
 * with the block-diagonal complex weights in the reference's layout: w1 [2, nb, bs, bs f], b1 [2, nb, bs f], w2 [2, nb, bs f, bs],
 * b2 [2, nb, bs] (bs = C / nb), ReLU and softshrink on real and imaginary parts separately.  `rows` = the n_rows retained rows
 * (the same bins on the way back), `extra` [B, h, w, C] | NULL is added to the result (Block's second skip, :185-187).
 * backward: dx = adjoint chain + dy (NULL: skipped; the gradient of `extra` is dy itself), dw1 / db1 / dw2 / db2 in the
 * parameters' own layouts.  The spectrum is recomputed from x, nothing is saved between the two calls.
 * workspace: dlwp_afno_filter_workspace bytes (device, 256-byte aligned). */
int dlwp_afno_filter_workspace(int batch, int C, int h, int w, int n_rows, int n_cols, int hidden_factor, size_t* bytes);
int dlwp_afno_filter_forward(const float* x, const float* extra, const float* w1, const float* b1, const float* w2, const float* b2,
                             float* y, int batch, int C, int h, int w, int n_rows, int n_cols, const int32_t* rows, int num_blocks,
                             int hidden_factor, float lambda, void* workspace, size_t workspace_bytes, void* stream);
int dlwp_afno_filter_backward(const float* x, const float* w1, const float* b1, const float* w2, const float* b2, const float* dy,
                              float* dx, float* dw1, float* db1, float* dw2, float* db2, int batch, int C, int h, int w, int n_rows,
                              int n_cols, const int32_t* rows, int num_blocks, int hidden_factor, float lambda, void* workspace,
                              size_t workspace_bytes, void* stream);
/* layout changes between the reference's NCHW tensors and the channels-last form above */
int dlwp_nchw_to_nhwc(const float* in, float* out, int B, int C, int HW, void* stream);
int dlwp_nhwc_to_nchw(const float* in, float* out, int B, int C, int HW, void* stream);

/* Number of kernels this library has launched in this process (for bench "gpu_launches"). */
int64_t dlwp_launch_count(void);

/* Per-kernel device timing for bench.py's roofline figure (new; no reference counterpart).
 * Between _begin and _end every kernel this library launches is followed by a CUDA event on
 * its stream; _end synchronises `stream` and aggregates event-to-event times by kernel name.
 * Not for the hot path: a profiled region is not graph-capturable. */
int dlwp_profile_begin(void* stream);
int dlwp_profile_end(void* stream);
int dlwp_profile_count(void);
int dlwp_profile_entry(int index, const char** name, double* total_ms, int64_t* launches);

/* Diagnostic: one tcgen05.mma tile D[M x N] = A[M x K] * B[N x K]^T (operands rounded to bf16,
 * fp32 accumulate in tensor memory); writes the raw 128-lane x N-column TMEM image to `d_raw`
 * (device, [128, N] fp32).  Used by the GPU tests to pin the descriptor encodings. */
int dlwp_selftest_umma(const float* a, const float* b, float* d_raw, int M, int N, int K, void* stream);
/* Same product with B given as [K, N] row-major and staged the way the kernels keep an activation
 * tile (k rows, 16-byte chunks of 8 consecutive n), read by tcgen05.mma as an MN-major operand.
 * variant 1: descriptor LBO = K stride, SBO = N stride of the 8x8 core matrices (what the library
 * uses); variant 2: swapped (diagnostic). */
int dlwp_selftest_umma_bmn(const float* a, const float* b_kn, float* d_raw, int M, int N, int K, int variant,
                           void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DLWP_B200_H */
