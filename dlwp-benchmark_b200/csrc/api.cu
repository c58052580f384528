// extern "C" surface of libdlwp_b200.so (include/dlwp_b200.h).  Nothing here throws.
#include <new>

#include <nvtx3/nvToolsExt.h>      // header-only; the ranges cost a null-pointer test unless a profiler is attached

#include "plan.cuh"

using namespace dlwp;

namespace {

int check_device(int device) {
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n <= 0) {
    cudaGetLastError();
    return fail(DLWP_ERR_NO_DEVICE, "no CUDA device is usable (%s); this library has no CPU fallback",
                e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
  }
  if (device < 0 || device >= n) return fail(DLWP_ERR_INVALID, "device %d out of range [0, %d)", device, n);
  cudaDeviceProp prop;
  DLWP_CUDA(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10)
    return fail(DLWP_ERR_NO_DEVICE, "device %d is sm_%d%d; this library is built for sm_100a only", device,
                prop.major, prop.minor);
  DLWP_CUDA(cudaSetDevice(device));
  return DLWP_OK;
}

void free_plan(dlwp_plan* p) {
  if (!p) return;
  cudaSetDevice(p->device);
  for (void* q : p->owned) cudaFree(q);
  for (auto& kv : p->tensors) cudaFree(kv.second.data);
  for (int i = 0; i < 4; ++i)
    if (p->stage[i]) cudaFree(p->stage[i]);
  for (cudaEvent_t e : p->step_done) cudaEventDestroy(e);
  if (p->copy_stream) cudaStreamDestroy(p->copy_stream);
  for (cudaEvent_t e : p->frame_ready) cudaEventDestroy(e);
  if (p->h2d_stream) cudaStreamDestroy(p->h2d_stream);
  for (cudaEvent_t e : p->part_ev) cudaEventDestroy(e);
  delete p;
}

// Frame bookkeeping of the autoregressive loop (fno.py:78-100 in the working form of
// fourcastnet.py:489-521): which NCHW planes make up x_t and prognostic_t[:, -1].
void rollout_io(const dlwp_plan* p, const float* constants, const float* prescribed, const float* prognostic,
                float* outs, int T, int t, StepIO* io) {
  const dlwp_model_desc& d = p->d;
  const long long HW = (long long)d.height * d.width;
  const int ctx = d.context_size, Cc = d.constant_channels, Cp = d.prescribed_channels,
            Cg = d.prognostic_channels, steps = T - ctx;
  int g = 0;
  if (constants && Cc > 0) io->grp[g++] = SrcGroup{constants, Cc * HW, Cc};
  if (prescribed && Cp > 0)
    io->grp[g++] = SrcGroup{prescribed + (long long)(t - ctx) * Cp * HW, (long long)T * Cp * HW, ctx * Cp};
  const int own = (t - ctx) < ctx ? (t - ctx) : ctx;  // frames taken from the model's own outputs
  const int gt = ctx - own;                           // ground-truth context frames still in use
  if (gt > 0)
    io->grp[g++] = SrcGroup{prognostic + (long long)(t - ctx) * Cg * HW, (long long)T * Cg * HW, gt * Cg};
  if (own > 0)
    io->grp[g++] = SrcGroup{outs + (long long)(t - ctx - own) * Cg * HW, (long long)steps * Cg * HW, own * Cg};
  io->ngrp = g;
  io->out = outs + (long long)(t - ctx) * Cg * HW;
  io->out_bstride = (long long)steps * Cg * HW;
  if (own > 0) {
    io->last = outs + (long long)(t - ctx - 1) * Cg * HW;
    io->last_bstride = (long long)steps * Cg * HW;
  } else {
    io->last = prognostic + (long long)(ctx - 1) * Cg * HW;
    io->last_bstride = (long long)T * Cg * HW;
  }
}

int grow(dlwp_plan* p, int slot, size_t bytes) {
  if (bytes <= p->stage_bytes[slot]) return DLWP_OK;
  if (p->stage[slot]) cudaFree(p->stage[slot]);
  p->stage[slot] = nullptr;
  p->stage_bytes[slot] = 0;
  DLWP_CUDA(cudaMalloc(&p->stage[slot], bytes));
  p->stage_bytes[slot] = bytes;
  return DLWP_OK;
}

}  // namespace

extern "C" {

int dlwp_abi_version(void) { return DLWP_ABI_VERSION; }
const char* dlwp_last_error(void) { return g_last_error.c_str(); }
int64_t dlwp_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

int dlwp_plan_create(const dlwp_model_desc* desc, int device, dlwp_plan** out) {
  if (!desc || !out) return fail(DLWP_ERR_INVALID, "dlwp_plan_create: null argument");
  *out = nullptr;
  DLWP_TRY(check_device(device));
  dlwp_plan* p = new (std::nothrow) dlwp_plan();
  if (!p) return fail(DLWP_ERR_INVALID, "out of host memory");
  p->d = *desc;
  p->device = device;
  int rc = plan_build(p);
  if (rc != DLWP_OK) {
    free_plan(p);
    return rc;
  }
  *out = p;
  return DLWP_OK;
}

int dlwp_plan_destroy(dlwp_plan* plan) {
  free_plan(plan);
  return DLWP_OK;
}

int dlwp_plan_workspace_bytes(const dlwp_plan* plan, size_t* bytes) {
  if (!plan || !bytes) return fail(DLWP_ERR_INVALID, "dlwp_plan_workspace_bytes: null argument");
  *bytes = plan->workspace_bytes;
  return DLWP_OK;
}

int dlwp_plan_uses_tensor_cores(const dlwp_plan* plan, int* out) {
  if (!plan || !out) return fail(DLWP_ERR_INVALID, "dlwp_plan_uses_tensor_cores: null argument");
  *out = (plan->tc || plan->tcw) ? 1 : 0;
  return DLWP_OK;
}

int dlwp_plan_check_guards(const dlwp_plan* plan, void* stream) {
  if (!plan) return fail(DLWP_ERR_INVALID, "dlwp_plan_check_guards: null plan");
  DLWP_CUDA(cudaSetDevice(plan->device));
  DLWP_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
  int bad = -1;
  return plan_check_guards(plan, &bad);
}

int dlwp_plan_debug_read(const dlwp_plan* plan, const char* name, void* dst, size_t max_bytes, void* stream) {
  if (!plan || !name || !dst) return fail(DLWP_ERR_INVALID, "dlwp_plan_debug_read: null argument");
  const std::string n(name);
  const void* src = nullptr;
  if (n == "tcw_in") src = plan->tcw_in;
  else if (n == "tcw_x0") src = plan->tcw_x[0];
  else if (n == "tcw_x1") src = plan->tcw_x[1];
  else if (n == "tcw_z") src = plan->tcw_z;
  else if (n == "tcw_sa") src = plan->tcw_sa;
  else if (n == "tcw_ys") src = plan->tcw_ys;
  else if (n == "specX") src = plan->specX;
  else if (n == "specA") src = plan->specA;
  else if (n == "specZ") src = plan->specZ;
  else if (n == "tc_img0") src = plan->tc_img[0];
  else if (n == "tc_img1") src = plan->tc_img[1];
  if (!src) return fail(DLWP_ERR_INVALID, "dlwp_plan_debug_read: no buffer '%s' in this plan", name);
  DLWP_CUDA(cudaMemcpyAsync(dst, src, max_bytes, cudaMemcpyDeviceToDevice, static_cast<cudaStream_t>(stream)));
  return DLWP_OK;
}

int dlwp_plan_set_tensor(dlwp_plan* plan, const char* name, const void* data, int ndim, const int64_t* shape,
                         void* stream) {
  if (!plan || !name || !data || ndim < 0 || (ndim > 0 && !shape))
    return fail(DLWP_ERR_INVALID, "dlwp_plan_set_tensor: null argument");
  DLWP_CUDA(cudaSetDevice(plan->device));
  int64_t n = 1;
  for (int i = 0; i < ndim; ++i) {
    DLWP_REQUIRE(shape[i] >= 0, "tensor '%s': negative dimension", name);
    n *= shape[i];
  }
  NamedTensor& t = plan->tensors[name];
  if (t.numel != n || !t.data) {
    if (t.data) cudaFree(t.data);
    t.data = nullptr;
    DLWP_CUDA(cudaMalloc((void**)&t.data, (size_t)(n > 0 ? n : 1) * sizeof(float)));
  }
  t.numel = n;
  t.shape.assign(shape, shape + ndim);
  if (n > 0)
    DLWP_CUDA(cudaMemcpyAsync(t.data, data, (size_t)n * sizeof(float), cudaMemcpyDefault, (cudaStream_t)stream));
  plan->committed = false;
  return DLWP_OK;
}

int dlwp_plan_commit_weights(dlwp_plan* plan, void* stream) {
  if (!plan) return fail(DLWP_ERR_INVALID, "dlwp_plan_commit_weights: null plan");
  DLWP_CUDA(cudaSetDevice(plan->device));
  return plan_commit(plan, (cudaStream_t)stream);
}

int dlwp_step(dlwp_plan* plan, const float* x, float* y, int batch, void* stream) {
  if (!plan || !x || !y) return fail(DLWP_ERR_INVALID, "dlwp_step: null argument");
  DLWP_CUDA(cudaSetDevice(plan->device));
  const long long HW = (long long)plan->d.height * plan->d.width;
  StepIO io;
  io.grp[0] = SrcGroup{x, plan->Cin * HW, plan->Cin};
  io.ngrp = 1;
  io.out = y;
  io.out_bstride = plan->d.prognostic_channels * HW;
  nvtxRangePushA("dlwp_step");
  const int rc = plan_backbone(plan, io, batch, (cudaStream_t)stream);
  nvtxRangePop();
  return rc;
}

int dlwp_rollout(dlwp_plan* plan, const float* constants, const float* prescribed, const float* prognostic,
                 int batch, int T, float* outs, void* stream) {
  if (!plan || !prognostic || !outs) return fail(DLWP_ERR_INVALID, "dlwp_rollout: null prognostic / outs");
  DLWP_CUDA(cudaSetDevice(plan->device));
  const dlwp_model_desc& d = plan->d;
  DLWP_REQUIRE(T > d.context_size, "sequence length %d must exceed context_size %d", T, d.context_size);
  DLWP_REQUIRE((d.constant_channels == 0) == (constants == nullptr),
               "constants pointer does not match constant_channels=%d", d.constant_channels);
  DLWP_REQUIRE((d.prescribed_channels == 0) == (prescribed == nullptr),
               "prescribed pointer does not match prescribed_channels=%d", d.prescribed_channels);
  nvtxRangePushA("dlwp_rollout");
  int rc = DLWP_OK;
  for (int t = d.context_size; t < T && rc == DLWP_OK; ++t) {
    nvtxRangePushA("dlwp_rollout: lead step");
    StepIO io;
    rollout_io(plan, constants, prescribed, prognostic, outs, T, t, &io);
    rc = plan_backbone(plan, io, batch, (cudaStream_t)stream);
    nvtxRangePop();
  }
  nvtxRangePop();
  return rc;
}

int dlwp_rollout_host(dlwp_plan* plan, const float* constants, const float* prescribed, const float* prognostic,
                      int batch, int T, float* outs, void* stream) {
  if (!plan || !prognostic || !outs) return fail(DLWP_ERR_INVALID, "dlwp_rollout_host: null prognostic / outs");
  const dlwp_model_desc& d = plan->d;
  DLWP_REQUIRE(batch > 0 && batch <= d.max_batch, "batch %d outside [1, max_batch=%d]", batch, d.max_batch);
  DLWP_REQUIRE(T > d.context_size, "sequence length %d must exceed context_size %d", T, d.context_size);
  DLWP_REQUIRE((d.constant_channels == 0) == (constants == nullptr),
               "constants pointer does not match constant_channels=%d", d.constant_channels);
  DLWP_REQUIRE((d.prescribed_channels == 0) == (prescribed == nullptr),
               "prescribed pointer does not match prescribed_channels=%d", d.prescribed_channels);
  DLWP_CUDA(cudaSetDevice(plan->device));
  cudaStream_t st = (cudaStream_t)stream;
  const size_t HW = (size_t)d.height * d.width, B = (size_t)batch, frame = HW * sizeof(float);
  const int ctx = d.context_size, steps = T - ctx;
  const size_t nb[4] = {constants ? B * d.constant_channels * frame : 0,
                        prescribed ? B * T * d.prescribed_channels * frame : 0,
                        B * T * d.prognostic_channels * frame,
                        B * steps * d.prognostic_channels * frame};
  for (int i = 0; i < 4; ++i) DLWP_TRY(grow(plan, i, nb[i]));
  if (!plan->copy_stream) DLWP_CUDA(cudaStreamCreateWithFlags(&plan->copy_stream, cudaStreamNonBlocking));
  while ((int)plan->step_done.size() < steps + 1) {
    cudaEvent_t e = nullptr;
    DLWP_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    plan->step_done.push_back(e);
  }
  // Large batches: the first and the last step run in K parts of the members, so that the later parts' initial frames
  // arrive under the first part's first step and the earlier parts' last frames leave under the later parts' last
  // step (the other steps run on the whole batch: every kernel launch has a fixed cost).
  const int K = (steps >= 2 && batch >= 64 && batch % 2 == 0) ? 2 : 1;      // (four parts at 256 members measured no better than two)
  const size_t Bp = B / K;      // members per part
  if (!plan->h2d_stream) DLWP_CUDA(cudaStreamCreateWithFlags(&plan->h2d_stream, cudaStreamNonBlocking));
  while ((int)plan->part_ev.size() < 2 * K) {      // [k]: part k's initial frames are on the device; [K + k]: part k's step is done
    cudaEvent_t e = nullptr;
    DLWP_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    plan->part_ev.push_back(e);
  }
  const size_t const_row = (size_t)d.constant_channels * frame;
  const size_t presc_pitch = (size_t)T * d.prescribed_channels * frame, presc_row = (size_t)d.prescribed_channels * frame;
  const size_t prog_pitch = (size_t)T * d.prognostic_channels * frame;
  // host -> device: only what the loop reads.  Of `prognostic` that is the first `ctx` frames of every sample
  // (fno.py:83,91-95 never look at later ground-truth frames), kept at their place in the [B, T, Cg, H, W] staging
  // layout so the frame bookkeeping is the device entry's.  Part 0 goes on the caller's stream, the others on the
  // copy-in stream ahead of the later prescribed frames.
  for (int k = 0; k < K; ++k) {
    cudaStream_t cs_ = k == 0 ? st : plan->h2d_stream;
    const size_t b0 = (size_t)k * Bp;
    if (nb[0])
      DLWP_CUDA(cudaMemcpyAsync((char*)plan->stage[0] + b0 * const_row, (const char*)constants + b0 * const_row, Bp * const_row,
                                cudaMemcpyHostToDevice, cs_));
    if (nb[1])
      DLWP_CUDA(cudaMemcpy2DAsync((char*)plan->stage[1] + b0 * presc_pitch, presc_pitch, (const char*)prescribed + b0 * presc_pitch,
                                  presc_pitch, (size_t)ctx * presc_row, Bp, cudaMemcpyHostToDevice, cs_));
    DLWP_CUDA(cudaMemcpy2DAsync((char*)plan->stage[2] + b0 * prog_pitch, prog_pitch, (const char*)prognostic + b0 * prog_pitch, prog_pitch,
                                (size_t)ctx * d.prognostic_channels * frame, Bp, cudaMemcpyHostToDevice, cs_));
    if (k > 0) DLWP_CUDA(cudaEventRecord(plan->part_ev[k], plan->h2d_stream));
  }
  // prescribed: step t reads frames [t - ctx, t) (fno.py:86,98).  Frame f >= ctx is first needed by step f + 1 and
  // arrives on the copy-in stream under the compute of the earlier steps (the last frame is never read).
  if (nb[1]) {
    while ((int)plan->frame_ready.size() < T) {
      cudaEvent_t e = nullptr;
      DLWP_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
      plan->frame_ready.push_back(e);
    }
    for (int f = ctx; f < T - 1; ++f) {
      const size_t off = (size_t)f * presc_row;
      DLWP_CUDA(cudaMemcpy2DAsync((char*)plan->stage[1] + off, presc_pitch, (const char*)prescribed + off, presc_pitch, presc_row,
                                  B, cudaMemcpyHostToDevice, plan->h2d_stream));
      DLWP_CUDA(cudaEventRecord(plan->frame_ready[f], plan->h2d_stream));
    }
  }
  // the loop of dlwp_rollout; every lead time's frame starts its way back to the host on a second
  // stream as soon as its projection kernel is done, under the following steps' compute
  float* dev_outs = (float*)plan->stage[3];
  const size_t out_pitch = (size_t)steps * d.prognostic_channels * frame, out_row = (size_t)d.prognostic_channels * frame;
  const float* dc = constants ? (const float*)plan->stage[0] : nullptr;
  const float* dp = prescribed ? (const float*)plan->stage[1] : nullptr;
  const float* dg = (const float*)plan->stage[2];
  const size_t cs = (size_t)d.constant_channels * HW, ps = (size_t)T * d.prescribed_channels * HW,
               gs = (size_t)T * d.prognostic_channels * HW, os = (size_t)steps * d.prognostic_channels * HW;      // floats per member
  auto run = [&](int t, size_t b0, size_t nbm) -> int {      // step t on members [b0, b0 + nbm)
    StepIO io;
    rollout_io(plan, dc ? dc + b0 * cs : nullptr, dp ? dp + b0 * ps : nullptr, dg + b0 * gs, dev_outs + b0 * os, T, t, &io);
    return plan_backbone(plan, io, (int)nbm, st);
  };
  auto send_back = [&](int t, size_t b0, size_t nbm, cudaEvent_t ev) -> int {      // lead time t's frames of these members -> host
    DLWP_CUDA(cudaEventRecord(ev, st));
    DLWP_CUDA(cudaStreamWaitEvent(plan->copy_stream, ev, 0));
    const size_t off = b0 * os + (size_t)(t - ctx) * d.prognostic_channels * HW;
    DLWP_CUDA(cudaMemcpy2DAsync(outs + off, out_pitch, dev_outs + off, out_pitch, out_row, nbm, cudaMemcpyDeviceToHost,
                                plan->copy_stream));
    return DLWP_OK;
  };
  for (int t = ctx; t < T; ++t) {
    if (nb[1] && t - 1 >= ctx) DLWP_CUDA(cudaStreamWaitEvent(st, plan->frame_ready[t - 1], 0));
    if (K == 1 || (t != ctx && t != T - 1)) {
      DLWP_TRY(run(t, 0, B));
      DLWP_TRY(send_back(t, 0, B, plan->step_done[t - ctx]));
    } else {
      for (int k = 0; k < K; ++k) {
        if (t == ctx && k > 0) DLWP_CUDA(cudaStreamWaitEvent(st, plan->part_ev[k], 0));
        DLWP_TRY(run(t, (size_t)k * Bp, Bp));
        DLWP_TRY(send_back(t, (size_t)k * Bp, Bp, k + 1 < K ? plan->part_ev[K + k] : plan->step_done[t - ctx]));
      }
    }
  }
  cudaEvent_t fin = plan->step_done[steps];
  DLWP_CUDA(cudaEventRecord(fin, plan->copy_stream));
  DLWP_CUDA(cudaStreamWaitEvent(st, fin, 0));
  DLWP_CUDA(cudaStreamSynchronize(st));
  return DLWP_OK;
}

int dlwp_metrics_accumulate(const float* outputs, const float* targets, const float* climatology,
                            const float* lat_weights, int batch, int T, int C, int H, int W, double* sums,
                            void* stream) {
  if (!outputs || !targets || !lat_weights || !sums)
    return fail(DLWP_ERR_INVALID, "dlwp_metrics_accumulate: null argument");
  return launch_metrics(outputs, targets, climatology, lat_weights, batch, T, C, H, W, sums, (cudaStream_t)stream);
}

}  // extern "C"
