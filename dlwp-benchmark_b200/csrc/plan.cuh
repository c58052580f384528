// Plan object behind the C ABI (include/dlwp_b200.h): owns the model description, the
// truncated-DFT bases, the packed weights and every workspace buffer.
#pragma once
#include <map>
#include <string>
#include <utility>
#include <vector>

#include "kernels.cuh"
#include "tc_fno.cuh"
#include "tc_afno.cuh"
#include "tc_wide.cuh"

namespace dlwp {

struct NamedTensor {
  std::vector<int64_t> shape;
  float* data = nullptr;  // device, plan-owned
  int64_t numel = 0;
};

// Where one backbone evaluation reads its input from and writes its output to.
struct StepIO {
  SrcGroup grp[4];
  int ngrp = 0;
  float* out = nullptr;        // NCHW frame(s): out[b * out_bstride + (c, y, x)]
  long long out_bstride = 0;
  const float* last = nullptr; // residual frame (prognostic_t[:, -1]) or null
  long long last_bstride = 0;
};

}  // namespace dlwp

struct dlwp_plan {
  dlwp_model_desc d{};
  int device = 0;
  int Cin = 0;          // Cc + (Cp + Cg) * ctx
  int gh = 0, gw = 0;   // grid the spectral transform runs on (tokens: H/p1, W/p2)
  int C = 0;            // channel width of the spectral layers (hidden_channels | embed_dim)
  int n_spec = 0;       // number of spectral layers (n_layers | depth)
  std::vector<int32_t> rows_in, rows_out;
  bool committed = false;

  std::map<std::string, dlwp::NamedTensor> tensors;
  std::vector<void*> owned;   // every cudaMalloc of this plan except `tensors` and staging
  std::vector<std::pair<uint8_t*, size_t>> guarded;   // (buffer, bytes) of every plan_dev_alloc: guard zones on both sides
  size_t workspace_bytes = 0;

  // truncated-DFT bases (float2, built in fp64)
  float2 *Fw = nullptr, *Fh = nullptr, *Gh = nullptr, *Gw = nullptr;

  // FNO: mode-major spectral weights per layer [R, Kc, C, C]
  std::vector<float2*> wm;

  // workspace
  float *act0 = nullptr, *act1 = nullptr, *act2 = nullptr, *act3 = nullptr;  // [P, C]
  float* hid = nullptr;                                                    // [P, max hidden]
  float2 *specA = nullptr, *specZ = nullptr;                               // [B, gh, Kc, C]
  float2 *specX = nullptr, *specY = nullptr;                               // [R, Kc, B, C]

  // tensor-core path (tc_fno.cuh): used when precision == DLWP_PREC_FP32_TC and the shape fits
  bool tc = false;
  dlwp::TcGeom tcg{};
  uint8_t* tc_img[2] = {nullptr, nullptr};              // activation tile images (ping-pong)
  uint8_t *tc_G = nullptr, *tc_F = nullptr;             // inverse / forward W-DFT operand images
  uint8_t* tc_Z = nullptr;                              // Z' operand images between fno_mid_kernel and fno_layer_kernel
  uint8_t *tc_liftW1 = nullptr, *tc_liftW2 = nullptr, *tc_projW1 = nullptr, *tc_projW2 = nullptr,
          *tc_projW1x = nullptr;
  std::vector<uint8_t*> tc_skip;                        // per layer skip-conv images
  bool tc_mid2 = false;                                 // the retained-column products run on tcgen05 (fno_mid_tc_kernel, tc_mid2.cu)
  bool tc_layer_rev = false;                            // layer kernels walk the tile list alternately down and up
  bool tc_l2hint = false;                               // L2 eviction priorities on the activation images' bulk copies
  bool tc_lift4 = false;                                // lifting kernel with four workers per CTA (fno_lift4_kernel)
  bool tc_fuse_proj = false;                            // last layer + projection in one launch (fno_layerproj_kernel)
  uint8_t *tc_midF = nullptr, *tc_midG = nullptr;       // (re|im)-stacked forward / inverse H-DFT operand images
  std::vector<uint8_t*> tc_midW;                        // per layer mode-weight operand images [Kc][R][8 KB]

  // bf16 tensor-core path at hidden width 128 (tc_wide.cuh): precision == DLWP_PREC_BF16
  bool tcw = false;
  int tcw_TW = 0, tcw_KZ = 0, tcw_HB = 0;               // tiles per grid row; K extent of the inverse W-DFT operands; blocks of 64 grid rows
  uint8_t *tcw_in = nullptr, *tcw_z = nullptr;
  uint8_t* tcw_x[2] = {nullptr, nullptr};               // activation tile images (ping-pong)
  uint8_t *tcw_G = nullptr, *tcw_F = nullptr;           // inverse / forward W-DFT tiles
  uint8_t *tcw_Gs = nullptr, *tcw_Fs = nullptr;         // inverse / forward H-DFT tiles ((re | im)-stacked)
  uint8_t *tcw_sa = nullptr, *tcw_ys = nullptr;         // spectrum images after the W-DFT / after the mode mixing
  uint8_t *tcw_W1 = nullptr, *tcw_W2 = nullptr, *tcw_b2 = nullptr, *tcw_P1 = nullptr, *tcw_pb1 = nullptr,
          *tcw_P2 = nullptr, *tcw_pb2 = nullptr;
  std::vector<uint8_t*> tcw_skip, tcw_cbias;            // per layer skip-conv and bias images
  std::vector<uint32_t*> tcw_wm16;                      // per layer bf16 form of the mode-major spectral weights: (re, im) pairs, or
  bool tcw_mix_tc = false;                              //   (tcw_mix_tc) the operand images of mix_few_tc_kernel (same size)

  // tensor-core AFNO kernels (tc_afno.cuh): precision == DLWP_PREC_FP32_TC, embed_dim 64, mlp hidden 256
  bool atc = false;
  bool atc_spec = false;                                // the fused token-mixer kernels (afno_fused.cu) cover the grid / modes too
  std::vector<uint8_t*> atc_W1, atc_W2;                 // per block: fc1 * diag(norm2.weight), fc2 / 2 operand images
  std::vector<std::vector<float>> atc_b1h, atc_b2h;     // per block, host: fc1.bias + fc1 * norm2.bias; fc2.bias (kernel parameters)
  bool atc_wtc = false;                                 // the token mixer's W-DFTs run on tcgen05 (afno_fwd_tc / afno_inv_tc)
  int atc_RT = 0, atc_KZ = 0;                           // grid rows per 128-token tile; K extent of the inverse operand
  uint8_t *atc_F = nullptr, *atc_G = nullptr;           // forward / inverse W-DFT operand images
  bool atc_mid2 = false;                                // H-DFT + block MLP + inverse H-DFT on tcgen05 (afno_mid_tc_kernel)
  uint8_t *atc_midF = nullptr, *atc_midG = nullptr;     // its Fs / Gs operand images
  std::vector<uint8_t*> atc_midW1, atc_midW2;           // per block: (re|im)-stacked filter weights, operand images per channel block
  std::vector<float*> atc_midb;                         // per block: filter biases b1 | b2 in accumulator column order
  uint8_t* atc_Zimg = nullptr;                          // Z' operand images per 128-token tile (the A images overlay specA)

  // host-entry staging (dlwp_rollout_host)
  void* stage[4] = {nullptr, nullptr, nullptr, nullptr};
  size_t stage_bytes[4] = {0, 0, 0, 0};
  cudaStream_t copy_stream = nullptr;        // device -> host frames, under the following steps' compute
  std::vector<cudaEvent_t> step_done;        // one per lead time + one for the end of the copies
  cudaStream_t h2d_stream = nullptr;         // host -> device prescribed frames of later steps, under the earlier steps' compute
  std::vector<cudaEvent_t> part_ev;          // dlwp_rollout_host: per part of the members, initial frames there / step done
  std::vector<cudaEvent_t> frame_ready;      // one per prescribed frame
};

namespace dlwp {
int plan_build(dlwp_plan* p);                                  // bases + workspace
int plan_commit(dlwp_plan* p, cudaStream_t st);                // pack weights
int plan_backbone(dlwp_plan* p, const StepIO& io, int B, cudaStream_t st);
int plan_tc_build(dlwp_plan* p);                               // tc_plan.cu: eligibility + DFT operand images
int plan_tc_commit(dlwp_plan* p, cudaStream_t st);             // tc_plan.cu: weight images
int backbone_fno_tc(dlwp_plan* p, const StepIO& io, int B, cudaStream_t st);
int plan_tcw_build(dlwp_plan* p);                              // tc_wide_plan.cu
int plan_tcw_commit(dlwp_plan* p, cudaStream_t st);
int backbone_fno_tcw(dlwp_plan* p, const StepIO& io, int B, cudaStream_t st);
int plan_atc_build(dlwp_plan* p);                              // tc_afno_plan.cu
int plan_atc_commit(dlwp_plan* p, cudaStream_t st);
int atc_block_filter(dlwp_plan* p, int block, float* x, float* out, int B, cudaStream_t st, const dlwp::StepIO* embed);
int atc_block_mlp(dlwp_plan* p, int block, const float* x, float* out, int P, cudaStream_t st, const dlwp::StepIO* head);
int spectral_mid_fno(dlwp_plan* p, const float2* wm, int B, cudaStream_t st);
int plan_dev_alloc(dlwp_plan* p, void** out, size_t bytes);
int plan_check_guards(const dlwp_plan* p, int* bad_index);
constexpr size_t kGuardBytes = 256;      // keeps the 256-byte alignment of cudaMalloc
constexpr int kGuardByte = 0xA5;
int plan_need(const dlwp_plan* p, const std::string& name, int64_t numel, const float** out);
int spectral_layer(dlwp_plan* p, const float* x, float* y, const float2* wm_or_null, int layer,
                   const float* add1, const float* add2, const float* bias, int act, int B, cudaStream_t st);
}  // namespace dlwp
