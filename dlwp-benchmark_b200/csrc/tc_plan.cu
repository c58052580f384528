// Plan-side glue of the tensor-core FNO path: eligibility, constant operand images, weight
// images, and the per-step kernel sequence (reference: neuralop FNO.forward behind fno.py:103).
#include <cmath>
#include <cstdlib>
#include <vector>

#include "plan.cuh"

namespace dlwp {

// The tcgen05 kernels cover the reference's FNO2D / TFNO2D configurations (hidden width 32,
// 256-wide lifting / projection, grids whose rows divide a 128-pixel tile).  Anything else
// keeps the fp32 kernel set of kernels_fp32.cu.
int plan_tc_build(dlwp_plan* p) {
  const dlwp_model_desc& d = p->d;
  p->tc = false;
  if (d.kind != DLWP_MODEL_FNO || p->C != kTcC) return DLWP_OK;
  const int H = d.height, W = d.width, Kc = d.n_cols;
  if (W > kTcTile || kTcTile % W != 0 || (H * W) % kTcTile != 0) return DLWP_OK;
  const int RT = kTcTile / W;
  if (RT > 4) return DLWP_OK;
  const int KXP = 32 / RT;
  if (Kc < 1 || 2 * Kc > KXP || d.n_rows < 1) return DLWP_OK;
  if ((d.lifting_channels != 128 && d.lifting_channels != 256) ||
      (d.projection_channels != 128 && d.projection_channels != 256))
    return DLWP_OK;
  if (p->Cin < 1 || p->Cin > 15 || d.prognostic_channels > 16) return DLWP_OK;
  // fno_mid_kernel: one warp per retained row, 8 grid rows per output item; the layer kernel's
  // inverse W-DFT operands (2 K slots per retained column and grid row of the tile) must fit beside its tile buffers
  const int KZ = (2 * RT * Kc + 15) & ~15;
  if (d.n_rows > 16 || H % 8 != 0 || KZ > 48 || tc_mid_smem_bytes(H, d.n_rows) > 220 * 1024) return DLWP_OK;

  p->tcg = TcGeom{0, H * W, H, W, RT, KXP, Kc, KZ, 0, 0, H / RT};
  double s = 1.0, t = 1.0;
  const double hw = (double)H * (double)W;
  if (d.fft_norm == DLWP_FFT_NORM_BACKWARD) t = 1.0 / hw;
  else if (d.fft_norm == DLWP_FFT_NORM_FORWARD) s = 1.0 / hw;
  else s = t = 1.0 / std::sqrt(hw);

  // inverse W-DFT operand G'[(h2, w), (kx, h2', re|im)]: y[w] = sum_kx Re(Gw[w, kx] * Z[kx]) with
  // Gw = t * c_kx * exp(+2 pi i kx w / W) (c = 1 for the self-conjugate columns, else 2), K index
  // (kx * RT + h2') * 2 + (re|im) as fno_mid_kernel lays Z' out;
  // forward operand F''[(h2, kx'), (h2', w)] with Fw = s * exp(-2 pi i kx w / W)
  std::vector<double> G((size_t)128 * KZ, 0.0), F((size_t)32 * 128, 0.0);
  for (int h2 = 0; h2 < RT; ++h2)
    for (int w = 0; w < W; ++w)
      for (int kx = 0; kx < Kc; ++kx) {
        const long long r = ((long long)kx * w) % W;
        const double ang = 2.0 * M_PI * (double)r / (double)W;
        const bool self_conj = (kx == 0) || (W % 2 == 0 && kx == W / 2);
        const double ck = t * (self_conj ? 1.0 : 2.0);
        const int m = h2 * W + w, k = h2 * KXP + 2 * kx, kz = kx * 2 * RT + h2 * 2;
        G[(size_t)m * KZ + kz] = ck * std::cos(ang);
        G[(size_t)m * KZ + kz + 1] = -ck * std::sin(ang);
        F[(size_t)k * 128 + m] = s * std::cos(ang);
        F[(size_t)(k + 1) * 128 + m] = -s * std::sin(ang);
      }
  std::vector<uint8_t> gi((size_t)2 * 128 * KZ * 2), fi(2 * 32 * 128 * 2);
  tc_pack_images_f64(G.data(), 128, KZ, KZ, 128, KZ, gi.data());
  tc_pack_images_f64(F.data(), 32, 128, 128, 32, 128, fi.data());
  DLWP_TRY(plan_dev_alloc(p, (void**)&p->tc_G, gi.size()));
  DLWP_TRY(plan_dev_alloc(p, (void**)&p->tc_F, fi.size()));
  DLWP_CUDA(cudaMemcpy(p->tc_G, gi.data(), gi.size(), cudaMemcpyHostToDevice));
  DLWP_CUDA(cudaMemcpy(p->tc_F, fi.data(), fi.size(), cudaMemcpyHostToDevice));

  const size_t img = (size_t)d.max_batch * H * W * kTcC * 4;
  DLWP_TRY(plan_dev_alloc(p, (void**)&p->tc_img[0], img));
  DLWP_TRY(plan_dev_alloc(p, (void**)&p->tc_img[1], img));
  // Z' operand images: fno_mid_kernel writes the K slots of the retained columns; the slots that
  // pad K to a multiple of 16 stay zero from here on
  const size_t zimg = (size_t)d.max_batch * H * W / kTcTile * KZ * 128;
  DLWP_TRY(plan_dev_alloc(p, (void**)&p->tc_Z, zimg));
  DLWP_CUDA(cudaMemset(p->tc_Z, 0, zimg));
  const int L = d.lifting_channels, Pj = d.projection_channels;
  DLWP_TRY(plan_dev_alloc(p, (void**)&p->tc_liftW1, (size_t)L * 16 * 4));
  DLWP_TRY(plan_dev_alloc(p, (void**)&p->tc_liftW2, (size_t)32 * L * 4));
  DLWP_TRY(plan_dev_alloc(p, (void**)&p->tc_projW1, (size_t)Pj * 32 * 4));
  DLWP_TRY(plan_dev_alloc(p, (void**)&p->tc_projW2, (size_t)16 * Pj * 4));
  DLWP_TRY(plan_dev_alloc(p, (void**)&p->tc_projW1x, (size_t)Pj * 16 * 2));
  p->tc_skip.assign(d.n_layers, nullptr);
  for (int l = 0; l < d.n_layers; ++l) DLWP_TRY(plan_dev_alloc(p, (void**)&p->tc_skip[l], 4096));
  // H-DFT, mode mixing and inverse H-DFT on tcgen05 where the geometry allows it (DLWP_MID_FMA=1 keeps the fp32 FMA kernel: A/B runs)
  p->tc_mid2 = tc_mid2_samples(H, d.n_rows) != 0 && std::getenv("DLWP_MID_FMA") == nullptr;
  // DLWP_LIFT4=0 keeps the two-worker lifting kernel at every batch size (A/B runs)
  {
    const char* e = std::getenv("DLWP_LIFT4");
    p->tc_lift4 = e == nullptr || std::atoi(e) != 0;
    e = std::getenv("DLWP_LAYER_REV");
    p->tc_layer_rev = e == nullptr || std::atoi(e) != 0;
    e = std::getenv("DLWP_L2_HINTS");      // 0: the bulk copies of the activation images carry no L2 eviction priority (A/B runs)
    p->tc_l2hint = e == nullptr || std::atoi(e) != 0;
  }
  // the last layer and the projection as one kernel (DLWP_NO_FUSE_PROJ=1 keeps the two launches: A/B runs)
  p->tc_fuse_proj = d.n_layers >= 1 && tc_layerproj_fits(KZ) && std::getenv("DLWP_NO_FUSE_PROJ") == nullptr;
  if (p->tc_mid2) {
    const int R = d.n_rows;
    std::vector<double> fr((size_t)R * H), fi((size_t)R * H), gr((size_t)H * R), gi2((size_t)H * R);
    for (int j = 0; j < R; ++j)
      for (int h = 0; h < H; ++h) {
        long long a = ((long long)p->rows_in[j] * h) % H, b = ((long long)p->rows_out[j] * h) % H;
        const double af = -2.0 * M_PI * (double)a / (double)H, ag = 2.0 * M_PI * (double)b / (double)H;
        fr[(size_t)j * H + h] = std::cos(af);
        fi[(size_t)j * H + h] = std::sin(af);
        gr[(size_t)h * R + j] = std::cos(ag);
        gi2[(size_t)h * R + j] = std::sin(ag);
      }
    std::vector<uint8_t> fimg((size_t)2 * 64 * 2 * H * 2), gimg((size_t)2 * (2 * H < 64 ? 64 : 2 * H) * 32 * 2);
    tc_mid2_pack_bases(fr.data(), fi.data(), gr.data(), gi2.data(), H, R, fimg.data(), gimg.data());
    DLWP_TRY(plan_dev_alloc(p, (void**)&p->tc_midF, fimg.size()));
    DLWP_TRY(plan_dev_alloc(p, (void**)&p->tc_midG, gimg.size()));
    DLWP_CUDA(cudaMemcpy(p->tc_midF, fimg.data(), fimg.size(), cudaMemcpyHostToDevice));
    DLWP_CUDA(cudaMemcpy(p->tc_midG, gimg.data(), gimg.size(), cudaMemcpyHostToDevice));
    p->tc_midW.assign(d.n_layers, nullptr);
    for (int l = 0; l < d.n_layers; ++l) {      // [Kc][groups of 4 modes][hi 16 KB | lo 16 KB]; the modes that pad the last group stay zero
      const size_t wb = (size_t)Kc * ((R + 3) / 4) * 32768;
      DLWP_TRY(plan_dev_alloc(p, (void**)&p->tc_midW[l], wb));
      DLWP_CUDA(cudaMemset(p->tc_midW[l], 0, wb));
    }
  }
  p->tc = true;
  return DLWP_OK;
}

static int fetch(const dlwp_plan* p, const std::string& name, int64_t numel, std::vector<float>* out) {
  const float* dev;
  DLWP_TRY(plan_need(p, name, numel, &dev));
  out->resize((size_t)numel);
  DLWP_CUDA(cudaMemcpy(out->data(), dev, (size_t)numel * sizeof(float), cudaMemcpyDeviceToHost));
  return DLWP_OK;
}

static int upload(uint8_t* dst, const std::vector<uint8_t>& src) {
  DLWP_CUDA(cudaMemcpy(dst, src.data(), src.size(), cudaMemcpyHostToDevice));
  return DLWP_OK;
}

int plan_tc_commit(dlwp_plan* p, cudaStream_t st) {
  const dlwp_model_desc& d = p->d;
  const int L = d.lifting_channels, Pj = d.projection_channels, Cin = p->Cin, Cg = d.prognostic_channels;
  DLWP_CUDA(cudaStreamSynchronize(st));
  std::vector<float> w, b;
  std::vector<uint8_t> img;
  // lifting fcs.0: [L, Cin] weights with the bias as column Cin (the activations carry a 1 there)
  DLWP_TRY(fetch(p, "lifting.fcs.0.weight", (int64_t)L * Cin, &w));
  DLWP_TRY(fetch(p, "lifting.fcs.0.bias", L, &b));
  {
    std::vector<float> m((size_t)L * 16, 0.f);
    for (int h = 0; h < L; ++h) {
      for (int c = 0; c < Cin; ++c) m[(size_t)h * 16 + c] = w[(size_t)h * Cin + c];
      m[(size_t)h * 16 + Cin] = b[h];
    }
    img.assign((size_t)L * 16 * 4, 0);
    tc_pack_images(m.data(), L, 16, 16, L, 16, img.data());
    DLWP_TRY(upload(p->tc_liftW1, img));
  }
  // the MLP epilogues hand 2 * gelu(hidden) to the second GEMM (tc_fno.cu: gelu2x_8): its weights carry the 1/2
  DLWP_TRY(fetch(p, "lifting.fcs.1.weight", (int64_t)kTcC * L, &w));
  for (float& v : w) v *= 0.5f;
  img.assign((size_t)32 * L * 4, 0);
  tc_pack_images(w.data(), kTcC, L, L, 32, L, img.data());
  DLWP_TRY(upload(p->tc_liftW2, img));
  for (int l = 0; l < d.n_layers; ++l) {
    DLWP_TRY(fetch(p, "fno_blocks.fno_skips." + std::to_string(l) + ".weight", (int64_t)kTcC * kTcC, &w));
    img.assign(4096, 0);
    tc_pack_images(w.data(), 32, 32, 32, 32, 32, img.data());
    DLWP_TRY(upload(p->tc_skip[l], img));
  }
  DLWP_TRY(fetch(p, "projection.fcs.0.weight", (int64_t)Pj * kTcC, &w));
  img.assign((size_t)Pj * 32 * 4, 0);
  tc_pack_images(w.data(), Pj, 32, 32, Pj, 32, img.data());
  DLWP_TRY(upload(p->tc_projW1, img));
  DLWP_TRY(fetch(p, "projection.fcs.0.bias", Pj, &b));
  img.assign((size_t)Pj * 16 * 2, 0);
  tc_pack_bias_image(b.data(), Pj, Pj, img.data());
  DLWP_TRY(upload(p->tc_projW1x, img));
  DLWP_TRY(fetch(p, "projection.fcs.1.weight", (int64_t)Cg * Pj, &w));
  for (float& v : w) v *= 0.5f;
  img.assign((size_t)16 * Pj * 4, 0);
  tc_pack_images(w.data(), Cg, Pj, Pj, 16, Pj, img.data());
  DLWP_TRY(upload(p->tc_projW2, img));
  if (p->tc_mid2)      // the mode-major fp32 weights (packed by commit_fno on `st`) as tcgen05 operand images
    for (int l = 0; l < d.n_layers; ++l) DLWP_TRY(launch_tc_mid2_pack(p->wm[l], p->tc_midW[l], d.n_rows, d.n_cols, st));
  return DLWP_OK;
}

// One model evaluation on the tensor-core path.
int backbone_fno_tc(dlwp_plan* p, const StepIO& io, int B, cudaStream_t st) {
  const dlwp_model_desc& d = p->d;
  const int H = d.height, C = kTcC, R = d.n_rows, Kc = d.n_cols, N = d.n_layers;
  TcGeom g = p->tcg;
  g.ntiles = (int)((long long)B * g.HW / kTcTile);
  g.abf = p->tc_mid2 ? 1 : 0;
  g.Bn = B;
  g.l2hint = p->tc_l2hint ? 1 : 0;
  const float *b2, *cbias, *pb2;
  DLWP_TRY(plan_need(p, "lifting.fcs.1.bias", C, &b2));
  DLWP_TRY(plan_need(p, "fno_blocks.convs.bias", (int64_t)N * C, &cbias));
  DLWP_TRY(plan_need(p, "projection.fcs.1.bias", d.prognostic_channels, &pb2));

  TcLiftArgs la{};
  la.g = g;
  la.ngrp = io.ngrp;
  for (int i = 0; i < io.ngrp; ++i) la.grp[i] = io.grp[i];
  la.Cin = p->Cin;
  la.hid = d.lifting_channels;
  la.workers4 = p->tc_lift4 ? 1 : 0;
  la.W1img = p->tc_liftW1;
  la.W2img = p->tc_liftW2;
  la.b2 = b2;
  la.Fimg = p->tc_F;
  la.y_img = p->tc_img[0];
  la.Aout = reinterpret_cast<float*>(p->specA);
  DLWP_TRY(launch_tc_lift(la, st));

  int cur = 0;
  for (int l = 0; l < N; ++l) {
    // H-DFT on the retained columns, per-mode channel mixing, inverse H-DFT -> Z' operand images
    if (p->tc_mid2) {
      TcMid2Args ma{};
      ma.A = reinterpret_cast<const uint8_t*>(p->specA); ma.Wimg = p->tc_midW[l]; ma.Fimg = p->tc_midF; ma.Gimg = p->tc_midG;
      ma.Zimg = p->tc_Z; ma.B = B; ma.H = H; ma.R = R; ma.Kc = Kc; ma.RT = g.RT; ma.KZ = g.KZ;
      ma.l2hint = g.l2hint;
      // its spectrum tiles were written by the lifting kernel (upwards) or by layer l - 1: the sample groups go the other way
      ma.rev = p->tc_layer_rev && (l == 0 || !((N - l) & 1)) ? 1 : 0;
      DLWP_TRY(launch_tc_mid2(ma, st));
    } else {
    TcMidArgs ma{};
    ma.A = reinterpret_cast<const float*>(p->specA); ma.Wm = p->wm[l]; ma.Fh = p->Fh; ma.Gh = p->Gh; ma.Zimg = p->tc_Z;
    ma.B = B; ma.H = H; ma.R = R; ma.Kc = Kc; ma.RT = g.RT; ma.KZ = g.KZ;
    DLWP_TRY(launch_tc_mid(ma, st));
    }

    TcLayerArgs ya{};
    ya.g = g;
    ya.x_img = p->tc_img[cur];
    ya.y_img = p->tc_img[cur ^ 1];
    ya.Zimg = p->tc_Z;
    ya.Aout = l < N - 1 ? reinterpret_cast<float*>(p->specA) : nullptr;
    ya.Gimg = p->tc_G;
    ya.Fimg = p->tc_F;
    ya.Wsimg = p->tc_skip[l];
    ya.bias = cbias + (size_t)l * C;
    ya.gelu = l < N - 1 ? 1 : 0;
    // Consecutive kernels walk the tile list in opposite directions: the lifting kernel and the fused tail upwards, the layer
    // kernels alternately down and up, so that a kernel starts on the tiles its producer wrote last -- the ones the L2 still
    // holds (an activation image set is larger than what survives the kernel in between).  DLWP_LAYER_REV=0: all upwards.
    ya.g.rev = p->tc_layer_rev && ((N - 1 - l) & 1) ? 1 : 0;
    if (l == N - 1 && p->tc_fuse_proj) {      // last layer + projection in one kernel: the layer's output tile never leaves the SM
      TcLayerProjArgs fa{};
      fa.layer = ya;
      fa.proj.g = g;
      fa.proj.hid = d.projection_channels;
      fa.proj.Cg = d.prognostic_channels;
      fa.proj.W1img = p->tc_projW1;
      fa.proj.W1ximg = p->tc_projW1x;
      fa.proj.W2img = p->tc_projW2;
      fa.proj.b2 = pb2;
      fa.proj.out = io.out;
      fa.proj.out_bstride = io.out_bstride;
      fa.proj.last = io.last;
      fa.proj.last_bstride = io.last_bstride;
      return launch_tc_layerproj(fa, st);
    }
    DLWP_TRY(launch_tc_layer(ya, st));
    cur ^= 1;
  }

  TcProjArgs pa{};
  pa.g = g;
  pa.x_img = p->tc_img[cur];
  pa.hid = d.projection_channels;
  pa.Cg = d.prognostic_channels;
  pa.W1img = p->tc_projW1;
  pa.W1ximg = p->tc_projW1x;
  pa.W2img = p->tc_projW2;
  pa.b2 = pb2;
  pa.out = io.out;
  pa.out_bstride = io.out_bstride;
  pa.last = io.last;
  pa.last_bstride = io.last_bstride;
  DLWP_TRY(launch_tc_proj(pa, st));
  return DLWP_OK;
}

}  // namespace dlwp
