// Plan-side glue of the bf16 tensor-core path at hidden width 128 (tc_wide.cuh): eligibility,
// W-DFT operand tiles, weight images and the per-step kernel sequence (reference: neuralop
// FNO.forward behind fno.py:103, at BASELINE config 5's widths).
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "plan.cuh"

namespace dlwp {

int plan_tcw_build(dlwp_plan* p) {
  const dlwp_model_desc& d = p->d;
  p->tcw = false;
  DLWP_REQUIRE(d.kind == DLWP_MODEL_FNO && p->C == kWideC && d.lifting_channels == kWideHid &&
                   d.projection_channels == kWideHid,
               "DLWP_PREC_BF16 is built for FNO2D with hidden_channels 128 and 256-wide lifting / projection "
               "(got hidden %d, lifting %d, projection %d)",
               p->C, d.lifting_channels, d.projection_channels);
  DLWP_REQUIRE(p->Cin >= 1 && p->Cin <= 14 && d.prognostic_channels <= 16,
               "DLWP_PREC_BF16: %d input / %d output channels exceed 14 / 16", p->Cin, d.prognostic_channels);
  const int H = d.height, W = d.width, Kc = d.n_cols;
  const int KZ = (2 * Kc + 15) & ~15;
  DLWP_REQUIRE(Kc >= 1 && KZ <= 80 && d.n_rows >= 1 && d.n_rows <= 64,
               "DLWP_PREC_BF16: %d x %d retained modes (at most 64 rows, 40 columns)", d.n_rows, Kc);
  const int TW = (W + 127) / 128;
  p->tcw_TW = TW;
  p->tcw_KZ = KZ;

  double s = 1.0, t = 1.0;
  const double hw = (double)H * (double)W;
  if (d.fft_norm == DLWP_FFT_NORM_BACKWARD) t = 1.0 / hw;
  else if (d.fft_norm == DLWP_FFT_NORM_FORWARD) s = 1.0 / hw;
  else s = t = 1.0 / std::sqrt(hw);
  // per tile of 128 slots (slot -> w = tile * 128 + slot; slots past W are zero rows / columns):
  //   G[slot, k] : y[w] = sum_kx Re(Gw[w, kx] Z[kx]),  Gw = t c_kx exp(+2 pi i kx w / W)  (plan.cu: build_bases)
  //   F[k, slot] : A[kx] = sum_w Fw[kx, w] x[w],       Fw = s exp(-2 pi i kx w / W),      k = 2 kx + (re | im)
  const size_t gbytes = (size_t)128 * KZ * 2, fbytes = (size_t)128 * 128 * 2;
  std::vector<uint8_t> gi(TW * gbytes), fi(TW * fbytes);
  std::vector<double> G((size_t)128 * KZ), F((size_t)128 * 128);
  for (int tile = 0; tile < TW; ++tile) {
    std::fill(G.begin(), G.end(), 0.0);
    std::fill(F.begin(), F.end(), 0.0);
    for (int slot = 0; slot < 128; ++slot) {
      const int w = tile * 128 + slot;
      if (w >= W) continue;
      for (int kx = 0; kx < Kc; ++kx) {
        const long long r = ((long long)kx * w) % W;
        const double ang = 2.0 * M_PI * (double)r / (double)W;
        const bool self_conj = (kx == 0) || (W % 2 == 0 && kx == W / 2);
        const double ck = t * (self_conj ? 1.0 : 2.0);
        G[(size_t)slot * KZ + 2 * kx] = ck * std::cos(ang);
        G[(size_t)slot * KZ + 2 * kx + 1] = -ck * std::sin(ang);
        F[(size_t)(2 * kx) * 128 + slot] = s * std::cos(ang);
        F[(size_t)(2 * kx + 1) * 128 + slot] = -s * std::sin(ang);
      }
    }
    tcw_pack_image_f64(G.data(), 128, KZ, KZ, 128, KZ, gi.data() + tile * gbytes);
    tcw_pack_image_f64(F.data(), 128, 128, 128, 128, 128, fi.data() + tile * fbytes);
  }
  DLWP_TRY(plan_dev_alloc(p, (void**)&p->tcw_G, gi.size()));
  DLWP_TRY(plan_dev_alloc(p, (void**)&p->tcw_F, fi.size()));
  DLWP_CUDA(cudaMemcpy(p->tcw_G, gi.data(), gi.size(), cudaMemcpyHostToDevice));
  DLWP_CUDA(cudaMemcpy(p->tcw_F, fi.data(), fi.size(), cudaMemcpyHostToDevice));

  // H-DFT tiles, one per block of 64 grid rows; complex products as real GEMMs on (re | im)-stacked indices:
  //   Fs[(r, re|im), (h, re|im)] : X = Fh A,  Fh[r, h] = exp(-2 pi i rows_in[r] h / H)
  //   Gs[(h, re|im), (r, re|im)] : Z = Gh Y,  Gh[h, r] = exp(+2 pi i rows_out[r] h / H)
  const int HB = (H + 63) / 64, R = d.n_rows;
  p->tcw_HB = HB;
  std::vector<uint8_t> fsi((size_t)HB * fbytes), gsi((size_t)HB * fbytes);
  std::vector<double> Fs((size_t)128 * 128), Gs((size_t)128 * 128);
  for (int hb = 0; hb < HB; ++hb) {
    std::fill(Fs.begin(), Fs.end(), 0.0);
    std::fill(Gs.begin(), Gs.end(), 0.0);
    for (int hh = 0; hh < 64; ++hh) {
      const int h = hb * 64 + hh;
      if (h >= H) continue;
      for (int r = 0; r < R; ++r) {
        const double af = -2.0 * M_PI * (double)(((long long)p->rows_in[r] * h) % H) / (double)H;
        const double ag = 2.0 * M_PI * (double)(((long long)p->rows_out[r] * h) % H) / (double)H;
        const double fr = std::cos(af), fi2 = std::sin(af), gr = std::cos(ag), gi2 = std::sin(ag);
        Fs[(size_t)(2 * r) * 128 + 2 * hh] = fr;       Fs[(size_t)(2 * r) * 128 + 2 * hh + 1] = -fi2;
        Fs[(size_t)(2 * r + 1) * 128 + 2 * hh] = fi2;  Fs[(size_t)(2 * r + 1) * 128 + 2 * hh + 1] = fr;
        Gs[(size_t)(2 * hh) * 128 + 2 * r] = gr;       Gs[(size_t)(2 * hh) * 128 + 2 * r + 1] = -gi2;
        Gs[(size_t)(2 * hh + 1) * 128 + 2 * r] = gi2;  Gs[(size_t)(2 * hh + 1) * 128 + 2 * r + 1] = gr;
      }
    }
    tcw_pack_image_f64(Fs.data(), 128, 128, 128, 128, 128, fsi.data() + hb * fbytes);
    tcw_pack_image_f64(Gs.data(), 128, 128, 128, 128, 128, gsi.data() + hb * fbytes);
  }
  DLWP_TRY(plan_dev_alloc(p, (void**)&p->tcw_Fs, fsi.size()));
  DLWP_TRY(plan_dev_alloc(p, (void**)&p->tcw_Gs, gsi.size()));
  DLWP_CUDA(cudaMemcpy(p->tcw_Fs, fsi.data(), fsi.size(), cudaMemcpyHostToDevice));
  DLWP_CUDA(cudaMemcpy(p->tcw_Gs, gsi.data(), gsi.size(), cudaMemcpyHostToDevice));
  // spectrum images between the W- and H-DFTs (rows past H stay zero) and after the mode mixing (rows past 2 R stay zero)
  const size_t sab = (size_t)d.max_batch * Kc * HB * fbytes, ysb = (size_t)d.max_batch * Kc * fbytes;
  DLWP_TRY(plan_dev_alloc(p, (void**)&p->tcw_sa, sab));
  DLWP_TRY(plan_dev_alloc(p, (void**)&p->tcw_ys, ysb));
  DLWP_CUDA(cudaMemset(p->tcw_sa, 0, sab));
  DLWP_CUDA(cudaMemset(p->tcw_ys, 0, ysb));

  const size_t nt = (size_t)d.max_batch * H * TW;
  DLWP_TRY(plan_dev_alloc(p, (void**)&p->tcw_in, nt * 128 * 32 * 2));
  DLWP_TRY(plan_dev_alloc(p, (void**)&p->tcw_x[0], nt * 128 * kWideC * 2));
  DLWP_TRY(plan_dev_alloc(p, (void**)&p->tcw_x[1], nt * 128 * kWideC * 2));
  // per grid row [KZ rows (kx, re | im)] x [128 channels]; the rows past 2 Kc stay zero from here on
  const size_t zb = (size_t)d.max_batch * H * KZ * 256;
  DLWP_TRY(plan_dev_alloc(p, (void**)&p->tcw_z, zb));
  DLWP_CUDA(cudaMemset(p->tcw_z, 0, zb));

  DLWP_TRY(plan_dev_alloc(p, (void**)&p->tcw_W1, (size_t)kWideHid * 32 * 2));
  DLWP_TRY(plan_dev_alloc(p, (void**)&p->tcw_W2, (size_t)kWideC * kWideHid * 2));
  DLWP_TRY(plan_dev_alloc(p, (void**)&p->tcw_b2, (size_t)kWideC * 32));
  DLWP_TRY(plan_dev_alloc(p, (void**)&p->tcw_P1, (size_t)kWideHid * kWideC * 2));
  DLWP_TRY(plan_dev_alloc(p, (void**)&p->tcw_pb1, (size_t)kWideHid * 32));
  DLWP_TRY(plan_dev_alloc(p, (void**)&p->tcw_P2, (size_t)16 * kWideHid * 2));
  DLWP_TRY(plan_dev_alloc(p, (void**)&p->tcw_pb2, (size_t)16 * 32));
  p->tcw_skip.assign(d.n_layers, nullptr);
  p->tcw_cbias.assign(d.n_layers, nullptr);
  for (int l = 0; l < d.n_layers; ++l) {
    DLWP_TRY(plan_dev_alloc(p, (void**)&p->tcw_skip[l], (size_t)kWideC * kWideC * 2));
    DLWP_TRY(plan_dev_alloc(p, (void**)&p->tcw_cbias[l], (size_t)kWideC * 32));
  }
  p->tcw = true;
  return DLWP_OK;
}

static int fetch(const dlwp_plan* p, const std::string& name, int64_t numel, std::vector<float>* out) {
  const float* dev;
  DLWP_TRY(plan_need(p, name, numel, &dev));
  out->resize((size_t)numel);
  DLWP_CUDA(cudaMemcpy(out->data(), dev, (size_t)numel * sizeof(float), cudaMemcpyDeviceToHost));
  return DLWP_OK;
}

// `scale` * w [rows, cols] -> [R x K] image on the device
static int put_weight(uint8_t* dst, const std::vector<float>& w, float scale, int rows, int cols, int R, int K) {
  std::vector<float> m(w.begin(), w.begin() + (size_t)rows * cols);
  for (float& v : m) v *= scale;
  std::vector<uint8_t> img((size_t)R * K * 2);
  tcw_pack_image(m.data(), rows, cols, cols, R, K, img.data());
  DLWP_CUDA(cudaMemcpy(dst, img.data(), img.size(), cudaMemcpyHostToDevice));
  return DLWP_OK;
}

// bf16 rounding of v as float (round to nearest even), for the hi / lo split of the biases
static float bf16_round(float v) {
  uint32_t u;
  std::memcpy(&u, &v, 4);
  u += 0x7FFFu + ((u >> 16) & 1u);
  u &= 0xFFFF0000u;
  float r;
  std::memcpy(&r, &u, 4);
  return r;
}

// bias [n] -> [R x 16] image: column 0 = hi part, column 1 = lo part (both meet a 1.0 of the ones operand)
static int put_bias(uint8_t* dst, const std::vector<float>& b, int n, int R) {
  std::vector<float> m((size_t)R * 16, 0.f);
  for (int i = 0; i < n; ++i) {
    const float hi = bf16_round(b[i]);
    m[(size_t)i * 16] = hi;
    m[(size_t)i * 16 + 1] = b[i] - hi;
  }
  std::vector<uint8_t> img((size_t)R * 32);
  tcw_pack_image(m.data(), R, 16, 16, R, 16, img.data());
  DLWP_CUDA(cudaMemcpy(dst, img.data(), img.size(), cudaMemcpyHostToDevice));
  return DLWP_OK;
}

int plan_tcw_commit(dlwp_plan* p, cudaStream_t st) {
  const dlwp_model_desc& d = p->d;
  const int L = kWideHid, C = kWideC, Cin = p->Cin, Cg = d.prognostic_channels, N = d.n_layers;
  DLWP_CUDA(cudaStreamSynchronize(st));
  std::vector<float> w, b;
  // lifting fcs.0 against the [hi | ones | lo] input image: [L x 32]
  DLWP_TRY(fetch(p, "lifting.fcs.0.weight", (int64_t)L * Cin, &w));
  DLWP_TRY(fetch(p, "lifting.fcs.0.bias", L, &b));
  {
    std::vector<float> m((size_t)L * 32, 0.f);
    for (int h = 0; h < L; ++h) {
      for (int c = 0; c < Cin; ++c) m[(size_t)h * 32 + c] = m[(size_t)h * 32 + 16 + c] = w[(size_t)h * Cin + c];
      const float hi = bf16_round(b[h]);
      m[(size_t)h * 32 + Cin] = hi;
      m[(size_t)h * 32 + 15] = b[h] - hi;
    }
    DLWP_TRY(put_weight(p->tcw_W1, m, 1.f, L, 32, L, 32));
  }
  // the GELU epilogues store 2 * gelu: every consumer of such an image carries the 1/2
  DLWP_TRY(fetch(p, "lifting.fcs.1.weight", (int64_t)C * L, &w));
  DLWP_TRY(put_weight(p->tcw_W2, w, 0.5f, C, L, C, L));
  DLWP_TRY(fetch(p, "lifting.fcs.1.bias", C, &b));
  DLWP_TRY(put_bias(p->tcw_b2, b, C, C));
  DLWP_TRY(fetch(p, "fno_blocks.convs.bias", (int64_t)N * C, &b));
  for (int l = 0; l < N; ++l) {
    DLWP_TRY(fetch(p, "fno_blocks.fno_skips." + std::to_string(l) + ".weight", (int64_t)C * C, &w));
    DLWP_TRY(put_weight(p->tcw_skip[l], w, l >= 1 ? 0.5f : 1.f, C, C, C, C));
    std::vector<float> bl(b.begin() + (size_t)l * C, b.begin() + (size_t)(l + 1) * C);
    DLWP_TRY(put_bias(p->tcw_cbias[l], bl, C, C));
  }
  // bf16 copy of the mode-major spectral weights (plan.cu: commit_fno packed them before this runs):
  // with a few samples per GPU the mode mixing is bound by reading them
  {
    const long long n = (long long)d.n_rows * d.n_cols * C * C;
    if (p->tcw_wm16.empty()) {
      p->tcw_wm16.assign(N, nullptr);
      for (int l = 0; l < N; ++l) DLWP_TRY(plan_dev_alloc(p, (void**)&p->tcw_wm16[l], (size_t)n * 4));
    }
    // few samples on tcgen05 (DLWP_MIX_FMA=1 keeps the CUDA-core kernel: A/B runs): the same bytes as operand images
    p->tcw_mix_tc = mix_few_tc_covers(d.n_rows * d.n_cols, 1, C, d.n_cols) && std::getenv("DLWP_MIX_FMA") == nullptr;
    for (int l = 0; l < N; ++l) {
      if (p->tcw_mix_tc) DLWP_TRY(launch_mix_few_pack(p->wm[l], reinterpret_cast<uint8_t*>(p->tcw_wm16[l]), d.n_rows * d.n_cols, st));
      else DLWP_TRY(launch_narrow_bf16(p->wm[l], p->tcw_wm16[l], n, st));
    }
    DLWP_CUDA(cudaStreamSynchronize(st));
  }
  // the last block applies no GELU, so the projection reads a plain image
  DLWP_TRY(fetch(p, "projection.fcs.0.weight", (int64_t)L * C, &w));
  DLWP_TRY(put_weight(p->tcw_P1, w, 1.f, L, C, L, C));
  DLWP_TRY(fetch(p, "projection.fcs.0.bias", L, &b));
  DLWP_TRY(put_bias(p->tcw_pb1, b, L, L));
  DLWP_TRY(fetch(p, "projection.fcs.1.weight", (int64_t)Cg * L, &w));
  DLWP_TRY(put_weight(p->tcw_P2, w, 0.5f, Cg, L, 16, L));
  DLWP_TRY(fetch(p, "projection.fcs.1.bias", Cg, &b));
  DLWP_TRY(put_bias(p->tcw_pb2, b, Cg, 16));
  return DLWP_OK;
}

// One model evaluation on the bf16 tensor-core path.
int backbone_fno_tcw(dlwp_plan* p, const StepIO& io, int B, cudaStream_t st) {
  const dlwp_model_desc& d = p->d;
  const int H = d.height, W = d.width, N = d.n_layers, TW = p->tcw_TW;
  TcwArgs base{};
  base.nitems = B * H * TW;
  base.TW = TW; base.H = H; base.W = W; base.Kc = d.n_cols; base.KZ = p->tcw_KZ;
  base.B = B; base.R = d.n_rows; base.HB = p->tcw_HB;
  base.scale = 1.f;

  WideGatherArgs ga{};
  ga.ngrp = io.ngrp;
  for (int i = 0; i < io.ngrp; ++i) ga.grp[i] = io.grp[i];
  ga.Cin = p->Cin; ga.ntiles = base.nitems; ga.TW = TW; ga.H = H; ga.W = W; ga.in_img = p->tcw_in;
  DLWP_TRY(launch_wide_gather(ga, st));

  TcwArgs a = base;
  a.a_img = p->tcw_in; a.b_res = p->tcw_W1; a.w2_img = p->tcw_W2; a.bias2_img = p->tcw_b2; a.out_img = p->tcw_x[0];
  DLWP_TRY(launch_tcw_mlp(0, a, st));

  int cur = 0;
  for (int l = 0; l < N; ++l) {
    // forward W-DFT -> spectrum images; forward H-DFT -> X; mode mixing -> Y images; inverse H-DFT -> Z images
    a = base;
    a.nitems = B * H;
    a.a_img = p->tcw_x[cur]; a.g_img = p->tcw_F; a.out_img = p->tcw_sa; a.scale = l >= 1 ? 0.5f : 1.f;
    DLWP_TRY(launch_tcw(TCW_FWD, a, st));
    a = base;
    a.nitems = B * d.n_cols;
    a.a_img = p->tcw_sa; a.g_img = p->tcw_Fs; a.spec = p->specX;
    DLWP_TRY(launch_tcw(TCW_HFWD, a, st));
    if (p->tcw_mix_tc && B <= 8)
      DLWP_TRY(launch_mix_few_tc(p->specX, reinterpret_cast<const uint8_t*>(p->tcw_wm16[l]), p->tcw_ys, d.n_rows * d.n_cols, B, d.n_cols, st));
    else
      DLWP_TRY(launch_mix_dense(p->specX, p->wm[l], p->specY, d.n_rows * d.n_cols, B, kWideC, kWideC, st,
                                MixWideOpts{p->tcw_ys, d.n_cols, (B <= 4 && !p->tcw_mix_tc) ? p->tcw_wm16[l] : nullptr}));
    a = base;
    a.nitems = B * d.n_cols * p->tcw_HB;
    a.a_img = p->tcw_ys; a.g_img = p->tcw_Gs; a.out_img = p->tcw_z;
    DLWP_TRY(launch_tcw(TCW_HINV, a, st));
    a = base;
    a.a_img = p->tcw_x[cur]; a.b_res = p->tcw_skip[l]; a.bias_img = p->tcw_cbias[l]; a.g_img = p->tcw_G;
    a.z_img = p->tcw_z; a.out_img = p->tcw_x[cur ^ 1]; a.gelu = l < N - 1 ? 1 : 0;
    DLWP_TRY(launch_tcw(TCW_LAYER, a, st));
    cur ^= 1;
  }

  a = base;
  a.a_img = p->tcw_x[cur]; a.b_res = p->tcw_P1; a.bias_img = p->tcw_pb1; a.w2_img = p->tcw_P2; a.bias2_img = p->tcw_pb2;
  a.out = io.out; a.out_bstride = io.out_bstride; a.last = io.last; a.last_bstride = io.last_bstride;
  a.Cg = d.prognostic_channels;
  DLWP_TRY(launch_tcw_mlp(1, a, st));
  return DLWP_OK;
}

}  // namespace dlwp
