// Plan-side glue of the tensor-core AFNO kernels: eligibility, weight operand images
// (reference: Block.forward / Mlp.forward, fourcastnet.py:41-57,176-193).
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "plan.cuh"

namespace dlwp {

static std::string bname(int i, const char* tail) { return "blocks." + std::to_string(i) + tail; }

// The kernels are written for the reference configuration (configs/model/fourcastnet.yaml:
// embed_dim 64, mlp_ratio 4) on token counts that fill 128-token tiles.  Anything else keeps the
// fp32 kernel set of kernels_fp32.cu.
int plan_atc_build(dlwp_plan* p) {
  const dlwp_model_desc& d = p->d;
  p->atc = false;
  if (d.kind != DLWP_MODEL_AFNO || d.embed_dim != kAfnoE || d.mlp_hidden != kAfnoHid) return DLWP_OK;
  if ((p->gh * p->gw) % kTcTile != 0) return DLWP_OK;
  p->atc_W1.assign(d.depth, nullptr);
  p->atc_W2.assign(d.depth, nullptr);
  p->atc_b1h.assign(d.depth, std::vector<float>());
  p->atc_b2h.assign(d.depth, std::vector<float>());
  for (int i = 0; i < d.depth; ++i) {
    DLWP_TRY(plan_dev_alloc(p, (void**)&p->atc_W1[i], (size_t)kAfnoHid * kAfnoE * 4));
    DLWP_TRY(plan_dev_alloc(p, (void**)&p->atc_W2[i], (size_t)kAfnoE * kAfnoHid * 4));
  }
  p->atc = true;
  const int bs = d.embed_dim / d.num_blocks;
  p->atc_spec = d.afno_filter == DLWP_AFNO_FILTER_AFNO2D && p->gw <= 64 && p->gh <= 32 && d.n_rows >= 1 && d.n_rows <= 32 && d.n_cols >= 1 && d.n_cols <= 36 &&
                d.hidden_size_factor == 1 && (bs == 8 || bs == 16) &&
                p->gw % 2 == 0 && (d.n_rows * p->gh) % 2 == 0;   // bulk copies move multiples of 16 bytes
  // W-DFTs on tcgen05: whole grid rows per 128-token tile, all retained (column, re | im) pairs of a tile in one M = 128 UMMA
  const int W = p->gw, Kc = d.n_cols;
  if (p->atc_spec && W <= kTcTile && kTcTile % W == 0 && (kTcTile / W) * 2 * Kc <= 128 && p->gh % (kTcTile / W) == 0) {
    const int RT = kTcTile / W, KZ = (RT * 2 * Kc + 15) & ~15;
    double sc = 1.0, tc = 1.0;
    const double hw = (double)p->gh * (double)W;
    if (d.fft_norm == DLWP_FFT_NORM_BACKWARD) tc = 1.0 / hw;
    else if (d.fft_norm == DLWP_FFT_NORM_FORWARD) sc = 1.0 / hw;
    else sc = tc = 1.0 / std::sqrt(hw);
    // F''[(h2, kx, re | im), (h2', w)] = [h2 == h2'] s exp(-2 pi i kx w / W);  G'[(h2, w), (h2', kx, re | im)]: y[w] =
    // sum_kx Re(Gw[w, kx] Z[kx]), Gw = t c_kx exp(+2 pi i kx w / W), c = 1 for the self-conjugate columns, else 2
    std::vector<double> F((size_t)128 * 128, 0.0), G((size_t)128 * KZ, 0.0);
    for (int h2 = 0; h2 < RT; ++h2)
      for (int w = 0; w < W; ++w)
        for (int kx = 0; kx < Kc; ++kx) {
          const long long r = ((long long)kx * w) % W;
          const double ang = 2.0 * M_PI * (double)r / (double)W;
          const bool self_conj = (kx == 0) || (W % 2 == 0 && kx == W / 2);
          const double ck = tc * (self_conj ? 1.0 : 2.0);
          const int row = (h2 * Kc + kx) * 2, px = h2 * W + w;
          F[(size_t)row * 128 + px] = sc * std::cos(ang);
          F[(size_t)(row + 1) * 128 + px] = -sc * std::sin(ang);
          G[(size_t)px * KZ + row] = ck * std::cos(ang);
          G[(size_t)px * KZ + row + 1] = -ck * std::sin(ang);
        }
    std::vector<uint8_t> fi((size_t)2 * 128 * 128 * 2), gi((size_t)2 * 128 * KZ * 2);
    tc_pack_images_f64(F.data(), 128, 128, 128, 128, 128, fi.data());
    tc_pack_images_f64(G.data(), 128, KZ, KZ, 128, KZ, gi.data());
    DLWP_TRY(plan_dev_alloc(p, (void**)&p->atc_F, fi.size()));
    DLWP_TRY(plan_dev_alloc(p, (void**)&p->atc_G, gi.size()));
    DLWP_CUDA(cudaMemcpy(p->atc_F, fi.data(), fi.size(), cudaMemcpyHostToDevice));
    DLWP_CUDA(cudaMemcpy(p->atc_G, gi.data(), gi.size(), cudaMemcpyHostToDevice));
    p->atc_RT = RT;
    p->atc_KZ = KZ;
    p->atc_wtc = true;
    // H-DFT, block MLP and inverse H-DFT on tcgen05 where the geometry allows it (DLWP_MID_FMA=1 keeps the fp32 FMA kernel: A/B runs)
    if (afno_mid_tc_covers(p->gh, d.n_rows, bs) && std::getenv("DLWP_MID_FMA") == nullptr) {
      const int H = p->gh, R = d.n_rows;
      std::vector<double> fr((size_t)R * H), fi((size_t)R * H), gr((size_t)H * R), gi2((size_t)H * R);
      for (int j = 0; j < R; ++j)
        for (int h = 0; h < H; ++h) {
          long long r1 = ((long long)p->rows_in[j] * h) % H, r2 = ((long long)p->rows_out[j] * h) % H;
          if (r1 < 0) r1 += H;
          if (r2 < 0) r2 += H;
          const double a1 = -2.0 * M_PI * (double)r1 / (double)H, a2 = 2.0 * M_PI * (double)r2 / (double)H;
          fr[(size_t)j * H + h] = std::cos(a1);
          fi[(size_t)j * H + h] = std::sin(a1);
          gr[(size_t)h * R + j] = std::cos(a2);
          gi2[(size_t)h * R + j] = std::sin(a2);
        }
      std::vector<uint8_t> fimg(16384), gimg(16384);
      afno_mid_tc_pack_bases(fr.data(), fi.data(), gr.data(), gi2.data(), H, R, fimg.data(), gimg.data());
      DLWP_TRY(plan_dev_alloc(p, (void**)&p->atc_midF, fimg.size()));
      DLWP_TRY(plan_dev_alloc(p, (void**)&p->atc_midG, gimg.size()));
      DLWP_CUDA(cudaMemcpy(p->atc_midF, fimg.data(), fimg.size(), cudaMemcpyHostToDevice));
      DLWP_CUDA(cudaMemcpy(p->atc_midG, gimg.data(), gimg.size(), cudaMemcpyHostToDevice));
      p->atc_midW1.assign(d.depth, nullptr);
      p->atc_midW2.assign(d.depth, nullptr);
      p->atc_midb.assign(d.depth, nullptr);
      const size_t wbytes = (size_t)d.num_blocks * 2 * (2 * bs) * (2 * bs) * 2;
      for (int i = 0; i < d.depth; ++i) {
        DLWP_TRY(plan_dev_alloc(p, (void**)&p->atc_midW1[i], wbytes));
        DLWP_TRY(plan_dev_alloc(p, (void**)&p->atc_midW2[i], wbytes));
        DLWP_TRY(plan_dev_alloc(p, (void**)&p->atc_midb[i], (size_t)4 * kAfnoE * sizeof(float)));
      }
      // the K rows that pad an image to a multiple of 16 are never written: they stay zero
      const size_t zbytes = (size_t)d.max_batch * p->gh / RT * 2 * KZ * 128;
      DLWP_TRY(plan_dev_alloc(p, (void**)&p->atc_Zimg, zbytes));
      DLWP_CUDA(cudaMemset(p->atc_Zimg, 0, zbytes));
      p->atc_mid2 = true;
    }
  }
  return DLWP_OK;
}

static int fetch(const dlwp_plan* p, const std::string& name, int64_t numel, std::vector<float>* out) {
  const float* dev;
  DLWP_TRY(plan_need(p, name, numel, &dev));
  out->resize((size_t)numel);
  DLWP_CUDA(cudaMemcpy(out->data(), dev, (size_t)numel * sizeof(float), cudaMemcpyDeviceToHost));
  return DLWP_OK;
}

int plan_atc_commit(dlwp_plan* p, cudaStream_t st) {
  const dlwp_model_desc& d = p->d;
  DLWP_CUDA(cudaStreamSynchronize(st));
  std::vector<float> w, b, g, be;
  std::vector<uint8_t> img((size_t)kAfnoHid * kAfnoE * 4);
  for (int i = 0; i < d.depth; ++i) {
    // LayerNorm(x) W1^T + b1 = xhat (W1 diag(gamma))^T + (b1 + W1 beta): the kernel only normalises
    DLWP_TRY(fetch(p, bname(i, ".norm2.weight"), kAfnoE, &g));
    DLWP_TRY(fetch(p, bname(i, ".norm2.bias"), kAfnoE, &be));
    DLWP_TRY(fetch(p, bname(i, ".mlp.fc1.weight"), (int64_t)kAfnoHid * kAfnoE, &w));
    DLWP_TRY(fetch(p, bname(i, ".mlp.fc1.bias"), kAfnoHid, &b));
    std::vector<float> b1(kAfnoHid);
    for (int h = 0; h < kAfnoHid; ++h) {
      double acc = b[h];
      for (int c = 0; c < kAfnoE; ++c) {
        acc += (double)w[(size_t)h * kAfnoE + c] * (double)be[c];
        w[(size_t)h * kAfnoE + c] = (float)((double)w[(size_t)h * kAfnoE + c] * (double)g[c]);
      }
      b1[h] = (float)acc;
    }
    tc_pack_images(w.data(), kAfnoHid, kAfnoE, kAfnoE, kAfnoHid, kAfnoE, img.data());
    DLWP_CUDA(cudaMemcpy(p->atc_W1[i], img.data(), img.size(), cudaMemcpyHostToDevice));
    p->atc_b1h[i] = b1;
    DLWP_TRY(fetch(p, bname(i, ".mlp.fc2.bias"), kAfnoE, &p->atc_b2h[i]));
    // the GELU epilogue hands 2 * gelu(hidden) to the second GEMM: its weights carry the 1/2
    DLWP_TRY(fetch(p, bname(i, ".mlp.fc2.weight"), (int64_t)kAfnoE * kAfnoHid, &w));
    for (float& v : w) v *= 0.5f;
    tc_pack_images(w.data(), kAfnoE, kAfnoHid, kAfnoHid, kAfnoE, kAfnoHid, img.data());
    DLWP_CUDA(cudaMemcpy(p->atc_W2[i], img.data(), img.size(), cudaMemcpyHostToDevice));
    if (p->atc_mid2) {      // the block MLP of the filter as (re|im)-stacked operand images (fourcastnet.py:73-76,96-119)
      const int nb = d.num_blocks, bs = kAfnoE / nb;
      std::vector<uint8_t> wimg((size_t)nb * 2 * (2 * bs) * (2 * bs) * 2);
      std::vector<float> bias((size_t)4 * kAfnoE);
      DLWP_TRY(fetch(p, bname(i, ".filter.w1"), (int64_t)2 * nb * bs * bs, &w));
      DLWP_TRY(fetch(p, bname(i, ".filter.b1"), (int64_t)2 * nb * bs, &b));
      afno_mid_tc_pack_layer(w.data(), b.data(), nb, bs, wimg.data(), bias.data());
      DLWP_CUDA(cudaMemcpy(p->atc_midW1[i], wimg.data(), wimg.size(), cudaMemcpyHostToDevice));
      DLWP_TRY(fetch(p, bname(i, ".filter.w2"), (int64_t)2 * nb * bs * bs, &w));
      DLWP_TRY(fetch(p, bname(i, ".filter.b2"), (int64_t)2 * nb * bs, &b));
      afno_mid_tc_pack_layer(w.data(), b.data(), nb, bs, wimg.data(), bias.data() + 2 * kAfnoE);
      DLWP_CUDA(cudaMemcpy(p->atc_midW2[i], wimg.data(), wimg.size(), cudaMemcpyHostToDevice));
      DLWP_CUDA(cudaMemcpy(p->atc_midb[i], bias.data(), bias.size() * sizeof(float), cudaMemcpyHostToDevice));
    }
  }
  return DLWP_OK;
}

// out = filter(norm1(x)) + norm1(x) + x   (fourcastnet.py:176-187 with AFNO2D.forward :78-127)
// `embed` (first block, 1x1 patches): x is not read but computed here from the NCHW sources of the step
// (patch embedding + position embedding, fourcastnet.py:285-289) and stored for the residual uses.
int atc_block_filter(dlwp_plan* p, int block, float* x, float* out, int B, cudaStream_t st, const StepIO* embed) {
  const dlwp_model_desc& d = p->d;
  const int nb = d.num_blocks, bs = kAfnoE / nb;
  const float *g, *be;
  DLWP_TRY(plan_need(p, bname(block, ".norm1.weight"), kAfnoE, &g));
  DLWP_TRY(plan_need(p, bname(block, ".norm1.bias"), kAfnoE, &be));
  AfnoFwdArgs fa{};
  fa.x = x; fa.A = p->specA; fa.Fw = p->Fw; fa.gamma = g; fa.beta = be; fa.eps = d.layernorm_eps;
  fa.B = B; fa.H = p->gh; fa.W = p->gw; fa.Kc = d.n_cols;
  if (embed != nullptr) {
    fa.ngrp = embed->ngrp;
    for (int i = 0; i < embed->ngrp; ++i) fa.grp[i] = embed->grp[i];
    fa.Cin = p->Cin;
    DLWP_TRY(plan_need(p, "patch_embed.proj.weight", (int64_t)kAfnoE * p->Cin, &fa.We));
    DLWP_TRY(plan_need(p, "patch_embed.proj.bias", kAfnoE, &fa.be));
    if (d.use_pos_embed) DLWP_TRY(plan_need(p, "pos_embed", (int64_t)p->gh * p->gw * kAfnoE, &fa.pos));
    fa.xout = x;
  }
  AfnoWTcArgs wa{};
  wa.ntiles = B * p->gh * p->gw / kTcTile; wa.H = p->gh; wa.W = p->gw; wa.RT = p->atc_RT; wa.Kc = d.n_cols;
  wa.x = x; wa.gamma = g; wa.beta = be; wa.eps = d.layernorm_eps;
  wa.Fimg = p->atc_F; wa.A = reinterpret_cast<float*>(p->specA);
  wa.Gimg = p->atc_G; wa.KZ = p->atc_KZ; wa.Z = p->specZ; wa.out = out;
  const bool fwd_tc = p->atc_wtc && (embed == nullptr || p->Cin <= 8);
  if (embed != nullptr) {
    wa.ngrp = fa.ngrp;
    for (int i = 0; i < fa.ngrp; ++i) wa.grp[i] = fa.grp[i];
    wa.Cin = fa.Cin; wa.We = fa.We; wa.be = fa.be; wa.pos = fa.pos; wa.xout = fa.xout;
  }
  const bool mid2 = fwd_tc && p->atc_mid2 && afno_inv_tc_smem_bytes(p->atc_RT, d.n_cols, p->atc_KZ) <= 226 * 1024;
  if (mid2) {      // operand images between the three kernels instead of the fp32 spectra
    wa.Aimg = reinterpret_cast<uint8_t*>(p->specA);
    wa.ncols = B * d.n_cols;
    wa.Zimg = p->atc_Zimg;
    DLWP_TRY(launch_afno_fwd_tc(wa, st));
    AfnoMid2Args m2{};
    m2.Aimg = wa.Aimg; m2.Zimg = p->atc_Zimg; m2.Fimg = p->atc_midF; m2.Gimg = p->atc_midG;
    m2.W1img = p->atc_midW1[block]; m2.W2img = p->atc_midW2[block];
    m2.b1 = p->atc_midb[block]; m2.b2 = p->atc_midb[block] + 2 * kAfnoE;
    m2.lambda = d.sparsity_threshold;
    m2.ncols = wa.ncols; m2.H = p->gh; m2.R = d.n_rows; m2.Kc = d.n_cols; m2.bs = bs; m2.RT = p->atc_RT; m2.KZ = p->atc_KZ;
    DLWP_TRY(launch_afno_mid_tc(m2, st));
    return launch_afno_inv_tc(wa, st);
  }
  if (fwd_tc) DLWP_TRY(launch_afno_fwd_tc(wa, st));
  else DLWP_TRY(launch_afno_fwd(fa, st));
  AfnoMidArgs ma{};
  ma.planarA = fwd_tc ? 1 : 0;
  ma.A = p->specA; ma.Z = p->specZ; ma.Fh = p->Fh; ma.Gh = p->Gh;
  DLWP_TRY(plan_need(p, bname(block, ".filter.w1"), (int64_t)2 * nb * bs * bs, &ma.w1));
  DLWP_TRY(plan_need(p, bname(block, ".filter.b1"), (int64_t)2 * nb * bs, &ma.b1));
  DLWP_TRY(plan_need(p, bname(block, ".filter.w2"), (int64_t)2 * nb * bs * bs, &ma.w2));
  DLWP_TRY(plan_need(p, bname(block, ".filter.b2"), (int64_t)2 * nb * bs, &ma.b2));
  ma.lambda = d.sparsity_threshold;
  ma.B = B; ma.H = p->gh; ma.R = d.n_rows; ma.Kc = d.n_cols; ma.bs = bs;
  DLWP_TRY(launch_afno_mid(ma, st));
  if (p->atc_wtc && afno_inv_tc_smem_bytes(p->atc_RT, d.n_cols, p->atc_KZ) <= 226 * 1024) return launch_afno_inv_tc(wa, st);
  AfnoInvArgs ia{p->specZ, x, out, p->Gw, g, be, d.layernorm_eps, B, p->gh, p->gw, d.n_cols};
  return launch_afno_inv(ia, st);
}

// x = x + mlp(norm2(x))   (fourcastnet.py:189-192)
// `head` (last block, 1x1 patches): the network head + residual frame are applied in the epilogue and the NCHW
// frame of the step is written instead of `out` (fourcastnet.py:349-358).
int atc_block_mlp(dlwp_plan* p, int block, const float* x, float* out, int P, cudaStream_t st, const StepIO* head) {
  AfnoMlpArgs a{};
  if (head != nullptr) {
    a.Cg = p->d.prognostic_channels;
    a.HW = p->gh * p->gw;
    DLWP_TRY(plan_need(p, "head.weight", (int64_t)a.Cg * kAfnoE, &a.Wh));
    a.frame = head->out; a.frame_bstride = head->out_bstride;
    a.last = head->last; a.last_bstride = head->last_bstride;
  }
  a.ntiles = P / kTcTile;
  a.x = x;
  a.out = out;
  a.W1img = p->atc_W1[block];
  DLWP_REQUIRE((int)p->atc_b1h[block].size() == kAfnoHid && (int)p->atc_b2h[block].size() == kAfnoE, "afno_mlp_kernel: weights not committed");
  std::memcpy(a.b1c, p->atc_b1h[block].data(), sizeof(a.b1c));
  std::memcpy(a.b2c, p->atc_b2h[block].data(), sizeof(a.b2c));
  a.W2img = p->atc_W2[block];
  a.eps = p->d.layernorm_eps;
  return launch_afno_mlp(a, st);
}

}  // namespace dlwp
