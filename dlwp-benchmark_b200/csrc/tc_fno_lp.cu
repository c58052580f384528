// fno_layerproj_kernel: the LAST spectral layer and the projection MLP of the FNO / TFNO step in one kernel
//   y   = irfft_W(Z') + skip(x) + bias            (neuralop FNOBlocks: no activation after the last block)
//   out = last + W2 * GELU(W1 * y + b1) + b2      (neuralop MLP projection + fno.py:103)
// fno_layer_kernel wrote y as a 16 KB tile image that fno_proj_kernel read straight back (67 MB each way per evaluation at
// C4, a launch boundary in between).  Here the layer's accumulator tile D0 [128 x 32] goes accumulator -> + bias -> split-bf16
// pairs written back into the same tensor-memory columns -> A operand of the projection's first GEMM (tcgen05.mma with the A
// operand in tensor memory), and from there the projection runs as in fno_proj_kernel (hidden layer in tensor memory, output
// taken over one tile late, residual frame rows by bulk copy).
//
// Per worker (two per CTA, 256 tensor-memory columns each): D1 at +0 / +64, D2 at +128 / +160, D0 at +192 / +224.
//   control warp A : input bulk copies (activation image + Z' image, two slots) and the layer's UMMAs -> D0[u & 1]
//   control warp B : the projection's UMMAs (first GEMM from D0's columns, second from D1's) and the residual-frame rows
//   8 epilogue warps: D0 -> operand (one tile ahead, after pass P - 2 of the tile before), the MLP epilogue, the output
#include "tc_mlp.cuh"

namespace dlwp {

using namespace tc;

namespace {

constexpr int kLpSlots = 2;
constexpr int kCtaLp = 2 * kEpi + 4 * 32;      // 2 x 8 epilogue warps + 2 x 2 control warps
enum { PB_XFULL = 0, PB_XFREE = kLpSlots, PB_D0FULL = 2 * kLpSlots, PB_PER = 2 * kLpSlots + 2 };

struct LpSmem {   // byte offsets; sized for hid = 256
  int W1, W2, W1x, Ones, B2, G, Ws, Bias, Worker, X, Z, Last, WorkerBytes, Total;
  __host__ __device__ explicit LpSmem(int KZ) {
    W1 = 0;
    W2 = W1 + 256 * 32 * 4;
    W1x = W2 + 16 * 256 * 4;
    Ones = W1x + 256 * 16 * 2;
    B2 = Ones + 4096;
    G = B2 + 128;
    Ws = G + 2 * 128 * KZ * 2;
    Bias = Ws + 4096;
    Worker = Bias + 128;
    X = 0;
    Z = X + kLpSlots * kTileBytes;
    Last = Z + kLpSlots * KZ * 128;
    WorkerBytes = Last + 2 * 16 * 512;
    Total = Worker + 2 * WorkerBytes;
  }
};
constexpr int kLpSmemMax = 226 * 1024;

}  // namespace

__global__ void __launch_bounds__(kCtaLp, 1) fno_layerproj_kernel(const TcLayerProjArgs a) {
  extern __shared__ __align__(128) uint8_t smem[];
  pdl_launch_dependents();
  __shared__ uint64_t bars[2 * MB_PER + 2 * PB_PER + 1];
  __shared__ uint32_t tmem_s;
  const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0);   // provably warp-uniform
  const TcGeom g = a.layer.g;
  const LpSmem L(g.KZ);
  const int hid = a.proj.hid, P = hid / 64;
  float* sB2 = reinterpret_cast<float*>(smem + L.B2);
  float* sBias = reinterpret_cast<float*>(smem + L.Bias);
  uint64_t* lbars = bars + 2 * MB_PER;
  uint64_t* bar_const = &bars[2 * MB_PER + 2 * PB_PER];

  if (tid == 0) {
    mlp_init_barriers(bars, kEpi);      // A1FULL: the 256 epilogue threads hand the operand over
    mlp_init_barriers(bars + MB_PER, kEpi);
    for (int w = 0; w < 2; ++w) {
      uint64_t* lb = lbars + w * PB_PER;
      for (int i = 0; i < kLpSlots; ++i) {
        mbar_init(&lb[PB_XFULL + i], 1);
        mbar_init(&lb[PB_XFREE + i], 1);
      }
      mbar_init(&lb[PB_D0FULL], 1);
      mbar_init(&lb[PB_D0FULL + 1], 1);
    }
    mbar_init(bar_const, 1);
    fence_mbar_init();
  }
  if (warp == kCtlWarp) {
    tmem_alloc(&tmem_s, 512);
    tmem_relinquish();
  }
  {   // Ones[128, 16]: columns 0 and 1 are 1 (they multiply the hi and lo halves of the bias in W1x)
    uint32_t* o = reinterpret_cast<uint32_t*>(smem + L.Ones);
    for (int i = tid; i < 1024; i += kCtaLp) o[i] = (i < 512 && (i & 3) == 0) ? 0x3F803F80u : 0u;
  }
  fence_async_smem();
  if (tid < 16) sB2[tid] = tid < a.proj.Cg ? a.proj.b2[tid] : 0.f;
  if (tid >= 32 && tid < 64) sBias[tid - 32] = a.layer.bias ? a.layer.bias[tid - 32] : 0.f;
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t w1_bytes = (uint32_t)hid * 32 * 4, w2_bytes = 16u * hid * 4, w1x_bytes = (uint32_t)hid * 16 * 2;
  const uint32_t g_bytes = 2u * 128u * (uint32_t)g.KZ * 2u, z_bytes = (uint32_t)g.KZ * 128u;
  if (tid == 0) {   // constant operands: requested before the programmatic-dependent-launch wait
    mbar_arrive_expect_tx(bar_const, w1_bytes + w2_bytes + w1x_bytes + g_bytes + 4096);
    bulk_g2s(smem + L.W1, a.proj.W1img, w1_bytes, bar_const);
    bulk_g2s(smem + L.W2, a.proj.W2img, w2_bytes, bar_const);
    bulk_g2s(smem + L.W1x, a.proj.W1ximg, w1x_bytes, bar_const);
    bulk_g2s(smem + L.G, a.layer.Gimg, g_bytes, bar_const);
    bulk_g2s(smem + L.Ws, a.layer.Wsimg, 4096, bar_const);
  }
  pdl_wait();   // everything below reads what the previous kernel wrote (or overwrites what it read)
  const int stride = gridDim.x * 2;

  if (warp >= kCtlWarp + 2) {
    // ---------------- control warp A: input copies and the layer's UMMAs ----------------
    const uint32_t leader = elect_one();
    const int w = warp - kCtlWarp - 2;
    uint64_t* b = bars + w * MB_PER;
    uint64_t* lb = lbars + w * PB_PER;
    uint8_t* ws = smem + L.Worker + w * L.WorkerBytes;
    const uint32_t tmem = tmem_s + 256u * w;
    const int first = (int)blockIdx.x + w * (int)gridDim.x;
    auto load_inputs = [&](int tile, uint32_t u) {
      const int s = u % kLpSlots;
      mbar_wait(&lb[PB_XFREE + s], ((u / kLpSlots) & 1u) ^ 1u);   // the UMMAs that read this slot last are done
      mbar_arrive_expect_tx_p(leader, &lb[PB_XFULL + s], kTileBytes + z_bytes);
      const uint64_t pol_in = g.l2hint ? l2_evict_first() : 0;      // dead after this read
      bulk_g2s_hp(leader, ws + L.X + s * kTileBytes, a.layer.x_img + (size_t)tile * kTileBytes, kTileBytes, &lb[PB_XFULL + s], pol_in);
      bulk_g2s_hp(leader, ws + L.Z + s * z_bytes, a.layer.Zimg + (size_t)tile * z_bytes, z_bytes, &lb[PB_XFULL + s], pol_in);
    };
    for (int i = 0; i < kLpSlots; ++i)
      if (first + i * stride < g.ntiles) load_inputs(first + i * stride, i);
    mbar_wait(bar_const, 0);
    const uint32_t aG = smem_u32(smem + L.G), aW = smem_u32(smem + L.Ws);
    const uint64_t dG = chunk_desc(aG, 128, 0), dW = chunk_desc(aW, 32, 0);
    const uint32_t idesc = idesc_bf16(128, 32);
    const uint32_t g_lo = 128u * (uint32_t)g.KZ * 2u, z_lo = (uint32_t)g.KZ * 64u;
    uint32_t u = 0;
    for (int tile = first; tile < g.ntiles; tile += stride, ++u) {
      const int s = u % kLpSlots, db = u & 1;
      mbar_wait(&lb[PB_XFULL + s], (u / kLpSlots) & 1u);
      mbar_wait(&b[MB_A1EMPTY + db], empty_par(u));      // the projection's first GEMM has read the operand in these columns
      tc_fence_after();
      // Z': MN-major B operand (K rows of 16 bytes = 8 channels; 8-row core matrices 128 B apart along K, KZ * 16 B along the channels)
      const uint64_t dX = chunk_desc(smem_u32(ws + L.X + s * kTileBytes), 128, 0),
                     dZ = smem_desc(smem_u32(ws + L.Z + s * z_bytes), 128, (uint32_t)g.KZ * 16u);
      const uint32_t d = tmem + 192u + 32u * db;
      uint32_t acc = 0;
#pragma unroll
      for (int term = 0; term < 3; ++term) {      // hi*hi, lo*hi, hi*lo
        const uint32_t ga = term == 1 ? g_lo : 0u, zb = term == 2 ? z_lo : 0u;
        const uint32_t xa = term == 1 ? (uint32_t)kImg : 0u, wb = term == 2 ? 2048u : 0u;
#pragma unroll
        for (int k0 = 0; k0 < 128; k0 += 16) {      // (unrolled with a uniform guard: descriptors stay in uniform registers)
          if (k0 < g.KZ) {
            umma_bf16_p(leader, d, desc_off(dG, ga + (k0 >> 3) * 2048), desc_off(dZ, zb + (uint32_t)k0 * 16u), idesc | kIdescBMnMajor, acc);
            acc = 1;
          }
        }
#pragma unroll
        for (int k0 = 0; k0 < 32; k0 += 16)
          umma_bf16_p(leader, d, desc_off(dX, xa + (k0 >> 3) * 2048), desc_off(dW, wb + (k0 >> 3) * 512), idesc, 1);
      }
      umma_commit_p(leader, &lb[PB_D0FULL + db]);
      umma_commit_p(leader, &lb[PB_XFREE + s]);
      __syncwarp();
      if (tile + kLpSlots * stride < g.ntiles) load_inputs(tile + kLpSlots * stride, u + kLpSlots);
    }
  } else if (warp >= kCtlWarp) {
    // ---------------- control warp B: the projection's UMMAs, the residual-frame rows ----------------
    const uint32_t leader = elect_one();
    const int w = warp - kCtlWarp;
    uint64_t* b = bars + w * MB_PER;
    uint8_t* ws = smem + L.Worker + w * L.WorkerBytes;
    MlpIssuer is{b, 0u, 0u, smem_u32(smem + L.W1), smem_u32(smem + L.W2), tmem_s + 256u * w, hid, P, 0u,
                 smem_u32(smem + L.Ones), smem_u32(smem + L.W1x), leader};
    const uint32_t tA1 = tmem_s + 256u * w + 192u;
    const int first = (int)blockIdx.x + w * (int)gridDim.x;
    mbar_wait(bar_const, 0);
    // the tile's rows of the residual frame (128 consecutive pixels per channel) for the epilogue: bulk copies,
    // because a generic global load in flight in an epilogue thread would be waited for by its next proxy fence
    auto load_last = [&](int tile, uint32_t u) {
      const int s = u & 1;
      const long long bb = tile / g.TPS, pix = (long long)(tile - (int)bb * g.TPS) * kTcTile;
      mbar_wait(&b[MB_INEMPTY + s], empty_par(u));
      mbar_arrive_expect_tx_p(leader, &b[MB_INFULL + s], (uint32_t)a.proj.Cg * 512u);
      for (int c = 0; c < a.proj.Cg; ++c)
        bulk_g2s_p(leader, ws + L.Last + s * 8192 + c * 512, a.proj.last + bb * a.proj.last_bstride + (long long)c * g.HW + pix, 512u,
                   &b[MB_INFULL + s]);
    };
    const bool bulk_last = a.proj.last != nullptr && a.proj.last_bulk;
    if (first < g.ntiles) is.mma1_ts<32>(tA1, 0, 0, 0);
    uint32_t u = 0;
    for (int tile = first; tile < g.ntiles; tile += stride, ++u) {
      const bool has_next = tile + stride < g.ntiles;
      if (bulk_last) load_last(tile, u);
      for (int p = 0; p < P; ++p) {
        const uint32_t gp = u * P + p;
        if (p + 1 < P) is.mma1_ts<32>(tA1, u, p + 1, gp + 1);
        else if (has_next) is.mma1_ts<32>(tA1, u + 1, 0, gp + 1);
        is.mma2<16>(u, p);
      }
    }
  } else {
    // ------------------------------ epilogue group ------------------------------
    const EpiId id(tid);
    const int w = id.w;
    uint64_t* b = bars + w * MB_PER;
    uint64_t* lb = lbars + w * PB_PER;
    uint8_t* ws = smem + L.Worker + w * L.WorkerBytes;
    const uint32_t tmem_lane = tmem_s + 256u * w + ((uint32_t)(id.q4 * 32) << 16) + 16u * id.half;
    const int c0 = 8 * id.half;      // this thread's 8 output channels
    uint32_t u = 0, gp = 0, nc = 0;
    // the layer's accumulator tile -> + bias (GELU when the layer has one) -> operand of the first projection GEMM, in place:
    // this thread's 16 columns become 8 columns of hi pairs and 8 of lo pairs = one K step of the operand
    auto make_a1 = [&](uint32_t v) {
      const int db = v & 1;
      mbar_wait(&lb[PB_D0FULL + db], full_par(v));
      tc_fence_after();
      uint32_t r[16], h[8], l[8];
      tmem_ld16(tmem_lane + 192u + 32u * db, r);
      tmem_ld_wait();
      const float2* bias2 = reinterpret_cast<const float2*>(sBias + 16 * id.half);
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        const float2 bb = bias2[c];
        const float2 x = make_float2(__uint_as_float(r[2 * c]) + bb.x, __uint_as_float(r[2 * c + 1]) + bb.y);
        split2p(a.layer.gelu ? gelu2(x) : x, h[c], l[c]);
      }
      tmem_st8(tmem_lane + 192u + 32u * db, h);
      tmem_st8(tmem_lane + 192u + 32u * db + 8u, l);
      tmem_st_wait();
      tc_fence_before();
      mbar_arrive(&b[MB_A1FULL + db]);
    };
    // A tile is finished one tile late, after the first pass of the next one: its second GEMM completes under
    // that pass instead of being waited for (D2 is double buffered in TMEM).
    float lastv[8];
    long long obase = 0;
    const bool bulk_last = a.proj.last != nullptr && a.proj.last_bulk;
    const float* sLast = reinterpret_cast<const float*>(ws + L.Last);
    auto finish = [&](uint32_t v) {
      if (bulk_last) {
        mbar_wait(&b[MB_INFULL + (v & 1)], full_par(v));
#pragma unroll
        for (int c = 0; c < 8; ++c) lastv[c] = c0 + c < a.proj.Cg ? sLast[(v & 1) * 2048 + (c0 + c) * 128 + id.row] : 0.f;
        mbar_arrive(&b[MB_INEMPTY + (v & 1)]);
      }
      mbar_wait(&b[MB_D2FULL + (v & 1)], full_par(v));
      tc_fence_after();
      uint32_t r[8];
      tmem_ld8(tmem_s + 256u * w + ((uint32_t)(id.q4 * 32) << 16) + 128u + 32u * (v & 1) + (uint32_t)c0, r);
      tmem_ld_wait();
      tc_fence_before();
      mbar_arrive(&b[MB_D2EMPTY + (v & 1)]);
      float* o = a.proj.out + obase;
#pragma unroll
      for (int c = 0; c < 8; ++c)
        if (c0 + c < a.proj.Cg) o[(long long)(c0 + c) * g.HW] = __uint_as_float(r[c]) + sB2[c0 + c] + lastv[c];
    };
    auto load_last = [&](int tile) {      // output position; without bulk copies also the residual frame values
      const long long bb = tile / g.TPS, pix = (long long)(tile - (int)bb * g.TPS) * kTcTile + id.row;
      if (!bulk_last) {
#pragma unroll
        for (int c = 0; c < 8; ++c)
          lastv[c] = (a.proj.last != nullptr && c0 + c < a.proj.Cg) ? a.proj.last[bb * a.proj.last_bstride + (long long)(c0 + c) * g.HW + pix] : 0.f;
      }
      obase = bb * a.proj.out_bstride + pix;
    };
    int prev = -1;
    const int first = (int)blockIdx.x + w * (int)gridDim.x;
    if (first < g.ntiles) make_a1(0);
    const int a1_pass = P >= 3 ? P - 2 : 0;
    for (int tile = first; tile < g.ntiles; tile += stride, ++u) {
      const bool has_next = tile + stride < g.ntiles;
      if (prev >= 0) load_last(prev);
      epi_mlp_tile(b, id, tmem_lane, P, gp, nc, [&](int pass) {
        if (pass == 0 && prev >= 0) finish(u - 1);
        if (pass == a1_pass && has_next) make_a1(u + 1);
      });
      prev = tile;
    }
    if (prev >= 0) {
      load_last(prev);
      finish(u - 1);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == kCtlWarp) tmem_dealloc(tmem_s, 512);
}

int launch_tc_layerproj(const TcLayerProjArgs& a_in, cudaStream_t st) {
  TcLayerProjArgs a = a_in;
  a.proj.last_bulk = a.proj.last != nullptr && (reinterpret_cast<uintptr_t>(a.proj.last) & 15) == 0 && a.proj.last_bstride % 4 == 0 &&
                     a.layer.g.HW % 4 == 0;
  int sms = 0;
  DLWP_TRY(current_sms(&sms));
  const LpSmem L(a.layer.g.KZ);
  DLWP_REQUIRE(a.proj.hid == 128 || a.proj.hid == 256, "tensor-core projection needs width 128 or 256, got %d", a.proj.hid);
  DLWP_REQUIRE(a.layer.g.KZ >= 16 && a.layer.g.KZ % 16 == 0 && L.Total <= kLpSmemMax, "fno_layerproj_kernel: KZ=%d does not fit", a.layer.g.KZ);
  DLWP_TRY(ensure_dynamic_smem((const void*)fno_layerproj_kernel, kLpSmemMax));
  const int grid = a.layer.g.ntiles < sms ? a.layer.g.ntiles : sms;      // (tile t of a CTA's list goes to worker (t / grid) % 2)
  DLWP_CUDA(launch_pdl(fno_layerproj_kernel, dim3(grid), dim3(kCtaLp), (size_t)L.Total, st, a));
  DLWP_LAUNCHED("fno_layerproj_kernel");
  return DLWP_OK;
}

// largest KZ the fused kernel has shared memory for
bool tc_layerproj_fits(int KZ) { return KZ >= 16 && KZ % 16 == 0 && KZ <= 128 && LpSmem(KZ).Total <= kLpSmemMax; }

}  // namespace dlwp
