// tcgen05 kernels of the AFNO block (see tc_afno.cuh).
#include "tc_afno.cuh"

#include "tc_mlp.cuh"

namespace dlwp {

using namespace tc;

#ifdef DLWP_TRACE
__device__ unsigned long long* g_trace_afno = nullptr;   // [16 tiles][8 events] clock64 stamps of CTA 0 (tools/trace_afno.py)
#define TRA(u, ev) do { if (g_trace_afno && blockIdx.x == 0 && (u) < 16) g_trace_afno[(u) * 8 + (ev)] = clock64(); } while (0)
#else
#define TRA(u, ev) do {} while (0)
#endif

// =============================================================================================
// afno_mlp_kernel: out = x + fc2(GELU(fc1(LN(x))))  on 128-token tiles, persistent, one CTA per SM
//   warps 0-7 / 8-15 : the epilogue groups of two workers (two threads per token: channel halves)
//   warps 16 / 17    : their control warps
// Both 64 KB weight operands ([hi | lo] split-bf16 images) stay in shared memory for the whole kernel and are
// shared by the two workers, which take the CTA's tiles alternately: the TMEM -> GELU -> MMA chain of one tile's
// four passes is a latency chain, so a second, independent tile per SM is what fills the pipes.  Per worker
// there is one A1 buffer (normalised tokens, 32 KB; the next tile's tokens are stored after the last first-layer
// MMA of this tile has read it, two passes before they are needed) and 16 KB of staging for the output rows.
// The 256-wide hidden layer never touches shared memory: the GELU epilogue writes the split-bf16 pairs back into
// the accumulator columns it has just read and the second GEMM takes its A operand from tensor memory
// (tc_mlp.cuh).  The hidden accumulator D1 [128 x 64] and the output accumulator D2 [128 x 64] are double-buffered
// in tensor memory (256 columns per worker).
// LayerNorm: a warp takes 8 tokens per step, lane = token + 8 x channel-octet pair (a load instruction touches 8
// lines of 128 bytes, not 32), statistics over the four lanes of a token by shuffles, two passes like torch.
// The biases travel as kernel parameters (constant bank).  What bounds the kernel is the L1 data pipe
// (DESIGN.md section 4d): every access pattern here is chosen for few 128-byte wavefronts.
// =============================================================================================
namespace {
constexpr int kAmCta = kCta;      // 2 x 256 epilogue threads + 2 control warps
struct AmSmem {
  static constexpr int W1 = 0, W2 = W1 + kAfnoHid * kAfnoE * 4, Worker = W2 + kAfnoE * kAfnoHid * 4, A1 = 0,
                       Stage = 32768,      // 16 KB: the output rows on their way to coalesced stores
                       WorkerBytes = Stage + kTileBytes, Total = Worker + 2 * WorkerBytes;
  static_assert(Total <= 224 * 1024, "afno_mlp_kernel shared memory");
};
}  // namespace

__global__ void __launch_bounds__(kAmCta, 1) afno_mlp_kernel(const AfnoMlpArgs a) {
  extern __shared__ __align__(128) uint8_t smem[];
  pdl_launch_dependents();
  __shared__ uint64_t bars[2 * MB_PER + 1];
  __shared__ uint32_t tmem_s;
  const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0);   // provably warp-uniform
  constexpr int hid = kAfnoHid, P = kAfnoHid / 64;
  uint64_t* bar_const = &bars[2 * MB_PER];

  if (tid == 0) {
    mlp_init_barriers(bars, kEpi);
    mlp_init_barriers(bars + MB_PER, kEpi);
    mbar_init(bar_const, 1);
    fence_mbar_init();
  }
  if (warp == kCtlWarp) {
    tmem_alloc(&tmem_s, 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (tid == 0) {   // constant operands: on their way before the predecessor has finished
    mbar_arrive_expect_tx(bar_const, 2u * 65536u);
    bulk_g2s(smem + AmSmem::W1, a.W1img, 65536, bar_const);
    bulk_g2s(smem + AmSmem::W2, a.W2img, 65536, bar_const);
  }
  pdl_wait();   // everything below reads what the previous kernel wrote (or overwrites what it read)
  const int stride = gridDim.x * 2;

  if (warp >= kCtlWarp) {
    const uint32_t leader = elect_one();
    const int w = warp - kCtlWarp;
    uint64_t* b = bars + w * MB_PER;
    uint8_t* ws = smem + AmSmem::Worker + w * AmSmem::WorkerBytes;
    MlpIssuer is{b, smem_u32(ws + AmSmem::A1), 32768u, smem_u32(smem + AmSmem::W1), smem_u32(smem + AmSmem::W2),
                 tmem_s + 256u * w, hid, P, 0u, 0u, 0u, leader, 1u};
    const int first = blockIdx.x + w * (int)gridDim.x;      // worker 1 starts on the CTA's second tile
    mbar_wait(bar_const, 0);
    if (first < a.ntiles) is.mma1<kAfnoE>(0, 0, 0);
    uint32_t u = 0;
    for (int tile = first; tile < a.ntiles; tile += stride, ++u) {
      const bool has_next = tile + stride < a.ntiles;
      for (int p = 0; p < P; ++p) {
        const uint32_t gp = u * P + p;
        if (p + 1 < P) is.mma1<kAfnoE>(u, p + 1, gp + 1);
        else if (has_next) is.mma1<kAfnoE>(u + 1, 0, gp + 1);
        is.mma2<kAfnoE>(u, p);
      }
    }
  } else {
    const EpiId id(tid);
    const int w = id.w;
    uint64_t* b = bars + w * MB_PER;
    uint8_t* ws = smem + AmSmem::Worker + w * AmSmem::WorkerBytes;
    const uint32_t tmem_w = tmem_s + 256u * w, tmem_row = tmem_w + ((uint32_t)(id.q4 * 32) << 16), tmem_lane = tmem_row + 16u * id.half;
    // gather / LayerNorm coordinates: a warp takes 8 tokens per step, lane = token t + 8 g; the thread holds channel octets
    // g and g + 4 of its token (a load instruction touches 8 lines of 128 bytes, not 32), statistics over the four lanes
    // of a token by shuffles (two passes like torch)
    const int lw = (tid >> 5) & (kEpiWarps - 1), ltid = tid & (kEpi - 1), gt = lw * 8 + (tid & 7), gg = (tid >> 3) & 3;

    auto gather = [&](int tile, float4 (&x)[8]) {
#pragma unroll
      for (int st = 0; st < 2; ++st) {
        const float4* src = reinterpret_cast<const float4*>(a.x + ((size_t)tile * kTcTile + st * 64 + gt) * kAfnoE);
        x[4 * st] = src[2 * gg];
        x[4 * st + 1] = src[2 * gg + 1];
        x[4 * st + 2] = src[8 + 2 * gg];
        x[4 * st + 3] = src[8 + 2 * gg + 1];
      }
    };
    auto store_a1 = [&](uint32_t u, float4 (&x)[8]) {
      float mean[2], rstd[2];
#pragma unroll
      for (int st = 0; st < 2; ++st) {
        float s = 0.f;
#pragma unroll
        for (int j = 0; j < 4; ++j) s += (x[4 * st + j].x + x[4 * st + j].y) + (x[4 * st + j].z + x[4 * st + j].w);
        s += __shfl_xor_sync(0xffffffffu, s, 8);
        s += __shfl_xor_sync(0xffffffffu, s, 16);
        mean[st] = s * (1.0f / 64.0f);
        float q = 0.f;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float d0 = x[4 * st + j].x - mean[st], d1 = x[4 * st + j].y - mean[st], d2 = x[4 * st + j].z - mean[st], d3 = x[4 * st + j].w - mean[st];
          q = fmaf(d0, d0, q); q = fmaf(d1, d1, q); q = fmaf(d2, d2, q); q = fmaf(d3, d3, q);
        }
        q += __shfl_xor_sync(0xffffffffu, q, 8);
        q += __shfl_xor_sync(0xffffffffu, q, 16);
        rstd[st] = rsqrtf(q * (1.0f / 64.0f) + a.eps);
      }
      mbar_wait(&b[MB_A1EMPTY], (u & 1u) ^ 1u);      // ring of depth 1
#pragma unroll
      for (int st = 0; st < 2; ++st) {
        uint8_t* d = ws + AmSmem::A1 + gg * 2048 + (st * 64 + gt) * 16;      // 8 channels = one 16-byte unit of the chunk-major image
#pragma unroll
        for (int o = 0; o < 2; ++o) {      // octets g and g + 4
          uint32_t h[4], l[4];
          const float4 v0 = x[4 * st + 2 * o], v1 = x[4 * st + 2 * o + 1];
          const float m = mean[st], r = rstd[st];
          split2p(make_float2((v0.x - m) * r, (v0.y - m) * r), h[0], l[0]);
          split2p(make_float2((v0.z - m) * r, (v0.w - m) * r), h[1], l[1]);
          split2p(make_float2((v1.x - m) * r, (v1.y - m) * r), h[2], l[2]);
          split2p(make_float2((v1.z - m) * r, (v1.w - m) * r), h[3], l[3]);
          *reinterpret_cast<uint4*>(d + o * 8192) = make_uint4(h[0], h[1], h[2], h[3]);
          *reinterpret_cast<uint4*>(d + 16384 + o * 8192) = make_uint4(l[0], l[1], l[2], l[3]);
        }
      }
      fence_async_smem();
      mbar_arrive(&b[MB_A1FULL]);
    };

    const int first = blockIdx.x + w * (int)gridDim.x;      // worker 1 starts on the CTA's second tile
    if (first < a.ntiles) {
      float4 nx[8];
      gather(first, nx);
      store_a1(0, nx);
    }
    // out = x + D2 + b2.  The accumulator rows pass through the worker's staging buffer, 32 channels
    // at a time: thread (row, half) writes its 128 bytes (16-byte units XOR-swizzled by the row: conflict-free), then the
    // 256 threads of the worker add x and the bias and store with a token's 8 units on consecutive lanes: 4 lines of 128
    // bytes per load / store instruction instead of 32 (the L1 data pipe is what bounds this kernel).
    // Fused head (last block, 1x1 patches; fourcastnet.py:349-358): instead of storing `out`, the 8 lanes of a token reduce
    // their share of the 64-channel dot products by shuffles, over both rounds; lane 0 adds the residual frame and writes NCHW.
    auto finish = [&](int tile, uint32_t v_) {
      uint8_t* stage = ws + AmSmem::Stage;
      mbar_wait(&b[MB_D2FULL + (v_ & 1)], full_par(v_));
      tc_fence_after();
      // head: [row of the thread][output channel], starting from the residual frame (lane 0 of a token's 8 lanes; the loads
      // are issued here, ahead of the accumulator wait, not one by one in front of the stores)
      // (a tile lies within one sample: HW is a multiple of the tile size)
      const long long p0 = (long long)tile * kTcTile, bb = a.Wh != nullptr ? p0 / a.HW : 0, pix0 = p0 - bb * a.HW;
      float hp[4][4];
#pragma unroll
      for (int it = 0; it < 4; ++it)
#pragma unroll
        for (int co = 0; co < 4; ++co)
          hp[it][co] = (a.Wh != nullptr && a.last != nullptr && co < a.Cg && (ltid & 7) == 0)
                           ? __ldg(a.last + bb * a.last_bstride + (long long)co * a.HW + pix0 + it * 32 + (ltid >> 3))
                           : 0.f;
      const int un = ltid & 7;
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        float4 wh[4];      // head: this thread's 4 channels of the round, per output channel
        if (a.Wh != nullptr) {
#pragma unroll
          for (int co = 0; co < 4; ++co)
            wh[co] = co < a.Cg ? __ldg(reinterpret_cast<const float4*>(a.Wh + co * kAfnoE + 32 * h + 4 * un)) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
        // the residual rows of the round: requested here, ahead of the accumulator round trip and the barrier
        float4 xv[4];
#pragma unroll
        for (int it = 0; it < 4; ++it)
          xv[it] = *reinterpret_cast<const float4*>(a.x + ((size_t)tile * kTcTile + it * 32 + (ltid >> 3)) * kAfnoE + 32 * h + 4 * un);
        if (id.half == h) {
          uint32_t r[32];
          tmem_ld32(tmem_row + 128u + 64u * (v_ & 1) + 32u * h, r);
          tmem_ld_wait();
          tc_fence_before();
          mbar_arrive(&b[MB_D2EMPTY + (v_ & 1)]);
          uint8_t* srow = stage + id.row * 128;
#pragma unroll
          for (int j = 0; j < 8; ++j)
            *reinterpret_cast<uint4*>(srow + ((j ^ (id.row & 7)) << 4)) = make_uint4(r[4 * j], r[4 * j + 1], r[4 * j + 2], r[4 * j + 3]);
        }
        bar_sync(1 + w, kEpi);
#pragma unroll
        for (int it = 0; it < 4; ++it) {
          const int row = it * 32 + (ltid >> 3), c = 32 * h + 4 * un;
          const float4 d = *reinterpret_cast<const float4*>(stage + row * 128 + ((un ^ (row & 7)) << 4));
          const size_t off = ((size_t)tile * kTcTile + row) * kAfnoE + c;
          const float4 v = make_float4(d.x + a.b2c[c] + xv[it].x, d.y + a.b2c[c + 1] + xv[it].y, d.z + a.b2c[c + 2] + xv[it].z,
                                       d.w + a.b2c[c + 3] + xv[it].w);
          if (a.Wh == nullptr) {
            *reinterpret_cast<float4*>(a.out + off) = v;
          } else {
#pragma unroll
            for (int co = 0; co < 4; ++co)
              hp[it][co] = fmaf(v.x, wh[co].x, fmaf(v.y, wh[co].y, fmaf(v.z, wh[co].z, fmaf(v.w, wh[co].w, hp[it][co]))));
          }
        }
        bar_sync(1 + w, kEpi);      // the staging rows are rewritten by the next round / the next tile's first chunk
      }
      if (a.Wh != nullptr) {
#pragma unroll
        for (int it = 0; it < 4; ++it) {
          const long long pix = pix0 + it * 32 + (ltid >> 3);
#pragma unroll
          for (int co = 0; co < 4; ++co)
            if (co < a.Cg) {
              float s = hp[it][co];
              s += __shfl_xor_sync(0xffffffffu, s, 1);
              s += __shfl_xor_sync(0xffffffffu, s, 2);
              s += __shfl_xor_sync(0xffffffffu, s, 4);
              if ((ltid & 7) == 0) a.frame[bb * a.frame_bstride + (long long)co * a.HW + pix] = s;
            }
        }
      }
    };
    uint32_t u = 0, gp = 0, nc = 0;
    for (int tile = first; tile < a.ntiles; tile += stride, ++u) {
      const bool has_next = tile + stride < a.ntiles;
      epi_mlp_tile(b, id, tmem_lane, P, gp, nc, [&](int pass) {
        if (pass == P - 2 && has_next) {      // (not prefetched over the passes: registers; the other worker runs meanwhile)
          float4 nx[8];
          gather(tile + stride, nx);
          store_a1(u + 1, nx);
        }
      }, [&](int p, int j, uint32_t (&r)[16]) {
        const int c = p * 64 + 32 * j + 16 * id.half;
#pragma unroll
        for (int e = 0; e < 16; ++e) r[e] = __float_as_uint(__uint_as_float(r[e]) + a.b1c[c + e]);
      });
      finish(tile, u);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == kCtlWarp) tmem_dealloc(tmem_s, 512);
}

int launch_afno_mlp(const AfnoMlpArgs& a, cudaStream_t st) {
  int sms = 0;
  DLWP_TRY(current_sms(&sms));
  DLWP_TRY(ensure_dynamic_smem((const void*)afno_mlp_kernel, AmSmem::Total));
  DLWP_REQUIRE(a.ntiles > 0, "afno_mlp_kernel: empty problem");
  DLWP_REQUIRE(a.Wh == nullptr || (a.Cg >= 1 && a.Cg <= 4 && a.HW > 0 && a.HW % kTcTile == 0 && a.frame != nullptr),
               "afno_mlp_kernel: bad head arguments");
  const int grid = a.ntiles < sms ? a.ntiles : sms;      // two workers per CTA: the second one has work when there are more tiles than CTAs
  DLWP_CUDA(launch_pdl(afno_mlp_kernel, dim3(grid), dim3(kAmCta), (size_t)AmSmem::Total, st, a));
  DLWP_LAUNCHED("afno_mlp_kernel");
  return DLWP_OK;
}

// =============================================================================================
// The truncated W-DFTs of the AFNO token mixer on tcgen05 (fourcastnet.py:84-85,124-127), 128-token tiles
// (= RT grid rows), persistent, one CTA per SM.
//
// afno_fwd_tc_kernel:  A[(h2, kx, re | im), c] = sum_px F''[(h2, kx, re | im), px] * LayerNorm(x)[px, c]
//   Three independent worker groups of 128 threads take the CTA's tiles in turn, one thread per token: the
//   token's 64 channels come straight from global memory into registers (the next tile's loads are issued
//   before this tile's accumulator is drained), LayerNorm is thread-local, and the normalised token goes
//   into the group's split-bf16 image, which the tensor core reads in place as an MN-major B operand
//   (tokens = K rows; a warp's lanes are 32 consecutive tokens, so every operand store is 512 contiguous
//   bytes).  A proxy fence waits for the global accesses its thread has in flight, so a group stalls for
//   about a memory latency per tile: the other two groups work meanwhile.  The control warp issues the
//   UMMAs of the tiles in order; a group drains its accumulator rows (one retained column of one grid
//   row, Re or Im) from TMEM lanes to their 256-byte rows of the planar spectrum.
// =============================================================================================
namespace {
constexpr int kFwGroups = 3, kFwGrp = 128, kFwWork = kFwGroups * kFwGrp, kFwCta = kFwWork + 32, kFwCtlWarp = kFwWork / 32;
enum { WB_YFULL = 0, WB_DFULL = kFwGroups, WB_CONST = 2 * kFwGroups, WB_N = 2 * kFwGroups + 1 };
struct FwdSmem {
  static constexpr int F = 0, Y = F + 2 * 128 * 128 * 2, GB = Y + kFwGroups * 32768, Emb = GB + 2 * kAfnoE * 4,
                       In = Emb + (16 * kAfnoE + kAfnoE) * 4,      // first block: the step-input channels of a group's tile [8][128]
                       Total = In + kFwGroups * 8 * 128 * 4;
};
constexpr int kWtWork = 256, kWtCta = kWtWork + 32, kWtCtlWarp = kWtWork / 32;     // afno_inv_tc_kernel

// LayerNorm of one token spread over eight lanes (eight channels each); returns xhat * gamma + beta
__device__ __forceinline__ void ln8(float4& v0, float4& v1, const float4& g0, const float4& g1, const float4& b0, const float4& b1,
                                    float eps) {
  float s = (v0.x + v0.y) + (v0.z + v0.w) + (v1.x + v1.y) + (v1.z + v1.w);
#pragma unroll
  for (int o = 1; o < 8; o <<= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  const float mean = s * (1.0f / 64.0f);
  v0.x -= mean; v0.y -= mean; v0.z -= mean; v0.w -= mean; v1.x -= mean; v1.y -= mean; v1.z -= mean; v1.w -= mean;
  float q = v0.x * v0.x;
  q = fmaf(v0.y, v0.y, q); q = fmaf(v0.z, v0.z, q); q = fmaf(v0.w, v0.w, q);
  q = fmaf(v1.x, v1.x, q); q = fmaf(v1.y, v1.y, q); q = fmaf(v1.z, v1.z, q); q = fmaf(v1.w, v1.w, q);
#pragma unroll
  for (int o = 1; o < 8; o <<= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
  const float rstd = rsqrtf(q * (1.0f / 64.0f) + eps);
  v0.x = fmaf(v0.x * rstd, g0.x, b0.x); v0.y = fmaf(v0.y * rstd, g0.y, b0.y);
  v0.z = fmaf(v0.z * rstd, g0.z, b0.z); v0.w = fmaf(v0.w * rstd, g0.w, b0.w);
  v1.x = fmaf(v1.x * rstd, g1.x, b1.x); v1.y = fmaf(v1.y * rstd, g1.y, b1.y);
  v1.z = fmaf(v1.z * rstd, g1.z, b1.z); v1.w = fmaf(v1.w * rstd, g1.w, b1.w);
}
}  // namespace

__global__ void __launch_bounds__(kFwCta, 1) afno_fwd_tc_kernel(const AfnoWTcArgs a) {
  extern __shared__ __align__(128) uint8_t smem[];
  pdl_launch_dependents();
  __shared__ uint64_t bars[WB_N];
  __shared__ uint32_t tmem_s;
  const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;
  if (tid == 0) {
    for (int g = 0; g < kFwGroups; ++g) {
      mbar_init(&bars[WB_YFULL + g], kFwGrp);
      mbar_init(&bars[WB_DFULL + g], 1);
    }
    mbar_init(&bars[WB_CONST], 1);
    fence_mbar_init();
  }
  if (warp == kFwCtlWarp) {
    tmem_alloc(&tmem_s, 256);
    tmem_relinquish();
  }
  float* sGB = reinterpret_cast<float*>(smem + FwdSmem::GB);      // norm1 weight [64], bias [64]
  if (tid < kAfnoE) {
    sGB[tid] = a.gamma[tid];
    sGB[kAfnoE + tid] = a.beta[tid];
  }
  const bool embed = a.ngrp > 0;      // first block: the tokens are computed from the step input, not read
  float* sWe = reinterpret_cast<float*>(smem + FwdSmem::Emb);     // [Cin][64] embedding weights, transposed
  float* sBe = sWe + 16 * kAfnoE;                                 // [64] embedding bias
  if (embed) {
    for (int i = tid; i < a.Cin * kAfnoE; i += kFwCta) sWe[i] = a.We[(i & 63) * a.Cin + (i >> 6)];
    if (tid < kAfnoE) sBe[tid] = a.be[tid];
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint64_t* b = bars;
  const int stride = gridDim.x, first = blockIdx.x;

  if (warp == kFwCtlWarp) {
    const uint32_t leader = elect_one();
    mbar_arrive_expect_tx_p(leader, &b[WB_CONST], 65536u);
    bulk_g2s_p(leader, smem + FwdSmem::F, a.Fimg, 65536u, &b[WB_CONST]);
    mbar_wait(&b[WB_CONST], 0);
    const uint32_t idesc = idesc_bf16(128, 64) | kIdescBMnMajor;
    const uint64_t dF = chunk_desc(smem_u32(smem + FwdSmem::F), 128, 0);
    uint32_t t = 0;      // tiles of this CTA in order: tile t belongs to group t % kFwGroups
    for (int tile = first; tile < a.ntiles; tile += stride, ++t) {
      const int g = t % kFwGroups;
      mbar_wait(&b[WB_YFULL + g], (t / kFwGroups) & 1u);      // the group has drained its previous accumulator before it arrives here
      if (lane == 0) TRA(t, 6);
      tc_fence_after();
      // B: 8 x 8 core matrices 128 B apart along tokens (K), 2048 B along channels (N)
      const uint64_t dY = smem_desc(smem_u32(smem + FwdSmem::Y + g * 32768), 128, 2048);
      const uint32_t d = tmem_s + 64u * g;
#pragma unroll
      for (int term = 0; term < 3; ++term) {      // hi*hi, lo*hi, hi*lo
        const uint32_t fa = term == 1 ? 32768u : 0u, yp = term == 2 ? 16384u : 0u;
#pragma unroll
        for (int k0 = 0; k0 < 128; k0 += 16)
          umma_bf16_p(leader, d, desc_off(dF, fa + (k0 >> 3) * 2048), desc_off(dY, yp + k0 * 16), idesc, (term | k0) != 0);
      }
      umma_commit_p(leader, &b[WB_DFULL + g]);
      if (lane == 0) TRA(t, 7);
      __syncwarp();
    }
  } else {
    const int g = warp >> 2, q4 = warp & 3, row = q4 * 32 + lane;      // group, TMEM lane quarter, token of the tile = accumulator row
    const int rows_used = a.RT * 2 * a.Kc;
    const int HW = a.H * a.W;
    // Token loads and LayerNorm: a warp takes 8 tokens per step (4 steps per tile), lane = token t + 8 og; the thread holds
    // channel octets og and og + 4 of its token, so a load instruction touches 8 lines of 128 bytes (one thread per token
    // touched 32, and the L1 data pipe is what bounds this kernel); statistics over the four lanes of a token by shuffles.
    const int lt = q4 * 8 + (lane & 7), og = lane >> 3;      // token of step 0, octet pair
    uint8_t* ybase = smem + FwdSmem::Y + g * 32768 + og * 2048;

    auto load_x = [&](int tile, float4 (&x)[16]) {
#pragma unroll
      for (int st = 0; st < 4; ++st) {
        const long long p0 = (long long)tile * kTcTile + st * 32 + lt;
        const float4* src;
        if (!embed) {
          src = reinterpret_cast<const float4*>(a.x + (size_t)p0 * kAfnoE);
        } else {
          // first block: x = patch embedding + position embedding of the step input (fourcastnet.py:285-289,294-307);
          // here only the position embedding row
          const long long bb = p0 / HW, pix = p0 - bb * HW;
          src = reinterpret_cast<const float4*>(a.pos + pix * kAfnoE);
        }
        if (!embed || a.pos != nullptr) {
          x[4 * st] = src[2 * og];
          x[4 * st + 1] = src[2 * og + 1];
          x[4 * st + 2] = src[8 + 2 * og];
          x[4 * st + 3] = src[8 + 2 * og + 1];
        } else {
          x[4 * st] = x[4 * st + 1] = x[4 * st + 2] = x[4 * st + 3] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
      }
    };
    // first block: the input channels of the group's next tile travel global -> shared memory asynchronously (no registers
    // held over the drain); lane (t, og) copies channels og and og + 4 of its tokens, the four lanes of a token read all
    float* sIn = reinterpret_cast<float*>(smem + FwdSmem::In) + g * 8 * 128;
    auto issue_in = [&](int tile) {
#pragma unroll
      for (int st = 0; st < 4; ++st) {
        const long long p0 = (long long)tile * kTcTile + st * 32 + lt;
        const long long bb = p0 / HW, pix = p0 - bb * HW;
#pragma unroll
        for (int i = 0; i < 2; ++i) {
          const int k = og + 4 * i;
          int ch = k;
          const float* src = nullptr;
          if (k < a.Cin) {
#pragma unroll
            for (int gi = 0; gi < 4; ++gi)
              if (gi < a.ngrp) {
                if (ch >= 0 && ch < a.grp[gi].nch) src = a.grp[gi].base + bb * a.grp[gi].bstride + (long long)ch * HW + pix;
                ch -= a.grp[gi].nch;
              }
          }
          if (src != nullptr)
            asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(sIn + k * 128 + st * 32 + lt)), "l"(src) : "memory");
        }
      }
      asm volatile("cp.async.commit_group;" ::: "memory");
    };

    pdl_wait();   // x / the step input are what earlier kernels wrote
    float4 x[16];
    const int tile0 = first + g * stride;
    if (tile0 < a.ntiles) {
      load_x(tile0, x);
      if (embed) issue_in(tile0);
    }
    uint32_t u = 0;      // this group's tiles
    for (int tile = tile0; tile < a.ntiles; tile += kFwGroups * stride, ++u) {
      if (tid == 0) TRA(u, 0);
      if (embed) {
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        __syncwarp();
      }
      // the group's image is free: its previous tile's accumulator has been drained, so those MMAs are complete
#pragma unroll
      for (int st = 0; st < 4; ++st) {
        float4* xs = x + 4 * st;
        if (embed) {
#pragma unroll
          for (int k = 0; k < 8; ++k)
            if (k < a.Cin) {
              const float xk = sIn[k * 128 + st * 32 + lt];
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                const float4 w = *reinterpret_cast<const float4*>(sWe + k * kAfnoE + (j >> 1) * 32 + og * 8 + (j & 1) * 4);
                xs[j].x = fmaf(xk, w.x, xs[j].x); xs[j].y = fmaf(xk, w.y, xs[j].y);
                xs[j].z = fmaf(xk, w.z, xs[j].z); xs[j].w = fmaf(xk, w.w, xs[j].w);
              }
            }
          float4* xo = reinterpret_cast<float4*>(a.xout + ((size_t)tile * kTcTile + st * 32 + lt) * kAfnoE);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const int c4 = (j >> 1) * 8 + og * 2 + (j & 1);      // float4 index within the token
            const float4 bq = *reinterpret_cast<const float4*>(sBe + 4 * c4);
            xs[j] = make_float4(xs[j].x + bq.x, xs[j].y + bq.y, xs[j].z + bq.z, xs[j].w + bq.w);
            xo[c4] = xs[j];
          }
        }
        // LayerNorm of the token, two passes
        float sum = 0.f;
#pragma unroll
        for (int j = 0; j < 4; ++j) sum += (xs[j].x + xs[j].y) + (xs[j].z + xs[j].w);
        sum += __shfl_xor_sync(0xffffffffu, sum, 8);
        sum += __shfl_xor_sync(0xffffffffu, sum, 16);
        const float mean = sum * (1.0f / 64.0f);
        float q = 0.f;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float d0 = xs[j].x - mean, d1 = xs[j].y - mean, d2 = xs[j].z - mean, d3 = xs[j].w - mean;
          q = fmaf(d0, d0, q); q = fmaf(d1, d1, q); q = fmaf(d2, d2, q); q = fmaf(d3, d3, q);
        }
        q += __shfl_xor_sync(0xffffffffu, q, 8);
        q += __shfl_xor_sync(0xffffffffu, q, 16);
        const float rstd = rsqrtf(q * (1.0f / 64.0f) + a.eps);
#pragma unroll
        for (int o = 0; o < 2; ++o) {     // octets og and og + 4: 8 channels = one 16-byte unit of the chunk-major image
          const int c = 32 * o + 8 * og;
          const float4 ga = *reinterpret_cast<const float4*>(sGB + c), gb = *reinterpret_cast<const float4*>(sGB + c + 4);
          const float4 ba = *reinterpret_cast<const float4*>(sGB + kAfnoE + c), bb = *reinterpret_cast<const float4*>(sGB + kAfnoE + c + 4);
          const float4 v0 = xs[2 * o], v1 = xs[2 * o + 1];
          uint32_t h[4], l[4];
          split2p(make_float2(fmaf((v0.x - mean) * rstd, ga.x, ba.x), fmaf((v0.y - mean) * rstd, ga.y, ba.y)), h[0], l[0]);
          split2p(make_float2(fmaf((v0.z - mean) * rstd, ga.z, ba.z), fmaf((v0.w - mean) * rstd, ga.w, ba.w)), h[1], l[1]);
          split2p(make_float2(fmaf((v1.x - mean) * rstd, gb.x, bb.x), fmaf((v1.y - mean) * rstd, gb.y, bb.y)), h[2], l[2]);
          split2p(make_float2(fmaf((v1.z - mean) * rstd, gb.z, bb.z), fmaf((v1.w - mean) * rstd, gb.w, bb.w)), h[3], l[3]);
          uint8_t* yd = ybase + o * 8192 + (st * 32 + lt) * 16;
          *reinterpret_cast<uint4*>(yd) = make_uint4(h[0], h[1], h[2], h[3]);
          *reinterpret_cast<uint4*>(yd + 16384) = make_uint4(l[0], l[1], l[2], l[3]);
        }
      }
      fence_async_smem();
      mbar_arrive(&b[WB_YFULL + g]);
      if (tid == 0) TRA(u, 3);
      // the next tile's loads are in flight while this tile's MMAs run and its accumulator is drained
      const int next = tile + kFwGroups * stride;
      if (next < a.ntiles) {
        load_x(next, x);
        if (embed) {
          __syncwarp();      // every lane of the warp has read this tile's channels
          issue_in(next);
        }
      }
      mbar_wait(&b[WB_DFULL + g], u & 1u);
      if (tid == 0) TRA(u, 4);
      tc_fence_after();
      {
        const int h2 = row / (2 * a.Kc), rem = row - h2 * 2 * a.Kc, kx = rem >> 1, reim = rem & 1;
        const int grow = tile * a.RT + h2, bb = grow / a.H, h = grow - bb * a.H;
        float4* dst = reinterpret_cast<float4*>(a.A + ((((size_t)bb * a.Kc + kx) * a.H + h) * 2 + reim) * kAfnoE);
        // operand-image form (afno_mid_tc_kernel): K row k = 2h + (re|im) of the column's image, 16 bytes per channel octet;
        // the (re, im) rows of a mode sit on adjacent lanes: whole 32-byte sectors per store
        const uint32_t ostride = (uint32_t)a.H * 32u;      // 2H K rows x 16 bytes
        uint8_t* idst = a.Aimg + ((size_t)bb * a.Kc + kx) * (8u * ostride) + (size_t)(2 * h + reim) * 16;
        const size_t ilo = (size_t)a.ncols * (8u * ostride);
#pragma unroll
        for (int qc = 0; qc < 4; ++qc) {      // 16 columns at a time: the next tile's tokens are live in registers
          uint32_t r[16];
          tmem_ld16(tmem_s + ((uint32_t)(q4 * 32) << 16) + 64u * g + 16u * qc, r);
          tmem_ld_wait();
          if (a.Aimg != nullptr) {
            if (row < rows_used) {
#pragma unroll
              for (int o2 = 0; o2 < 2; ++o2) {
                uint32_t hi[4], lo[4];
#pragma unroll
                for (int e = 0; e < 4; ++e)
                  split2p(make_float2(__uint_as_float(r[8 * o2 + 2 * e]), __uint_as_float(r[8 * o2 + 2 * e + 1])), hi[e], lo[e]);
                uint8_t* d = idst + (size_t)(2 * qc + o2) * ostride;
                *reinterpret_cast<uint4*>(d) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
                *reinterpret_cast<uint4*>(d + ilo) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
              }
            }
          } else if (row < rows_used) {
#pragma unroll
            for (int j = 0; j < 4; ++j)
              dst[4 * qc + j] = make_float4(__uint_as_float(r[4 * j]), __uint_as_float(r[4 * j + 1]), __uint_as_float(r[4 * j + 2]),
                                            __uint_as_float(r[4 * j + 3]));
          }
        }
      }
      tc_fence_before();
      if (tid == 0) TRA(u, 5);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == kFwCtlWarp) tmem_dealloc(tmem_s, 256);
}

int launch_afno_fwd_tc(const AfnoWTcArgs& a, cudaStream_t st) {
  int sms = 0;
  DLWP_TRY(current_sms(&sms));
  DLWP_TRY(ensure_dynamic_smem((const void*)afno_fwd_tc_kernel, FwdSmem::Total));
  DLWP_REQUIRE(a.ntiles > 0 && a.RT * a.W == kTcTile && a.RT * 2 * a.Kc <= 128, "afno_fwd_tc_kernel: unsupported shape");
  DLWP_REQUIRE(a.ngrp == 0 || (a.Cin >= 1 && a.Cin <= 8 && a.We && a.be && a.xout), "afno_fwd_tc_kernel: bad embedding arguments");
  const int grid = a.ntiles < sms ? a.ntiles : sms;
  DLWP_CUDA(launch_pdl(afno_fwd_tc_kernel, dim3(grid), dim3(kFwCta), (size_t)FwdSmem::Total, st, a));
  DLWP_LAUNCHED("afno_fwd_tc_kernel");
  return DLWP_OK;
}

// =============================================================================================
// afno_inv_tc_kernel:  out[px, c] = sum_k G'[px, k] * Z'[c, k] + LayerNorm(x)[px, c] + x[px, c],  k = (h2, kx, re | im)
//   The tile's rows of Z (complex fp32, after the inverse H-DFT) arrive by one bulk copy and are split into
//   the K-major bf16 operand image Z' by the workers; the accumulator is staged row-padded in shared memory
//   so that the second phase can work token-major (four tokens per warp step as in the forward kernel):
//   LayerNorm(x) is recomputed from the token tile (bulk-copied, never stored normalised), the sum is
//   written with fully coalesced 1 KB warp stores.  A tile is finished one iteration late, under the
//   tensor-core work of the next one.
// =============================================================================================
namespace {
enum { IB_ZRFULL = 0, IB_ZREMPTY = 2, IB_ZIFULL = 4, IB_ZIEMPTY = 6, IB_XFULL = 8, IB_XEMPTY = 10, IB_DFULL = 12, IB_DEMPTY = 14,
       IB_CONST = 16, IB_N = 17 };
constexpr int kDstRow = 68;      // floats per staged accumulator row (64 + 4: conflict-free 16-byte accesses)
struct InvSmem {
  int G, X, Zraw, Zimg, Dst, zraw_bytes, zimg_bytes, Total;
  __host__ __device__ InvSmem(int RT, int Kc, int KZ) {
    zraw_bytes = RT * Kc * kAfnoE * 8;
    zimg_bytes = 2 * kAfnoE * KZ * 2;
    G = 0;
    X = G + 2 * 128 * KZ * 2;
    Zraw = X + 2 * 32768;
    Zimg = Zraw + 2 * zraw_bytes;
    Dst = Zimg + 2 * zimg_bytes;
    Total = Dst + 128 * kDstRow * 4;
  }
};
}  // namespace

__global__ void __launch_bounds__(kWtCta, 1) afno_inv_tc_kernel(const AfnoWTcArgs a) {
  extern __shared__ __align__(128) uint8_t smem[];
  pdl_launch_dependents();
  __shared__ uint64_t bars[IB_N];
  __shared__ uint32_t tmem_s;
  const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;
  const InvSmem L(a.RT, a.Kc, a.KZ);
  if (tid == 0) {
    for (int i = 0; i < 2; ++i) {
      mbar_init(&bars[IB_ZRFULL + i], 1);
      mbar_init(&bars[IB_ZREMPTY + i], kWtWork);
      mbar_init(&bars[IB_ZIFULL + i], kWtWork);
      mbar_init(&bars[IB_ZIEMPTY + i], 1);
      mbar_init(&bars[IB_XFULL + i], 1);
      mbar_init(&bars[IB_XEMPTY + i], kWtWork);
      mbar_init(&bars[IB_DFULL + i], 1);
      mbar_init(&bars[IB_DEMPTY + i], kWtWork);
    }
    mbar_init(&bars[IB_CONST], 1);
    fence_mbar_init();
  }
  if (warp == kWtCtlWarp) {
    tmem_alloc(&tmem_s, 128);
    tmem_relinquish();
  }
  // the K slots that pad the operand to a multiple of 16 stay zero in both Z' buffers
  for (int i = tid; i < 2 * L.zimg_bytes / 16; i += kWtCta) reinterpret_cast<uint4*>(smem + L.Zimg)[i] = make_uint4(0, 0, 0, 0);
  fence_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint64_t* b = bars;
  const int stride = gridDim.x, first = blockIdx.x;
  const uint32_t g_bytes = 2u * 128u * (uint32_t)a.KZ * 2u;

  if (warp == kWtCtlWarp) {
    const uint32_t leader = elect_one();
    mbar_arrive_expect_tx_p(leader, &b[IB_CONST], g_bytes);
    bulk_g2s_p(leader, smem + L.G, a.Gimg, g_bytes, &b[IB_CONST]);
    pdl_wait();   // Z and x are what the previous kernels wrote
    auto load_z = [&](int tile, int s) {
      mbar_arrive_expect_tx_p(leader, &b[IB_ZRFULL + s], (uint32_t)L.zraw_bytes);
      bulk_g2s_p(leader, smem + L.Zraw + s * L.zraw_bytes, a.Z + (size_t)tile * a.RT * a.Kc * kAfnoE, (uint32_t)L.zraw_bytes,
                 &b[IB_ZRFULL + s]);
    };
    auto load_x = [&](int tile, int s) {
      mbar_arrive_expect_tx_p(leader, &b[IB_XFULL + s], 32768u);
      bulk_g2s_p(leader, smem + L.X + s * 32768, a.x + (size_t)tile * kTcTile * kAfnoE, 32768u, &b[IB_XFULL + s]);
    };
    // ready operand images (afno_mid_tc_kernel): straight into the operand buffers, no conversion by the workers
    const bool zimg = a.Zimg != nullptr;
    auto load_zi = [&](int tile, int s) {
      mbar_arrive_expect_tx_p(leader, &b[IB_ZRFULL + s], (uint32_t)L.zimg_bytes);
      bulk_g2s_p(leader, smem + L.Zimg + s * L.zimg_bytes, a.Zimg + (size_t)tile * L.zimg_bytes, (uint32_t)L.zimg_bytes, &b[IB_ZRFULL + s]);
    };
    for (int i = 0; i < 2; ++i)
      if (first + i * stride < a.ntiles) {
        if (zimg) load_zi(first + i * stride, i);
        else load_z(first + i * stride, i);
        load_x(first + i * stride, i);
      }
    mbar_wait(&b[IB_CONST], 0);
    const uint32_t idesc = idesc_bf16(128, 64) | (zimg ? kIdescBMnMajor : 0u);
    const uint64_t dG = chunk_desc(smem_u32(smem + L.G), 128, 0);
    const uint32_t g_lo = 128u * (uint32_t)a.KZ * 2u, z_lo = (uint32_t)kAfnoE * (uint32_t)a.KZ * 2u;
    const uint32_t zstep = zimg ? 256u : 2048u;      // bytes per K step of 16: MN-major K rows of 16 bytes / K-major chunks of 64 rows
    uint32_t u = 0;
    for (int tile = first; tile < a.ntiles; tile += stride, ++u) {
      const int s = u & 1;
      const uint32_t ph = (u >> 1) & 1u;
      mbar_wait(&b[(zimg ? IB_ZRFULL : IB_ZIFULL) + s], ph);
      mbar_wait(&b[IB_DEMPTY + s], ph ^ 1u);
      tc_fence_after();
      const uint32_t zaddr = smem_u32(smem + L.Zimg + s * L.zimg_bytes);
      const uint64_t dZ = zimg ? smem_desc(zaddr, 128, (uint32_t)a.KZ * 16u) : chunk_desc(zaddr, kAfnoE, 0);
      const uint32_t d = tmem_s + 64u * s;
      uint32_t acc = 0;
#pragma unroll
      for (int term = 0; term < 3; ++term) {      // hi*hi, lo*hi, hi*lo
        const uint32_t ga = term == 1 ? g_lo : 0u, zb = term == 2 ? z_lo : 0u;
#pragma unroll
        for (int k0 = 0; k0 < 128; k0 += 16) {      // unrolled with a uniform guard (KZ <= 128: RT * 2 * Kc <= 128)
          if (k0 < a.KZ) {
            umma_bf16_p(leader, d, desc_off(dG, ga + (k0 >> 3) * 2048), desc_off(dZ, zb + (k0 >> 4) * zstep), idesc, acc);
            acc = 1;
          }
        }
      }
      umma_commit_p(leader, &b[IB_DFULL + s]);
      umma_commit_p(leader, &b[IB_ZIEMPTY + s]);
      if (zimg) {      // the other operand buffer is free when the previous tile's MMAs have completed
        if (u >= 1 && tile + stride < a.ntiles) {
          mbar_wait(&b[IB_ZIEMPTY + (s ^ 1)], ((u - 1) >> 1) & 1u);
          load_zi(tile + stride, s ^ 1);
        }
      } else if (tile + 2 * stride < a.ntiles) {      // the raw Z buffer was released when the workers had converted it
        mbar_wait(&b[IB_ZREMPTY + s], ph);
        load_z(tile + 2 * stride, s);
      }
      if (u >= 1 && tile + stride < a.ntiles) {   // the token buffer of tile u - 1 is released when that tile has been finished
        mbar_wait(&b[IB_XEMPTY + (s ^ 1)], ((u - 1) >> 1) & 1u);
        load_x(tile + stride, s ^ 1);
      }
      __syncwarp();
    }
  } else {
    // phase B: token of the step; the lane's channels 4 o .. 4 o + 3 and 32 + 4 o .. 32 + 4 o + 3 (the 8 lanes of a token read
    // 128 contiguous bytes per access: no bank conflicts, 4 lines per global store instruction)
    const int t4 = lane >> 3, o = lane & 7;
    const float4 g0 = __ldg(reinterpret_cast<const float4*>(a.gamma) + o), g1 = __ldg(reinterpret_cast<const float4*>(a.gamma) + 8 + o);
    const float4 e0 = __ldg(reinterpret_cast<const float4*>(a.beta) + o), e1 = __ldg(reinterpret_cast<const float4*>(a.beta) + 8 + o);
    const int q4 = warp & 3, half = warp >> 2, row = q4 * 32 + lane;     // phase A: accumulator row, column half
    float* dstg = reinterpret_cast<float*>(smem + L.Dst);
    const int n_el = a.RT * a.Kc * kAfnoE;

    auto finish = [&](int tile, uint32_t u) {
      const int s = u & 1;
      const uint32_t ph = (u >> 1) & 1u;
      mbar_wait(&b[IB_DFULL + s], ph);
      tc_fence_after();
      {
        uint32_t r[32];
        tmem_ld32(tmem_s + ((uint32_t)(q4 * 32) << 16) + 64u * s + 32u * half, r);
        tmem_ld_wait();
        tc_fence_before();
        mbar_arrive(&b[IB_DEMPTY + s]);
        float4* drow = reinterpret_cast<float4*>(dstg + row * kDstRow + 32 * half);
#pragma unroll
        for (int j = 0; j < 8; ++j)
          drow[j] = make_float4(__uint_as_float(r[4 * j]), __uint_as_float(r[4 * j + 1]), __uint_as_float(r[4 * j + 2]),
                                __uint_as_float(r[4 * j + 3]));
      }
      bar_sync(1, kWtWork);
      mbar_wait(&b[IB_XFULL + s], ph);
      const float* xs = reinterpret_cast<const float*>(smem + L.X + s * 32768);
      float* orow = a.out + (size_t)tile * kTcTile * kAfnoE;
#pragma unroll
      for (int st = 0; st < 4; ++st) {
        const int px = warp * 16 + st * 4 + t4;
        const float4 x0 = *reinterpret_cast<const float4*>(xs + px * kAfnoE + o * 4),
                     x1 = *reinterpret_cast<const float4*>(xs + px * kAfnoE + 32 + o * 4);
        const float4 d0 = *reinterpret_cast<const float4*>(dstg + px * kDstRow + o * 4),
                     d1 = *reinterpret_cast<const float4*>(dstg + px * kDstRow + 32 + o * 4);
        float4 y0 = x0, y1 = x1;
        ln8(y0, y1, g0, g1, e0, e1, a.eps);
        float4* op = reinterpret_cast<float4*>(orow + px * kAfnoE + o * 4);
        op[0] = make_float4(d0.x + y0.x + x0.x, d0.y + y0.y + x0.y, d0.z + y0.z + x0.z, d0.w + y0.w + x0.w);
        op[8] = make_float4(d1.x + y1.x + x1.x, d1.y + y1.y + x1.y, d1.z + y1.z + x1.z, d1.w + y1.w + x1.w);
      }
      mbar_arrive(&b[IB_XEMPTY + s]);
      bar_sync(1, kWtWork);      // the staging rows are rewritten by the next tile's first phase
    };

    uint32_t u = 0;
    int prev = -1;
    for (int tile = first; tile < a.ntiles; tile += stride, ++u) {
      const int s = u & 1;
      const uint32_t ph = (u >> 1) & 1u;
      if (a.Zimg == nullptr) {
        mbar_wait(&b[IB_ZRFULL + s], ph);
        mbar_wait(&b[IB_ZIEMPTY + s], ph ^ 1u);
        const float2* zr = reinterpret_cast<const float2*>(smem + L.Zraw + s * L.zraw_bytes);
        uint8_t* zi = smem + L.Zimg + s * L.zimg_bytes;
        const uint32_t lo_off = (uint32_t)kAfnoE * (uint32_t)a.KZ * 2u;
        for (int i = tid; i < n_el; i += kWtWork) {      // (grid row of the tile, column) x channel -> K slots (re, im) of that channel's row
          const int hk = i >> 6, c = i & 63, k = 2 * hk;
          uint32_t hi, lo;
          split2p(zr[i], hi, lo);
          uint8_t* dst = zi + (k >> 3) * (kAfnoE * 16) + c * 16 + (k & 7) * 2;
          *reinterpret_cast<uint32_t*>(dst) = hi;
          *reinterpret_cast<uint32_t*>(dst + lo_off) = lo;
        }
        mbar_arrive(&b[IB_ZREMPTY + s]);
        fence_async_smem();
        mbar_arrive(&b[IB_ZIFULL + s]);
      }
      if (prev >= 0) finish(prev, u - 1);
      prev = tile;
    }
    if (prev >= 0) finish(prev, u - 1);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == kWtCtlWarp) tmem_dealloc(tmem_s, 128);
}

int afno_inv_tc_smem_bytes(int RT, int Kc, int KZ) { return InvSmem(RT, Kc, KZ).Total; }

int launch_afno_inv_tc(const AfnoWTcArgs& a, cudaStream_t st) {
  int sms = 0;
  DLWP_TRY(current_sms(&sms));
  DLWP_TRY(ensure_dynamic_smem((const void*)afno_inv_tc_kernel, 226 * 1024));
  const InvSmem L(a.RT, a.Kc, a.KZ);
  DLWP_REQUIRE(a.ntiles > 0 && a.RT * a.W == kTcTile && a.RT * 2 * a.Kc <= a.KZ && a.KZ % 16 == 0 && a.KZ <= 128 && L.Total <= 226 * 1024,
               "afno_inv_tc_kernel: unsupported shape");
  const int grid = a.ntiles < sms ? a.ntiles : sms;
  DLWP_CUDA(launch_pdl(afno_inv_tc_kernel, dim3(grid), dim3(kWtCta), (size_t)L.Total, st, a));
  DLWP_LAUNCHED("afno_inv_tc_kernel");
  return DLWP_OK;
}

}  // namespace dlwp

#ifdef DLWP_TRACE
extern "C" int dlwp_debug_set_trace_afno(unsigned long long* p) {
  return cudaMemcpyToSymbol(dlwp::g_trace_afno, &p, sizeof(p)) == cudaSuccess ? 0 : -2;
}
#endif
