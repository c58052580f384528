// tcgen05 kernels of the AFNO block (see tc_afno.cuh).
#include "tc_afno.cuh"

#include "tc_mlp.cuh"

namespace dlwp {

using namespace tc;

// =============================================================================================
// afno_mlp_kernel: out = x + fc2(GELU(fc1(LN(x))))  on 128-token tiles, persistent, one CTA per SM
//   warps 0-7 : epilogue group (two threads per token: channel halves)     warp 8 : control warp
// Both 64 KB weight operands ([hi | lo] split-bf16 images) stay in shared memory for the whole
// kernel, which leaves room for one worker per CTA: one A1 buffer (normalised tokens, 32 KB; the
// next tile's tokens are stored after the last first-layer MMA of this tile has read it, two passes
// before they are needed) and a ring of four A2 buffers (32-column hidden chunks, 16 KB each), so the
// epilogue group never waits for the second-layer MMAs of the chunks it has just handed over; the hidden
// accumulator D1 [128 x 64] and the output accumulator D2 [128 x 64] are double-buffered in tensor memory.
// LayerNorm: each thread holds half a token (32 channels), computes its mean and centred sum of
// squares, and the two halves are combined with the pairwise update (exact like two passes).
// =============================================================================================
namespace {
constexpr int kAmEpi = 256, kAmCta = kAmEpi + 32, kAmCtlWarp = kAmEpi / 32;
struct AmSmem {
  static constexpr int W1 = 0, W2 = W1 + kAfnoHid * kAfnoE * 4, A1 = W2 + kAfnoE * kAfnoHid * 4, A2 = A1 + 32768,
                       Stats = A2 + 4 * kTileBytes, Total = Stats + 2 * 128 * 8;
  static_assert(Total <= 226 * 1024, "afno_mlp_kernel shared memory");
};
}  // namespace

__global__ void __launch_bounds__(kAmCta, 1) afno_mlp_kernel(const AfnoMlpArgs a) {
  extern __shared__ __align__(128) uint8_t smem[];
  pdl_launch_dependents();
  __shared__ uint64_t bars[MB_PER + 1];
  __shared__ uint32_t tmem_s;
  const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0);   // provably warp-uniform
  constexpr int hid = kAfnoHid, P = kAfnoHid / 64;
  uint64_t* bar_const = &bars[MB_PER];

  if (tid == 0) {
    mlp_init_barriers(bars, kAmEpi);
    mbar_init(bar_const, 1);
    fence_mbar_init();
  }
  if (warp == kAmCtlWarp) {
    tmem_alloc(&tmem_s, 256);
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
  const int stride = gridDim.x, first = blockIdx.x;
  uint64_t* b = bars;

  if (warp == kAmCtlWarp) {
    const uint32_t leader = elect_one();
    MlpIssuer is{b, smem_u32(smem + AmSmem::A1), 32768u, smem_u32(smem + AmSmem::W1), smem_u32(smem + AmSmem::W2),
                 smem_u32(smem + AmSmem::A2), tmem_s, hid, P, 0u, 0u, 0u, leader, 1u};
    mbar_wait(bar_const, 0);
    if (first < a.ntiles) is.mma1<kAfnoE>(0, 0, 0);
    uint32_t u = 0;
    for (int tile = first; tile < a.ntiles; tile += stride, ++u) {
      const bool has_next = tile + stride < a.ntiles;
      for (int p = 0; p < P; ++p) {
        const uint32_t gp = u * P + p;
        if (p + 1 < P) is.mma1<kAfnoE>(u, p + 1, gp + 1);
        else if (has_next) is.mma1<kAfnoE>(u + 1, 0, gp + 1);
        is.mma2<kAfnoE, 4>(u, p);
      }
    }
  } else {
    const EpiId id(tid);
    const uint32_t tmem_row = tmem_s + ((uint32_t)(id.q4 * 32) << 16), tmem_lane = tmem_row + 16u * id.half;
    float2* stats = reinterpret_cast<float2*>(smem + AmSmem::Stats);   // [half][row] = (mean, centred sum of squares)
    const int c0 = 32 * id.half;      // this thread's 32 channels of its token

    auto gather = [&](int tile, float4 (&x)[8]) {
      const float4* src = reinterpret_cast<const float4*>(a.x + ((size_t)tile * kTcTile + id.row) * kAfnoE + c0);
#pragma unroll
      for (int j = 0; j < 8; ++j) x[j] = src[j];
    };
    auto store_a1 = [&](uint32_t u, float4 (&x)[8]) {
      // the statistics are computed here, not where the loads were issued (the loads' latency hides behind the passes in between)
#pragma unroll
      for (int j = 0; j < 8; ++j) asm volatile("" : "+f"(x[j].x), "+f"(x[j].y), "+f"(x[j].z), "+f"(x[j].w));
      float s = 0.f;
#pragma unroll
      for (int j = 0; j < 8; ++j) s += (x[j].x + x[j].y) + (x[j].z + x[j].w);
      const float mh = s * (1.0f / 32.0f);
      float q = 0.f;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float d0 = x[j].x - mh, d1 = x[j].y - mh, d2 = x[j].z - mh, d3 = x[j].w - mh;
        q = fmaf(d0, d0, q); q = fmaf(d1, d1, q); q = fmaf(d2, d2, q); q = fmaf(d3, d3, q);
      }
      stats[id.half * 128 + id.row] = make_float2(mh, q);
      bar_sync(1, kAmEpi);
      const float2 o = stats[(id.half ^ 1) * 128 + id.row];
      const float mean = 0.5f * (mh + o.x), dm = mh - o.x;
      const float var = (q + o.y + 16.0f * dm * dm) * (1.0f / 64.0f);
      const float rstd = rsqrtf(var + a.eps);
      mbar_wait(&b[MB_A1EMPTY], (u & 1u) ^ 1u);      // ring of depth 1
      uint8_t* d = smem + AmSmem::A1 + (4 * id.half) * 2048 + id.row * 16;
#pragma unroll
      for (int j = 0; j < 4; ++j) {     // 8 channels = one 16-byte unit of the chunk-major image
        uint32_t h[4], l[4];
        const float4 v0 = x[2 * j], v1 = x[2 * j + 1];
        split2p(make_float2((v0.x - mean) * rstd, (v0.y - mean) * rstd), h[0], l[0]);
        split2p(make_float2((v0.z - mean) * rstd, (v0.w - mean) * rstd), h[1], l[1]);
        split2p(make_float2((v1.x - mean) * rstd, (v1.y - mean) * rstd), h[2], l[2]);
        split2p(make_float2((v1.z - mean) * rstd, (v1.w - mean) * rstd), h[3], l[3]);
        *reinterpret_cast<uint4*>(d + j * 2048) = make_uint4(h[0], h[1], h[2], h[3]);
        *reinterpret_cast<uint4*>(d + 16384 + j * 2048) = make_uint4(l[0], l[1], l[2], l[3]);
      }
      fence_async_smem();
      mbar_arrive(&b[MB_A1FULL]);
    };

    float4 b2v[8];      // this thread's 32 output biases
#pragma unroll
    for (int j = 0; j < 8; ++j) b2v[j] = __ldg(reinterpret_cast<const float4*>(a.b2 + c0) + j);
    float4 nx[8];
    if (first < a.ntiles) {
      gather(first, nx);
      store_a1(0, nx);
    }
    uint32_t u = 0, gp = 0, nc = 0;
    for (int tile = first; tile < a.ntiles; tile += stride, ++u) {
      const bool has_next = tile + stride < a.ntiles;
      if (has_next) gather(tile + stride, nx);      // loads in flight while the first passes are processed
      epi_mlp_tile<4>(b, id, tmem_lane, P, smem + AmSmem::A2, gp, nc, [&] {
        if (has_next) store_a1(u + 1, nx);
      }, a.b1, P - 2);
      // out = x + D2 + b2 on this thread's 32 channels (x is re-read: an L2 hit)
      const size_t off = ((size_t)tile * kTcTile + id.row) * kAfnoE + c0;
      const float4* xr = reinterpret_cast<const float4*>(a.x + off);
      float4 res[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) res[j] = xr[j];
      mbar_wait(&b[MB_D2FULL + (u & 1)], full_par(u));
      tc_fence_after();
      uint32_t r[32];
      tmem_ld32(tmem_row + 128u + 64u * (u & 1) + (uint32_t)c0, r);
      tmem_ld_wait();
      tc_fence_before();
      mbar_arrive(&b[MB_D2EMPTY + (u & 1)]);
      float4 v[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float4 bb = b2v[j];
        v[j] = make_float4(__uint_as_float(r[4 * j]) + bb.x + res[j].x, __uint_as_float(r[4 * j + 1]) + bb.y + res[j].y,
                           __uint_as_float(r[4 * j + 2]) + bb.z + res[j].z, __uint_as_float(r[4 * j + 3]) + bb.w + res[j].w);
      }
      if (a.Wh == nullptr) {
        float4* o = reinterpret_cast<float4*>(a.out + off);
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = v[j];
      } else {
        // fused head (fourcastnet.py:349-358, 1x1 patches): each thread's half of the 64-channel dot products, the upper
        // halves meet the lower ones in shared memory, the lower threads add the residual frame and write NCHW
        float part[4];
#pragma unroll
        for (int co = 0; co < 4; ++co) {
          part[co] = 0.f;
          if (co < a.Cg) {
            const float4* wq = reinterpret_cast<const float4*>(a.Wh + co * kAfnoE + c0);
            float acc = 0.f;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const float4 w = __ldg(wq + j);
              acc = fmaf(v[j].x, w.x, acc); acc = fmaf(v[j].y, w.y, acc); acc = fmaf(v[j].z, w.z, acc); acc = fmaf(v[j].w, w.w, acc);
            }
            part[co] = acc;
          }
        }
        float* sc = reinterpret_cast<float*>(stats);     // [4][128]; the statistics of the next tile were consumed two passes ago
        if (id.half == 1) {
#pragma unroll
          for (int co = 0; co < 4; ++co) sc[co * 128 + id.row] = part[co];
        }
        bar_sync(1, kAmEpi);
        if (id.half == 0) {
          const long long p = (long long)tile * kTcTile + id.row;
          const long long bb = p / a.HW, pix = p - bb * a.HW;
#pragma unroll
          for (int co = 0; co < 4; ++co)
            if (co < a.Cg) {
              const float lastv = a.last != nullptr ? a.last[bb * a.last_bstride + (long long)co * a.HW + pix] : 0.f;
              a.frame[bb * a.frame_bstride + (long long)co * a.HW + pix] = part[co] + sc[co * 128 + id.row] + lastv;
            }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == kAmCtlWarp) tmem_dealloc(tmem_s, 256);
}

int launch_afno_mlp(const AfnoMlpArgs& a, cudaStream_t st) {
  static int sms = 0;
  if (!sms) {
    int dev = 0;
    DLWP_CUDA(cudaGetDevice(&dev));
    DLWP_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    DLWP_CUDA(cudaFuncSetAttribute(afno_mlp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, AmSmem::Total));
  }
  DLWP_REQUIRE(a.ntiles > 0, "afno_mlp_kernel: empty problem");
  DLWP_REQUIRE(a.Wh == nullptr || (a.Cg >= 1 && a.Cg <= 4 && a.HW > 0 && a.frame != nullptr), "afno_mlp_kernel: bad head arguments");
  const int grid = a.ntiles < sms ? a.ntiles : sms;
  DLWP_CUDA(launch_pdl(afno_mlp_kernel, dim3(grid), dim3(kAmCta), (size_t)AmSmem::Total, st, a));
  DLWP_LAUNCHED("afno_mlp_kernel");
  return DLWP_OK;
}

}  // namespace dlwp
