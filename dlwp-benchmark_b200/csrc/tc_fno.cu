// tcgen05 kernels of the FNO / TFNO step (see tc_fno.cuh for the data layout and the split).
#include "tc_fno.cuh"

#include <cstdlib>
#include <cstring>

#include "tc_mlp.cuh"

namespace dlwp {

using namespace tc;

#ifdef DLWP_TRACE
__device__ unsigned long long* g_trace = nullptr;   // also used by tc_mid.cu
#define TR(w, u, ev) do { if (g_trace && tr_on && blockIdx.x == 0 && (u) < 16) g_trace[((w) * 16 + (u)) * 8 + (ev)] = clock64(); } while (0)
#else
#define TR(w, u, ev) do {} while (0)
#endif


// =============================================================================================
// CTA organisation shared by the three kernels (persistent, one CTA per SM, 576 threads)
//   warps 0-7  : epilogue group of worker 0      warp 16 : control warp of worker 0
//   warps 8-15 : epilogue group of worker 1      warp 17 : control warp of worker 1
// A worker owns every second tile of its CTA.  Its control warp issues the bulk copies and all
// tcgen05.mma of the worker (one lane) and never touches data.  Its 256 epilogue threads work
// in pairs on one accumulator row (= pixel): warps j and j + 4 of the group read the same TMEM
// lane quarter and take the lower / upper 16 columns of every 32-column block, move TMEM ->
// registers -> SMEM and never wait for an MMA round trip they did not have a full pass of
// work to hide behind (four epilogue warps per scheduler keep the FMA / MUFU pipes fed while
// others sit in tcgen05.ld or a barrier).  The two sides meet only on mbarriers:
//   *_full  : producer -> consumer ("data is there"): tcgen05.commit (count 1), a bulk copy's
//             complete_tx (count 1 + bytes) or 256 epilogue arrivals
//   *_empty : consumer -> producer ("buffer may be overwritten")
// Ring slot of use u on a ring of depth 2: u & 1; full parity (u >> 1) & 1; empty parity
// ((u >> 1) & 1) ^ 1 (so the first two uses pass immediately on fresh barriers).
// Constant operands (weights, DFT matrices) are loaded once per CTA and shared by both workers.
// TMEM: 512 columns per CTA, 256 per worker.
// =============================================================================================
namespace {


// ---- common tail of the lift and layer epilogues ------------------------------------------------
// v[32] = the 32 output channels of this thread's pixel (tile row t).  The epilogue group writes
// the split-bf16 tile image into a shared-memory buffer (same layout as in HBM) and hands it to
// the control lane (xt_full), which
//   * stores it to HBM with one bulk copy, and
//   * when `Aout` is set, runs the forward truncated W-DFT on it as M=64 UMMAs
//       D3[(h2, kx', re|im), c] = sum_{px} F''[(h2, kx', re|im), px] * v[px, c]   (accumulator rows 32..63 unused)
//     reading the image in place as an MN-major B operand (pixels = K rows, channels contiguous),
// then releases the buffer (y_free).  The accumulator of tile u is drained (drain_d3) while
// tile u + 1 is in the epilogue, so that UMMA round trip is never waited for: rows (Re | Im of
// one retained column of one grid row) go straight from TMEM lanes to their 128-byte rows of
// the planar spectrum  Aout[b][kx][h][re|im][32 channels]  (fp32; what fno_mid_kernel bulk-copies).
__device__ __forceinline__ void drain_d3(const EpiId& id, int tile, const TcGeom& g, float* Aout, uint32_t tmem_d3,
                                         uint64_t* d3_full, uint32_t n) {
  mbar_wait(d3_full, n & 1u);
  tc_fence_after();
  const int q4 = id.q4, lane = id.row & 31;
  if (q4 < 2) {   // an M=64 accumulator keeps row r in TMEM lane 32 * (r / 16) + r % 16
    uint32_t r[16];
    tmem_ld16(tmem_d3 + ((uint32_t)(q4 * 32) << 16) + 16u * id.half, r);
    tmem_ld_wait();
    const int row = q4 * 16 + lane;
    const int kxs = 31 - __clz(g.KXP);      // KXP = 32 / RT is a power of two
    const int h2 = row >> kxs, rem = row - (h2 << kxs), kx = rem >> 1, reim = rem & 1;
    if (lane < 16 && h2 < g.RT && kx < g.Kc) {
      const int b = tile / g.TPS, h = (tile - b * g.TPS) * g.RT + h2;
      if (g.abf) {
        // split-bf16 operand tiles of fno_mid_tc_kernel: [kx][hi | lo][b][channel group][k = 2h + (re|im)][8 channels]
        uint32_t hi[8], lo[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) split2p(make_float2(__uint_as_float(r[2 * e]), __uint_as_float(r[2 * e + 1])), hi[e], lo[e]);
        const size_t tile_b = (size_t)g.H * 128;
        uint8_t* d = reinterpret_cast<uint8_t*>(Aout) + ((size_t)(kx * 2) * g.Bn + b) * tile_b + (size_t)(2 * id.half) * (g.H * 32) +
                     (size_t)(2 * h + reim) * 16;
#pragma unroll
        for (int gq = 0; gq < 2; ++gq) {
          *reinterpret_cast<uint4*>(d + gq * (g.H * 32)) = make_uint4(hi[4 * gq], hi[4 * gq + 1], hi[4 * gq + 2], hi[4 * gq + 3]);
          *reinterpret_cast<uint4*>(d + (size_t)g.Bn * tile_b + gq * (g.H * 32)) = make_uint4(lo[4 * gq], lo[4 * gq + 1], lo[4 * gq + 2], lo[4 * gq + 3]);
        }
      } else {
      float4* dst = reinterpret_cast<float4*>(Aout + ((((size_t)b * g.Kc + kx) * g.H + h) * 2 + reim) * 32 + 16 * id.half);
#pragma unroll
      for (int j = 0; j < 4; ++j)
        dst[j] = make_float4(__uint_as_float(r[4 * j]), __uint_as_float(r[4 * j + 1]), __uint_as_float(r[4 * j + 2]),
                             __uint_as_float(r[4 * j + 3]));
      }
    }
  }
  tc_fence_before();
}

// n3 counts the tiles this worker has handed over.  v[8] = this thread's 16 output channels (16 * half ..) of its
// pixel, as pairs.  The tiles take `ny` image buffers (each with its xt_full / y_free barrier) and `nd` W-DFT
// accumulators (32 TMEM columns apart, each with its d3_full barrier) in turn:
//   nd = 1          the previous tile's accumulator is drained before this tile is handed over;
//   nd = 2 or 3     the accumulator of tile n - (nd - 1), which tile n + 1 overwrites, is drained AFTER tile n is
//                   handed over: the proxy fence of the hand-over waits for the thread's global stores in
//                   flight, so those of the drain must not be issued just ahead of it, and the chain of small
//                   W-DFT UMMAs of a tile (~3 k cycles) has nd - 1 tile periods to complete.
//   ny = 2          the epilogue of a tile does not wait for the bulk store of the tile before it either.
// (layer kernel: ny = 2, nd = 3 when the shared memory is there; lift kernel: ny = 1, nd = 2)
// prev_tile: nd = 3 keeps tile n in prev_tile[n & 1]; otherwise prev_tile[0] is the last tile.
struct TailCfg {
  int ny, nd;
  bool ext;      // the accumulators are drained by dedicated warps (layer kernel), not by the epilogue group;
                 // the W-DFT then runs as F_hi x [Y_hi | Y_lo] (N = 64) + F_lo x Y_hi (N = 32): 16 UMMAs instead of 24
                 // (shared-memory operand traffic bounds this kernel), accumulators 64 TMEM columns apart
};
__device__ __forceinline__ uint32_t d3_slot(uint32_t n, int nd) { return nd == 3 ? n % 3u : nd == 2 ? (n & 1u) : 0u; }
__device__ __forceinline__ uint32_t d3_use(uint32_t n, int nd) { return nd == 3 ? n / 3u : nd == 2 ? (n >> 1) : n; }

__device__ __forceinline__ void finish_tile(const float2 (&v)[8], const EpiId& id, int tile, const TcGeom& g, float* Aout,
                                            uint8_t* sY, uint32_t tmem_d3, uint64_t* xt_full, uint64_t* d3_full,
                                            uint64_t* y_free, uint32_t& n3, int (&prev_tile)[2], const TailCfg cfg,
                                            bool tr_on = false) {
  const int ny = cfg.ny, nd = cfg.nd;
  uint32_t hi[8], lo[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) split2p(v[j], hi[j], lo[j]);
  const uint32_t yb = ny == 2 ? (n3 & 1u) : 0u, use = ny == 2 ? (n3 >> 1) : n3;
  if (use > 0) mbar_wait(&y_free[yb], (use - 1) & 1u);   // the tile that used this image buffer last has left it (bulk store, UMMAs)
  if (!cfg.ext && nd == 1 && n3 > 0 && Aout != nullptr) drain_d3(id, prev_tile[0], g, Aout, tmem_d3, d3_full, n3 - 1);
  if ((threadIdx.x & 255) == 0) TR(id.w, n3, 7);
  uint8_t* dst = sY + yb * kTileBytes + id.row * 16 + id.half * 4096;
#pragma unroll
  for (int q = 0; q < 2; ++q) {
    *reinterpret_cast<uint4*>(dst + q * 2048) = make_uint4(hi[4 * q], hi[4 * q + 1], hi[4 * q + 2], hi[4 * q + 3]);
    *reinterpret_cast<uint4*>(dst + kImg + q * 2048) = make_uint4(lo[4 * q], lo[4 * q + 1], lo[4 * q + 2], lo[4 * q + 3]);
  }
  fence_async_smem();
  tc_fence_before();
  mbar_arrive(&xt_full[yb]);
  if ((threadIdx.x & 255) == 0) TR(id.w, n3, 4);
  const uint32_t back = (uint32_t)nd - 1u;
  if (!cfg.ext && nd > 1 && n3 >= back && Aout != nullptr) {
    const uint32_t k = n3 - back;
    const int kt = nd == 3 ? ((k & 1u) ? prev_tile[1] : prev_tile[0]) : prev_tile[0];
    drain_d3(id, kt, g, Aout, tmem_d3 + 32u * d3_slot(k, nd), &d3_full[d3_slot(k, nd)], d3_use(k, nd));
  }
  if (nd == 3 && (n3 & 1u)) prev_tile[1] = tile;
  else prev_tile[0] = tile;
  ++n3;
}

// after a worker's last tile: the accumulators that are still to be drained
__device__ __forceinline__ void drain_tail(const EpiId& id, const TcGeom& g, float* Aout, uint32_t tmem_d3, uint64_t* d3_full,
                                           uint32_t n3, const int (&prev_tile)[2], const TailCfg cfg) {
  if (Aout == nullptr) return;
  const int nd = cfg.nd;
  const uint32_t left = nd == 3 ? 2u : 1u;
  for (uint32_t k = n3 > left ? n3 - left : 0u; k < n3; ++k) {
    const int kt = nd == 3 ? ((k & 1u) ? prev_tile[1] : prev_tile[0]) : prev_tile[0];
    drain_d3(id, kt, g, Aout, tmem_d3 + 32u * d3_slot(k, nd), &d3_full[d3_slot(k, nd)], d3_use(k, nd));
  }
}

// control warp: n-th tile handed over by the epilogue group -> forward W-DFT UMMAs (hi*hi, lo*hi,
// hi*lo; when `dft`), bulk store of the image, release of the buffer
__device__ __forceinline__ void ctrl_finish_tile(uint32_t leader, bool dft, uint32_t aF, uint8_t* sY, uint8_t* y_dst,
                                                 uint32_t tmem_d3, uint64_t* xt_full, uint64_t* d3_full, uint64_t* y_free,
                                                 uint32_t n, const TailCfg cfg, bool tr_on = false, int tw = 0,
                                                 uint64_t* d3_empty = nullptr, uint64_t y_policy = 0) {
  const int ny = cfg.ny, nd = cfg.nd;
  const uint32_t yb = ny == 2 ? (n & 1u) : 0u, use = ny == 2 ? (n >> 1) : n;
  sY += yb * kTileBytes;
  mbar_wait(&xt_full[yb], use & 1u);
  TR(tw, n, 5);
  if (dft && cfg.ext && d3_use(n, nd) > 0) mbar_wait(&d3_empty[d3_slot(n, nd)], (d3_use(n, nd) - 1) & 1u);   // drained by the drain warps
  tc_fence_after();
  if (dft) {
    const uint32_t idesc = idesc_bf16(64, 32) | kIdescBMnMajor;
    // A: the 32-row image read as M = 64 (rows 32..63 alias the next K chunk: their accumulator
    // rows are never looked at); B: 8x8 core matrices 128 B apart along pixels, 2048 B along channels
    const uint64_t dF = chunk_desc(aF, 32, 0), dY = smem_desc(smem_u32(sY), 128, 2048);
    if (cfg.ext) {
      // the lo image follows the hi image 4 channel groups (4 x 2048 B) on: one N = 64 operand [Y_hi | Y_lo]
      const uint32_t idesc64 = idesc_bf16(64, 64) | kIdescBMnMajor, d = tmem_d3 + 64u * d3_slot(n, nd);
#pragma unroll
      for (int k0 = 0; k0 < 128; k0 += 16)
        umma_bf16_p(leader, d, desc_off(dF, (k0 >> 3) * 512), desc_off(dY, k0 * 16), idesc64, k0 != 0);
#pragma unroll
      for (int k0 = 0; k0 < 128; k0 += 16)
        umma_bf16_p(leader, d, desc_off(dF, 32u * 128u * 2u + (k0 >> 3) * 512), desc_off(dY, k0 * 16), idesc, 1);
    } else {
#pragma unroll
      for (int term = 0; term < 3; ++term) {
        const uint32_t fa = term == 1 ? 32u * 128u * 2u : 0u, yp = term == 2 ? (uint32_t)kImg : 0u;
#pragma unroll
        for (int k0 = 0; k0 < 128; k0 += 16)
          umma_bf16_p(leader, tmem_d3 + 32u * d3_slot(n, nd), desc_off(dF, fa + (k0 >> 3) * 512), desc_off(dY, yp + k0 * 16), idesc,
                      (term | k0) != 0);
      }
    }
    umma_commit_p(leader, &d3_full[d3_slot(n, nd)]);
  }
  bulk_s2g_hp(leader, y_dst, sY, kTileBytes, y_policy);
  bulk_commit();
  if (cfg.ext) {
    // y_free takes two arrivals per use: the UMMAs above have read the image / the bulk store has.  With two image
    // buffers the control warp does not sit out the store's read: it confirms the store of tile n - 1 here.
    if (dft) umma_commit_p(leader, &y_free[yb]);
    else mbar_arrive_p(leader, &y_free[yb]);
    if (ny == 2) {
      bulk_wait_read1();
      if (n > 0) mbar_arrive_p(leader, &y_free[yb ^ 1u]);
    } else {
      bulk_wait_read0();
      mbar_arrive_p(leader, &y_free[yb]);
    }
  } else {
    bulk_wait_read0();
    if (dft) umma_commit_p(leader, &y_free[yb]);   // arrives when the UMMAs above have read the image too
    else mbar_arrive_p(leader, &y_free[yb]);
  }
  __syncwarp();
}

}  // namespace

// =============================================================================================
// fno_layer_kernel: y = act( irfft_W(Z) + skip(x) + bias ) on 128-pixel tiles
//   D[128 px, 32 ch] = G'[128, KZ] * Z'^T  +  X[128, 32] * Ws^T      (split-bf16: 3 terms each)
// G' is block-diagonal over the grid rows of the tile; Z' (K index = (retained column, grid row of
// the tile, Re | Im)) arrives from fno_mid_kernel as a ready operand image.  Each worker cycles
// three input slots (activation image + Z' image, one bulk copy each): a slot is refilled as soon
// as the UMMAs that read it have completed, so up to three tiles per worker are in flight while
// the epilogue group works on the current one.
// =============================================================================================
namespace {
constexpr int kLayerSlots = 3;
struct LayerSmem {   // byte offsets; the inverse W-DFT operands scale with KZ
  int G, F, Ws, Bias, Worker, X, Z, Y, WorkerBytes, Total;
  __host__ __device__ explicit LayerSmem(int KZ, int ny) {
    G = 0;
    F = G + 2 * 128 * KZ * 2;
    Ws = F + kFBytes;
    Bias = Ws + 4096;
    Worker = Bias + 128;
    X = 0;
    Z = X + kLayerSlots * kTileBytes;
    Y = Z + kLayerSlots * KZ * 128;
    WorkerBytes = Y + ny * kTileBytes;
    Total = Worker + 2 * WorkerBytes;
  }
};
constexpr int kLayerSmemMax = 226 * 1024;   // 227 KB per CTA minus the kernel's static shared memory
enum { LB_XFULL = 0, LB_XFREE = 3, LB_DFULL = 6, LB_DEMPTY = 8, LB_XTFULL = 10, LB_D3FULL = 12, LB_YFREE = 15, LB_D3EMPTY = 17, LB_PER = 20 };
constexpr int kCtaLayerAll = kCtaLayer + 2 * 32;      // + two drain warps (TMEM lane quarters 0 and 1)
}  // namespace

__global__ void __launch_bounds__(kCtaLayerAll, 1) fno_layer_kernel(const TcLayerArgs a) {
  extern __shared__ __align__(128) uint8_t smem[];
  pdl_launch_dependents();
  __shared__ uint64_t bars[2 * LB_PER + 1];
  __shared__ uint32_t tmem_s;
  const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0);   // provably warp-uniform
  const TcGeom g = a.g;
  const LayerSmem L(g.KZ, a.ny);
  const TailCfg tail{a.ny, a.ny == 2 ? 3 : 1, true};
  [[maybe_unused]] const bool tr_on = a.gelu && a.Aout != nullptr;      // trace builds: a layer with GELU and forward W-DFT
  float* sBias = reinterpret_cast<float*>(smem + L.Bias);
  uint64_t* bar_const = &bars[2 * LB_PER];

  if (tid == 0) {
    for (int w = 0; w < 2; ++w) {
      uint64_t* b = bars + w * LB_PER;
      for (int i = 0; i < kLayerSlots; ++i) {
        mbar_init(&b[LB_XFULL + i], 1);
        mbar_init(&b[LB_XFREE + i], 1);
      }
      mbar_init(&b[LB_DFULL], 1);
      mbar_init(&b[LB_DFULL + 1], 1);
      mbar_init(&b[LB_DEMPTY], kEpi);
      mbar_init(&b[LB_DEMPTY + 1], kEpi);
      mbar_init(&b[LB_XTFULL], kEpi);
      mbar_init(&b[LB_XTFULL + 1], kEpi);
      mbar_init(&b[LB_D3FULL], 1);
      mbar_init(&b[LB_D3FULL + 1], 1);
      mbar_init(&b[LB_D3FULL + 2], 1);
      mbar_init(&b[LB_YFREE], 2);
      mbar_init(&b[LB_YFREE + 1], 2);
      for (int i = 0; i < 3; ++i) mbar_init(&b[LB_D3EMPTY + i], 2);
    }
    mbar_init(bar_const, 1);
    fence_mbar_init();
  }
  if (warp == kCtlWarp) {
    tmem_alloc(&tmem_s, 512);
    tmem_relinquish();
  }
  if (tid < 32) sBias[tid] = a.bias ? a.bias[tid] : 0.f;
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  // constant operands (packed at weight-commit time, which ends with a stream synchronisation): requested before the
  // programmatic-dependent-launch wait, so that on a partly filled machine they arrive under the previous kernel
  const uint32_t g_bytes = 2u * 128u * (uint32_t)g.KZ * 2u, z_bytes = (uint32_t)g.KZ * 128u;
  if (tid == 0) {
    mbar_arrive_expect_tx(bar_const, g_bytes + kFBytes + 4096);
    bulk_g2s(smem + L.G, a.Gimg, g_bytes, bar_const);
    bulk_g2s(smem + L.F, a.Fimg, kFBytes, bar_const);
    bulk_g2s(smem + L.Ws, a.Wsimg, 4096, bar_const);
  }
  pdl_wait();   // everything below reads what the previous kernel wrote (or overwrites what it read)
  const int stride = gridDim.x * 2;

  if (warp >= kCtlWarp + 4) {
    // ---------------- drain warps: the forward W-DFT accumulators of both workers -> planar spectrum ----------------
    // (an M = 64 accumulator keeps row r in TMEM lane 32 * (r / 16) + r % 16: warp q of the two reads rows 16 q .. 16 q + 15,
    //  one 128-byte row of Aout[b][kx][h][re|im][32] per lane.  In the epilogue group this drain sat on the
    //  critical path of every tile: a TMEM round trip, the address arithmetic and the stores, ~1.3 k cycles.)
    if (a.Aout != nullptr) {
      const int q = warp - kCtlWarp - 4, lane = tid & 31;
      const int row = q * 16 + lane;
      const int kxs = 31 - __clz(g.KXP);      // KXP = 32 / RT is a power of two
      const int h2 = row >> kxs, rem = row - (h2 << kxs), kx = rem >> 1, reim = rem & 1;
      const bool live = lane < 16 && h2 < g.RT && kx < g.Kc;
      const int nd = tail.nd;
      for (uint32_t u = 0;; ++u) {
        const int t0 = (int)blockIdx.x + (int)u * stride;
        if (t0 >= g.ntiles) break;
        for (int w = 0; w < 2; ++w) {
          const int slot = t0 + w * (int)gridDim.x;
          if (slot >= g.ntiles) break;
          const int tile = g.rev ? g.ntiles - 1 - slot : slot;
          uint64_t* b = bars + w * LB_PER;
          const uint32_t sl = d3_slot(u, nd);
          mbar_wait(&b[LB_D3FULL + sl], d3_use(u, nd) & 1u);
          tc_fence_after();
          uint32_t r[32], r2[32];      // columns 0..31: F_hi Y_hi + F_lo Y_hi, 32..63: F_hi Y_lo
          const uint32_t ta = tmem_s + 256u * w + 64u + 64u * sl + ((uint32_t)(q * 32) << 16);
          tmem_ld16(ta, r);
          tmem_ld16(ta + 16u, r + 16);
          tmem_ld16(ta + 32u, r2);
          tmem_ld16(ta + 48u, r2 + 16);
          tmem_ld_wait();
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&b[LB_D3EMPTY + sl]);
          if (live) {
            const int bb = tile / g.TPS, h = (tile - bb * g.TPS) * g.RT + h2;
            if (g.abf) {
              // split-bf16 operand tiles of fno_mid_tc_kernel: [kx][hi | lo][b][channel group][k = 2h + (re|im)][8 channels]
              uint32_t hi[16], lo[16];
#pragma unroll
              for (int e = 0; e < 16; ++e)
                split2p(make_float2(__uint_as_float(r[2 * e]) + __uint_as_float(r2[2 * e]), __uint_as_float(r[2 * e + 1]) + __uint_as_float(r2[2 * e + 1])),
                        hi[e], lo[e]);
              const size_t tile_b = (size_t)g.H * 128;
              uint8_t* d = reinterpret_cast<uint8_t*>(a.Aout) + ((size_t)(kx * 2) * g.Bn + bb) * tile_b + (size_t)(2 * h + reim) * 16;
#pragma unroll
              for (int gq = 0; gq < 4; ++gq) {
                *reinterpret_cast<uint4*>(d + gq * (g.H * 32)) = make_uint4(hi[4 * gq], hi[4 * gq + 1], hi[4 * gq + 2], hi[4 * gq + 3]);
                *reinterpret_cast<uint4*>(d + (size_t)g.Bn * tile_b + gq * (g.H * 32)) =
                    make_uint4(lo[4 * gq], lo[4 * gq + 1], lo[4 * gq + 2], lo[4 * gq + 3]);
              }
            } else {
            float4* dst = reinterpret_cast<float4*>(a.Aout + ((((size_t)bb * g.Kc + kx) * g.H + h) * 2 + reim) * 32);
#pragma unroll
            for (int j = 0; j < 8; ++j)
              dst[j] = make_float4(__uint_as_float(r[4 * j]) + __uint_as_float(r2[4 * j]),
                                   __uint_as_float(r[4 * j + 1]) + __uint_as_float(r2[4 * j + 1]),
                                   __uint_as_float(r[4 * j + 2]) + __uint_as_float(r2[4 * j + 2]),
                                   __uint_as_float(r[4 * j + 3]) + __uint_as_float(r2[4 * j + 3]));
            }
          }
        }
      }
    }
  } else if (warp >= kCtlWarp + 2) {
    // ---------------- loader / finisher lane: input bulk copies, forward W-DFT, image store ----------------
    // (its own warp so that a late input never delays handing a finished tile on, and vice versa)
    {
      const uint32_t leader = elect_one();
      const int w = warp - kCtlWarp - 2;
      uint64_t* b = bars + w * LB_PER;
      uint8_t* ws = smem + L.Worker + w * L.WorkerBytes;
      const uint32_t tmem = tmem_s + 256u * w;
      const uint32_t aF = smem_u32(smem + L.F);
      const int first = (int)blockIdx.x + w * (int)gridDim.x;

      const uint64_t pol_in = g.l2hint ? l2_evict_first() : 0, pol_out = g.l2hint ? l2_evict_last() : 0;
      auto load_inputs = [&](int tile, uint32_t u) {
        const int s = u % kLayerSlots;
        mbar_wait(&b[LB_XFREE + s], ((u / kLayerSlots) & 1u) ^ 1u);   // the UMMAs that read this slot last are done
        TR(w, u, 0);
        mbar_arrive_expect_tx_p(leader, &b[LB_XFULL + s], kTileBytes + z_bytes);
        const size_t t = (size_t)(g.rev ? g.ntiles - 1 - tile : tile);
        bulk_g2s_hp(leader, ws + L.X + s * kTileBytes, a.x_img + t * kTileBytes, kTileBytes, &b[LB_XFULL + s], pol_in);
        bulk_g2s_hp(leader, ws + L.Z + s * z_bytes, a.Zimg + t * z_bytes, z_bytes, &b[LB_XFULL + s], pol_in);
      };

      for (int i = 0; i < kLayerSlots; ++i)
        if (first + i * stride < g.ntiles) load_inputs(first + i * stride, i);
      mbar_wait(bar_const, 0);
      uint32_t u = 0;
      for (int tile = first; tile < g.ntiles; tile += stride, ++u) {
        if (tile + kLayerSlots * stride < g.ntiles) load_inputs(tile + kLayerSlots * stride, u + kLayerSlots);
        ctrl_finish_tile(leader, a.Aout != nullptr, aF, ws + L.Y, a.y_img + (size_t)(g.rev ? g.ntiles - 1 - tile : tile) * kTileBytes, tmem + 64,
                         &b[LB_XTFULL], &b[LB_D3FULL], &b[LB_YFREE], u, tail, tr_on, w, &b[LB_D3EMPTY], pol_out);
        TR(w, u, 6);
      }
      bulk_wait0();
    }
  } else if (warp >= kCtlWarp) {
    // ---------------- issuer lane: the main UMMAs, as soon as a tile's inputs and an accumulator are there ----------------
    {
      const uint32_t leader = elect_one();
      const int w = warp - kCtlWarp;
      uint64_t* b = bars + w * LB_PER;
      uint8_t* ws = smem + L.Worker + w * L.WorkerBytes;
      const uint32_t tmem = tmem_s + 256u * w;
      const uint32_t aG = smem_u32(smem + L.G), aW = smem_u32(smem + L.Ws);
      const uint64_t dG = chunk_desc(aG, 128, 0), dW = chunk_desc(aW, 32, 0);
      const int first = (int)blockIdx.x + w * (int)gridDim.x;
      mbar_wait(bar_const, 0);
      const uint32_t idesc = idesc_bf16(128, 32);
      const uint32_t g_lo = 128u * (uint32_t)g.KZ * 2u, z_lo = (uint32_t)g.KZ * 64u;
      uint32_t u = 0;
      for (int tile = first; tile < g.ntiles; tile += stride, ++u) {
        const int s = u % kLayerSlots, db = u & 1;
        mbar_wait(&b[LB_XFULL + s], (u / kLayerSlots) & 1u);
        TR(w, u, 1);
        mbar_wait(&b[LB_DEMPTY + db], empty_par(u));
        TR(w, u, 2);
        tc_fence_after();
        // Z': MN-major B operand (K rows of 16 bytes = 8 channels; 8-row core matrices 128 B apart along K, KZ * 16 B along the channels)
        const uint64_t dX = chunk_desc(smem_u32(ws + L.X + s * kTileBytes), 128, 0),
                       dZ = smem_desc(smem_u32(ws + L.Z + s * z_bytes), 128, (uint32_t)g.KZ * 16u);
        const uint32_t d = tmem + 32u * db;
        uint32_t acc = 0;
#pragma unroll
        for (int term = 0; term < 3; ++term) {      // hi*hi, lo*hi, hi*lo
          const uint32_t ga = term == 1 ? g_lo : 0u, zb = term == 2 ? z_lo : 0u;
          const uint32_t xa = term == 1 ? (uint32_t)kImg : 0u, wb = term == 2 ? 2048u : 0u;
#pragma unroll
          for (int k0 = 0; k0 < 128; k0 += 16) {     // operand chunks: 128 (32) rows x 16 bytes per 8 K; unrolled with a uniform guard
            if (k0 < g.KZ) {                          // (descriptors stay in uniform registers: no R2UR moves per instruction)
              umma_bf16_p(leader, d, desc_off(dG, ga + (k0 >> 3) * 2048), desc_off(dZ, zb + (uint32_t)k0 * 16u), idesc | kIdescBMnMajor, acc);
              acc = 1;
            }
          }
#pragma unroll
          for (int k0 = 0; k0 < 32; k0 += 16)
            umma_bf16_p(leader, d, desc_off(dX, xa + (k0 >> 3) * 2048), desc_off(dW, wb + (k0 >> 3) * 512), idesc, 1);
        }
        umma_commit_p(leader, &b[LB_DFULL + db]);
        umma_commit_p(leader, &b[LB_XFREE + s]);
        __syncwarp();
      }
    }
  } else {
    // ------------------------------ epilogue group ------------------------------
    const EpiId id(tid);
    const int w = id.w;
    uint64_t* b = bars + w * LB_PER;
    uint8_t* ws = smem + L.Worker + w * L.WorkerBytes;
    const uint32_t tmem = tmem_s + 256u * w + ((uint32_t)(id.q4 * 32) << 16) + 16u * id.half;
    uint32_t u = 0, n3 = 0;
    int prev_tile[2] = {0, 0};
    for (int tile = (int)blockIdx.x + w * (int)gridDim.x; tile < g.ntiles; tile += stride, ++u) {
      const int s = u & 1;
      mbar_wait(&b[LB_DFULL + s], full_par(u));
      if ((tid & 255) == 0) TR(w, u, 3);
      tc_fence_after();
      float2 v[8];
      {
        uint32_t r[16];
        tmem_ld16(tmem + 32u * s, r);
        tmem_ld_wait();
        tc_fence_before();
        mbar_arrive(&b[LB_DEMPTY + s]);
        const float2* bias2 = reinterpret_cast<const float2*>(sBias + 16 * id.half);
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          const float2 bb = bias2[c];
          const float2 x = make_float2(__uint_as_float(r[2 * c]) + bb.x, __uint_as_float(r[2 * c + 1]) + bb.y);
          v[c] = a.gelu ? gelu2(x) : x;
        }
      }
      finish_tile(v, id, g.rev ? g.ntiles - 1 - tile : tile, g, a.Aout, ws + L.Y, tmem_s + 256u * w + 64, &b[LB_XTFULL], &b[LB_D3FULL],
                  &b[LB_YFREE], n3, prev_tile, tail, tr_on);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == kCtlWarp) tmem_dealloc(tmem_s, 512);
}

int launch_tc_layer(const TcLayerArgs& a_in, cudaStream_t st) {
  TcLayerArgs a = a_in;
  a.ny = LayerSmem(a.g.KZ, 2).Total <= kLayerSmemMax ? 2 : 1;      // two output image buffers per worker when they fit
  int sms = 0;
  DLWP_TRY(current_sms(&sms));
  DLWP_TRY(ensure_dynamic_smem((const void*)fno_layer_kernel, kLayerSmemMax));
  const LayerSmem L(a.g.KZ, a.ny);
  DLWP_REQUIRE(a.g.KZ >= 16 && a.g.KZ % 16 == 0 && a.g.KZ <= 128 && L.Total <= kLayerSmemMax, "fno_layer_kernel: KZ=%d does not fit", a.g.KZ);
  // tile t of a CTA's list goes to worker (t / grid) % 2: with fewer tiles than SMs every tile has an SM (and its issue slots) to itself
  const int grid = a.g.ntiles < sms ? a.g.ntiles : sms;
  DLWP_CUDA(launch_pdl(fno_layer_kernel, dim3(grid), dim3(kCtaLayerAll), (size_t)L.Total, st, a));
  DLWP_LAUNCHED("fno_layer_kernel");
  return DLWP_OK;
}

// =============================================================================================
// Fused 2-layer pointwise MLP (neuralop MLP: 1x1 conv, GELU, 1x1 conv) on 128-pixel tiles
//   D1[128, 64]  = A1 * W1[pass]^T                 hidden columns, 64 per pass, double-buffered in TMEM
//   A2           = split(GELU(D1 + b1))            32 columns per chunk, double-buffered in SMEM
//   D2[128, N2] += A2 * W2[:, chunk]^T             double-buffered across tiles
// The hidden activations never leave the SM.  The control lane runs one pass ahead of the
// epilogue group, so the group only ever waits for data that had a full pass of time to arrive.
// Worker TMEM columns: D1 at +0 / +64, D2 at +128 / +160, D3 (lift only) at +192 / +224.
// =============================================================================================

// =============================================================================================
// fno_proj_kernel: out = last + W2 * GELU(W1 * x + b1) + b2, tile image in, NCHW frame out
// =============================================================================================
namespace {
struct ProjSmem {   // sized for hid = 256
  static constexpr int W1 = 0, W2 = W1 + 256 * 32 * 4, W1x = W2 + 16 * 256 * 4, Ones = W1x + 256 * 16 * 2, B2 = Ones + 4096,
                       Worker = B2 + 64;
  static constexpr int X = 0, Last = X + 2 * kTileBytes, WorkerBytes = Last + 2 * 16 * 512;
  static constexpr int Total = Worker + 2 * WorkerBytes;      // Last: two slots of [16 channels][128 pixels] fp32
};
}  // namespace

__global__ void __launch_bounds__(kCta, 1) fno_proj_kernel(const TcProjArgs a) {
  extern __shared__ __align__(128) uint8_t smem[];
  pdl_launch_dependents();
  __shared__ uint64_t bars[2 * MB_PER + 1];
  __shared__ uint32_t tmem_s;
  const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0);   // provably warp-uniform
  const TcGeom g = a.g;
  const int hid = a.hid, P = hid / 64;
  float* sB2 = reinterpret_cast<float*>(smem + ProjSmem::B2);
  uint64_t* bar_const = &bars[2 * MB_PER];

  if (tid == 0) {
    mlp_init_barriers(bars, 1);
    mlp_init_barriers(bars + MB_PER, 1);
    mbar_init(bar_const, 1);
    fence_mbar_init();
  }
  if (warp == kCtlWarp) {
    tmem_alloc(&tmem_s, 512);
    tmem_relinquish();
  }
  {   // Ones[128, 16]: columns 0 and 1 are 1 (they multiply the hi and lo halves of the bias in W1x)
    uint32_t* o = reinterpret_cast<uint32_t*>(smem + ProjSmem::Ones);
    for (int i = tid; i < 1024; i += kCta) o[i] = (i < 512 && (i & 3) == 0) ? 0x3F803F80u : 0u;
  }
  fence_async_smem();
  if (tid < 16) sB2[tid] = tid < a.Cg ? a.b2[tid] : 0.f;
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t w1_bytes = (uint32_t)hid * 32 * 4, w2_bytes = 16u * hid * 4, w1x_bytes = (uint32_t)hid * 16 * 2;
  if (tid == 0) {   // constant operands: requested before the programmatic-dependent-launch wait
    mbar_arrive_expect_tx(bar_const, w1_bytes + w2_bytes + w1x_bytes);
    bulk_g2s(smem + ProjSmem::W1, a.W1img, w1_bytes, bar_const);
    bulk_g2s(smem + ProjSmem::W2, a.W2img, w2_bytes, bar_const);
    bulk_g2s(smem + ProjSmem::W1x, a.W1ximg, w1x_bytes, bar_const);
  }
  pdl_wait();   // everything below reads what the previous kernel wrote (or overwrites what it read)
  const int stride = gridDim.x * 2;

  if (warp >= kCtlWarp) {
    {
      const uint32_t leader = elect_one();
      const int w = warp - kCtlWarp;
      uint64_t* b = bars + w * MB_PER;
      uint8_t* ws = smem + ProjSmem::Worker + w * ProjSmem::WorkerBytes;
      MlpIssuer is{b, smem_u32(ws + ProjSmem::X), (uint32_t)kTileBytes, smem_u32(smem + ProjSmem::W1),
                   smem_u32(smem + ProjSmem::W2), tmem_s + 256u * w, hid, P, 0u,
                   smem_u32(smem + ProjSmem::Ones), smem_u32(smem + ProjSmem::W1x), leader};
      const int first = (int)blockIdx.x + w * (int)gridDim.x;
      mbar_wait(bar_const, 0);
      auto load_x = [&](int tile, uint32_t u) {
        const int s = u & 1;
        mbar_wait(&b[MB_A1EMPTY + s], empty_par(u));
        mbar_arrive_expect_tx_p(leader, &b[MB_A1FULL + s], kTileBytes);
        bulk_g2s_p(leader, ws + ProjSmem::X + s * kTileBytes, a.x_img + (size_t)tile * kTileBytes, kTileBytes, &b[MB_A1FULL + s]);
      };
      // the tile's rows of the residual frame (128 consecutive pixels per channel) for the epilogue: bulk copies,
      // because a generic global load in flight in an epilogue thread would be waited for by its next proxy fence
      auto load_last = [&](int tile, uint32_t u) {
        const int s = u & 1;
        const long long bb = tile / g.TPS, pix = (long long)(tile - (int)bb * g.TPS) * kTcTile;
        mbar_wait(&b[MB_INEMPTY + s], empty_par(u));
        mbar_arrive_expect_tx_p(leader, &b[MB_INFULL + s], (uint32_t)a.Cg * 512u);
        for (int c = 0; c < a.Cg; ++c)
          bulk_g2s_p(leader, ws + ProjSmem::Last + s * 8192 + c * 512, a.last + bb * a.last_bstride + (long long)c * g.HW + pix, 512u,
                     &b[MB_INFULL + s]);
      };
      const bool bulk_last = a.last != nullptr && a.last_bulk;
      if (first < g.ntiles) {
        load_x(first, 0);
        is.mma1<32>(0, 0, 0);
      }
      uint32_t u = 0;
      for (int tile = first; tile < g.ntiles; tile += stride, ++u) {
        const bool has_next = tile + stride < g.ntiles;
        if (has_next) load_x(tile + stride, u + 1);
        if (bulk_last) load_last(tile, u);
        for (int p = 0; p < P; ++p) {
          const uint32_t gp = u * P + p;
          if (p + 1 < P) is.mma1<32>(u, p + 1, gp + 1);
          else if (has_next) is.mma1<32>(u + 1, 0, gp + 1);
          is.mma2<16>(u, p);
        }
      }
    }
  } else {
    const EpiId id(tid);
    const int w = id.w;
    uint64_t* b = bars + w * MB_PER;
    uint8_t* ws = smem + ProjSmem::Worker + w * ProjSmem::WorkerBytes;
    const uint32_t tmem_lane = tmem_s + 256u * w + ((uint32_t)(id.q4 * 32) << 16) + 16u * id.half;
    const int c0 = 8 * id.half;      // this thread's 8 output channels
    uint32_t u = 0, gp = 0, nc = 0;
    // A tile is finished one tile late, after the first pass of the next one: its second GEMM completes under
    // that pass instead of being waited for (D2 is double buffered in TMEM).
    float lastv[8];
    long long obase = 0;
    const bool bulk_last = a.last != nullptr && a.last_bulk;
    const float* sLast = reinterpret_cast<const float*>(ws + ProjSmem::Last);
    auto finish = [&](uint32_t v) {
      if (bulk_last) {
        mbar_wait(&b[MB_INFULL + (v & 1)], full_par(v));
#pragma unroll
        for (int c = 0; c < 8; ++c) lastv[c] = c0 + c < a.Cg ? sLast[(v & 1) * 2048 + (c0 + c) * 128 + id.row] : 0.f;
        mbar_arrive(&b[MB_INEMPTY + (v & 1)]);
      }
      mbar_wait(&b[MB_D2FULL + (v & 1)], full_par(v));
      tc_fence_after();
      uint32_t r[8];
      tmem_ld8(tmem_s + 256u * w + ((uint32_t)(id.q4 * 32) << 16) + 128u + 32u * (v & 1) + (uint32_t)c0, r);
      tmem_ld_wait();
      tc_fence_before();
      mbar_arrive(&b[MB_D2EMPTY + (v & 1)]);
      float* o = a.out + obase;
#pragma unroll
      for (int c = 0; c < 8; ++c)
        if (c0 + c < a.Cg) o[(long long)(c0 + c) * g.HW] = __uint_as_float(r[c]) + sB2[c0 + c] + lastv[c];
    };
    auto load_last = [&](int tile) {      // output position; without bulk copies also the residual frame values
      const long long bb = tile / g.TPS, pix = (long long)(tile - (int)bb * g.TPS) * kTcTile + id.row;
      if (!bulk_last) {
#pragma unroll
        for (int c = 0; c < 8; ++c)
          lastv[c] = (a.last != nullptr && c0 + c < a.Cg) ? a.last[bb * a.last_bstride + (long long)(c0 + c) * g.HW + pix] : 0.f;
      }
      obase = bb * a.out_bstride + pix;
    };
    int prev = -1;
    for (int tile = (int)blockIdx.x + w * (int)gridDim.x; tile < g.ntiles; tile += stride, ++u) {
      if (prev >= 0) load_last(prev);
      epi_mlp_tile(b, id, tmem_lane, P, gp, nc, [&](int pass) { if (pass == 0 && prev >= 0) finish(u - 1); });
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

int launch_tc_proj(const TcProjArgs& a_in, cudaStream_t st) {
  TcProjArgs a = a_in;
  a.last_bulk = a.last != nullptr && (reinterpret_cast<uintptr_t>(a.last) & 15) == 0 && a.last_bstride % 4 == 0 && a.g.HW % 4 == 0;
  int sms = 0;
  DLWP_TRY(current_sms(&sms));
  DLWP_TRY(ensure_dynamic_smem((const void*)fno_proj_kernel, ProjSmem::Total));
  DLWP_REQUIRE(a.hid == 128 || a.hid == 256, "tensor-core projection needs width 128 or 256, got %d", a.hid);
  // tile t of a CTA's list goes to worker (t / grid) % 2: with fewer tiles than SMs every tile has an SM (and its issue slots) to itself
  const int grid = a.g.ntiles < sms ? a.g.ntiles : sms;
  DLWP_CUDA(launch_pdl(fno_proj_kernel, dim3(grid), dim3(kCta), (size_t)ProjSmem::Total, st, a));
  DLWP_LAUNCHED("fno_proj_kernel");
  return DLWP_OK;
}

// =============================================================================================
// fno_lift_kernel: x_t gathered from the NCHW sources -> lifting MLP -> tile image + W-DFT
// =============================================================================================
namespace {
struct LiftSmem {   // sized for hid = 256
  static constexpr int W1 = 0, W2 = W1 + 256 * 16 * 4, F = W2 + 32 * 256 * 4, B2 = F + kFBytes, Worker = B2 + 128;
  static constexpr int A1 = 0, Y = A1 + 2 * 8192, WorkerBytes = Y + kTileBytes;
  static constexpr int Total = Worker + 2 * WorkerBytes;
  static_assert(Total <= 226 * 1024, "lift kernel shared memory");
};
}  // namespace

__global__ void __launch_bounds__(kCtaLayer, 1) fno_lift_kernel(const TcLiftArgs a) {
  extern __shared__ __align__(128) uint8_t smem[];
  pdl_launch_dependents();
  __shared__ uint64_t bars[2 * MB_PER + 1];
  __shared__ uint32_t tmem_s;
  __shared__ const float* kptr[16];     // per input channel: plane base and batch stride
  __shared__ long long kbs[16];
  const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0);   // provably warp-uniform
  const TcGeom g = a.g;
  const int hid = a.hid, P = hid / 64;
  float* sB2 = reinterpret_cast<float*>(smem + LiftSmem::B2);
  uint64_t* bar_const = &bars[2 * MB_PER];

  if (tid == 0) {
    mlp_init_barriers(bars, kEpi);
    mlp_init_barriers(bars + MB_PER, kEpi);
    mbar_init(bar_const, 1);
    fence_mbar_init();
    int k = 0;     // channel order: constants | prescribed (t, c) | prognostic (t, c)  (fno.py:58-62)
    for (int gi = 0; gi < a.ngrp; ++gi)
      for (int ch = 0; ch < a.grp[gi].nch && k < 16; ++ch, ++k) {
        kptr[k] = a.grp[gi].base + (long long)ch * g.HW;
        kbs[k] = a.grp[gi].bstride;
      }
    for (; k < 16; ++k) {
      kptr[k] = a.grp[0].base;
      kbs[k] = 0;
    }
  }
  if (warp == kCtlWarp) {
    tmem_alloc(&tmem_s, 512);
    tmem_relinquish();
  }
  if (tid < 32) sB2[tid] = a.b2[tid];
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t w1_bytes = (uint32_t)hid * 16 * 4, w2_bytes = 32u * hid * 4;
  if (tid == 0) {   // constant operands: requested before the programmatic-dependent-launch wait
    mbar_arrive_expect_tx(bar_const, w1_bytes + w2_bytes + kFBytes);
    bulk_g2s(smem + LiftSmem::W1, a.W1img, w1_bytes, bar_const);
    bulk_g2s(smem + LiftSmem::W2, a.W2img, w2_bytes, bar_const);
    bulk_g2s(smem + LiftSmem::F, a.Fimg, kFBytes, bar_const);
  }
  pdl_wait();   // everything below reads what the previous kernel wrote (or overwrites what it read)
  const int stride = gridDim.x * 2;

  if (warp >= kCtlWarp + 2) {
    // ---------------- finisher lane: forward W-DFT of a handed-over tile, image store ----------------
    // (its own warp: in the issuer's loop the wait for the hand-over, the chain of W-DFT UMMAs and the bulk store's
    //  read-completion wait delayed the MLP's UMMAs, and the epilogue group then waited for its next accumulator:
    //  12 % of this kernel's stall samples)
    const uint32_t leader = elect_one();
    const int w = warp - kCtlWarp - 2;
    uint64_t* b = bars + w * MB_PER;
    uint8_t* ws = smem + LiftSmem::Worker + w * LiftSmem::WorkerBytes;
    const uint32_t aF = smem_u32(smem + LiftSmem::F), d3 = tmem_s + 256u * w + 192u;
    const int first = (int)blockIdx.x + w * (int)gridDim.x;
    const bool dft = a.Aout != nullptr;
    mbar_wait(bar_const, 0);
    uint32_t u = 0;
    for (int tile = first; tile < g.ntiles; tile += stride, ++u)
      ctrl_finish_tile(leader, dft, aF, ws + LiftSmem::Y, a.y_img + (size_t)tile * kTileBytes, d3, &b[MB_XTFULL], &b[MB_D3FULL],
                       &b[MB_YFREE], u, TailCfg{1, 2, false}, false, 0, nullptr, g.l2hint ? l2_evict_last() : 0);
    bulk_wait0();
  } else if (warp >= kCtlWarp) {
    // ---------------- issuer lane: the MLP's UMMAs ----------------
    const uint32_t leader = elect_one();
    const int w = warp - kCtlWarp;
    uint64_t* b = bars + w * MB_PER;
    uint8_t* ws = smem + LiftSmem::Worker + w * LiftSmem::WorkerBytes;
    MlpIssuer is{b, smem_u32(ws + LiftSmem::A1), 8192u, smem_u32(smem + LiftSmem::W1), smem_u32(smem + LiftSmem::W2),
                 tmem_s + 256u * w, hid, P, 0u, 0u, 0u, leader};
    const int first = (int)blockIdx.x + w * (int)gridDim.x;
    mbar_wait(bar_const, 0);
    if (first < g.ntiles) is.mma1<16>(0, 0, 0);
    uint32_t u = 0;
    for (int tile = first; tile < g.ntiles; tile += stride, ++u) {
      const bool has_next = tile + stride < g.ntiles;
      for (int p = 0; p < P; ++p) {
        const uint32_t gp = u * P + p;
        if (p + 1 < P) is.mma1<16>(u, p + 1, gp + 1);
        else if (has_next) is.mma1<16>(u + 1, 0, gp + 1);
        is.mma2<32>(u, p);
      }
    }
  } else {
    const EpiId id(tid);
    const int w = id.w;
    uint64_t* b = bars + w * MB_PER;
    uint8_t* ws = smem + LiftSmem::Worker + w * LiftSmem::WorkerBytes;
    const uint32_t tmem_w = tmem_s + 256u * w, tmem_lane = tmem_w + ((uint32_t)(id.q4 * 32) << 16) + 16u * id.half;
    const int Cin = a.Cin, k0 = 8 * id.half;      // this thread's 8 input columns of its pixel

    auto gather = [&](int tile, float (&x)[8]) {
      const long long bb = tile / g.TPS, pix = (long long)(tile - (int)bb * g.TPS) * kTcTile + id.row;
#pragma unroll
      for (int k = 0; k < 8; ++k) x[k] = k0 + k < Cin ? kptr[k0 + k][bb * kbs[k0 + k] + pix] : (k0 + k == Cin ? 1.0f : 0.0f);
    };
    auto store_a1 = [&](uint32_t u, const float (&x)[8]) {   // column Cin carries the 1 that multiplies the bias
      const int s = u & 1;
      mbar_wait(&b[MB_A1EMPTY + s], empty_par(u));
      // the gathered values are first touched here: without this the compiler hoists the splits to right behind the loads and
      // the epilogue group sits out the gather's latency at the top of every tile (6 % of the kernel's stall samples)
      float xs[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) xs[k] = x[k];
      asm volatile("" : "+f"(xs[0]), "+f"(xs[1]), "+f"(xs[2]), "+f"(xs[3]), "+f"(xs[4]), "+f"(xs[5]), "+f"(xs[6]), "+f"(xs[7]));
      uint32_t h[4], l[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) split2(xs[2 * j], xs[2 * j + 1], h[j], l[j]);
      uint8_t* d = ws + LiftSmem::A1 + s * 8192 + id.half * 2048 + id.row * 16;
      *reinterpret_cast<uint4*>(d) = make_uint4(h[0], h[1], h[2], h[3]);
      *reinterpret_cast<uint4*>(d + 4096) = make_uint4(l[0], l[1], l[2], l[3]);
      fence_async_smem();
      mbar_arrive(&b[MB_A1FULL + s]);
    };

    const int first = (int)blockIdx.x + w * (int)gridDim.x;
    float nx[8];
    if (first < g.ntiles) {
      gather(first, nx);
      store_a1(0, nx);
    }
    uint32_t u = 0, gp = 0, nc = 0, n3 = 0;
    int prev_tile[2] = {0, 0};
    // tile v's second GEMM completes under the first pass of tile v + 1: its output is taken over after that pass
    auto finish_ep = [&](int tile_f, uint32_t v_) {
      mbar_wait(&b[MB_D2FULL + (v_ & 1)], full_par(v_));
      tc_fence_after();
      float2 v[8];
      {
        uint32_t r[16];
        tmem_ld16(tmem_lane + 128u + 32u * (v_ & 1), r);
        tmem_ld_wait();
        tc_fence_before();
        mbar_arrive(&b[MB_D2EMPTY + (v_ & 1)]);
        const float2* bias2 = reinterpret_cast<const float2*>(sB2 + 16 * id.half);
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          const float2 bb = bias2[c];
          v[c] = make_float2(__uint_as_float(r[2 * c]) + bb.x, __uint_as_float(r[2 * c + 1]) + bb.y);
        }
      }
      finish_tile(v, id, tile_f, g, a.Aout, ws + LiftSmem::Y, tmem_w + 192u, &b[MB_XTFULL], &b[MB_D3FULL], &b[MB_YFREE], n3,
                  prev_tile, TailCfg{1, 2, false});
    };
    const int a1_pass = P >= 3 ? P - 2 : 0;      // the next tile's gathered inputs are consumed as late as the issuer allows
    for (int tile = first; tile < g.ntiles; tile += stride, ++u) {
      const bool has_next = tile + stride < g.ntiles;
      if (has_next) gather(tile + stride, nx);      // loads in flight while the first passes are processed
      epi_mlp_tile(b, id, tmem_lane, P, gp, nc, [&](int pass) {
        if (pass == 0 && u > 0) finish_ep(tile - stride, u - 1);
        if (pass == a1_pass && has_next) store_a1(u + 1, nx);
      });
    }
    if (u > 0) finish_ep(first + (int)(u - 1) * stride, u - 1);
    drain_tail(id, g, a.Aout, tmem_w + 192u, &b[MB_D3FULL], n3, prev_tile, TailCfg{1, 2, false});
  }
  tc_fence_before();
  __syncthreads();
  if (warp == kCtlWarp) tmem_dealloc(tmem_s, 512);
}

// =============================================================================================
// fno_lift4_kernel: the same lifting MLP with FOUR workers of four epilogue warps per CTA.
// What bounds the two-worker kernel is the FMA pipe while both workers are in their GELU passes, and nothing while one of
// them gathers, hands a tile over or waits for an accumulator (a worker is in the GELU arithmetic 48 % of the time: the pipe
// averages 0.23 x 1 + 0.5 x 0.5 = 48 %, ncu: 49.5 %).  Four workers with independent phases keep it fed: a thread owns its
// pixel's whole 32-column chunk (one chunk per pass, two accumulator buffers), the tile's second accumulator is single
// buffered and read at the end of the tile, and the W-DFT accumulator is drained before the next hand-over: 128
// tensor-memory columns per worker.
//   worker w: epilogue warps 4 w .. 4 w + 3, issuer warp 16 + w, finisher warp 20 + w;  tiles blockIdx.x + (w + 4 n) gridDim.x
// =============================================================================================
namespace {
constexpr int kL4Workers = 4, kL4Epi = 128, kL4Iss = 16, kL4Fin = 20, kL4Cta = 24 * 32;
enum { L4_A1FULL = 0 /* two */, L4_A1EMPTY = 2 /* two */, L4_D1FULL = 4 /* two */, L4_A2FULL = 6 /* two */, L4_D2FULL = 8, L4_D2EMPTY = 9,
       L4_XTFULL = 10, L4_D3FULL = 11, L4_YFREE = 12, L4_PER = 13 };
struct Lift4Smem {   // sized for hid = 256
  static constexpr int W1 = 0, W2 = W1 + 256 * 16 * 4, F = W2 + 32 * 256 * 4, B2 = F + kFBytes, Worker = B2 + 128;
  static constexpr int A1 = 0, Y = A1 + 2 * 8192, WorkerBytes = Y + kTileBytes;
  static constexpr int Total = Worker + kL4Workers * WorkerBytes;
  static_assert(Total <= 226 * 1024, "lift4 kernel shared memory");
};
}  // namespace

__global__ void __launch_bounds__(kL4Cta, 1) fno_lift4_kernel(const TcLiftArgs a) {
  extern __shared__ __align__(128) uint8_t smem[];
  pdl_launch_dependents();
  __shared__ uint64_t bars[kL4Workers * L4_PER + 1];
  __shared__ uint32_t tmem_s;
  __shared__ const float* kptr[16];     // per input channel: plane base and batch stride
  __shared__ long long kbs[16];
  const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0);   // provably warp-uniform
  const TcGeom g = a.g;
  const int hid = a.hid, P = hid / 32;      // passes of one 32-column chunk
  float* sB2 = reinterpret_cast<float*>(smem + Lift4Smem::B2);
  uint64_t* bar_const = &bars[kL4Workers * L4_PER];

  if (tid == 0) {
    for (int w = 0; w < kL4Workers; ++w) {
      uint64_t* b = bars + w * L4_PER;
      for (int i = 0; i < 2; ++i) {
        mbar_init(&b[L4_A1FULL + i], kL4Epi);
        mbar_init(&b[L4_A1EMPTY + i], 1);
        mbar_init(&b[L4_D1FULL + i], 1);
        mbar_init(&b[L4_A2FULL + i], kL4Epi);
      }
      mbar_init(&b[L4_D2FULL], 1);
      mbar_init(&b[L4_D2EMPTY], kL4Epi);
      mbar_init(&b[L4_XTFULL], 2 * kL4Epi);      // finish_tile arrives once per 16-channel half
      mbar_init(&b[L4_D3FULL], 1);
      mbar_init(&b[L4_YFREE], 1);
    }
    mbar_init(bar_const, 1);
    fence_mbar_init();
    int k = 0;     // channel order: constants | prescribed (t, c) | prognostic (t, c)  (fno.py:58-62)
    for (int gi = 0; gi < a.ngrp; ++gi)
      for (int ch = 0; ch < a.grp[gi].nch && k < 16; ++ch, ++k) {
        kptr[k] = a.grp[gi].base + (long long)ch * g.HW;
        kbs[k] = a.grp[gi].bstride;
      }
    for (; k < 16; ++k) {
      kptr[k] = a.grp[0].base;
      kbs[k] = 0;
    }
  }
  if (warp == kL4Iss) {
    tmem_alloc(&tmem_s, 512);
    tmem_relinquish();
  }
  if (tid < 32) sB2[tid] = a.b2[tid];
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t w1_bytes = (uint32_t)hid * 16 * 4, w2_bytes = 32u * hid * 4;
  if (tid == 0) {   // constant operands: requested before the programmatic-dependent-launch wait
    mbar_arrive_expect_tx(bar_const, w1_bytes + w2_bytes + kFBytes);
    bulk_g2s(smem + Lift4Smem::W1, a.W1img, w1_bytes, bar_const);
    bulk_g2s(smem + Lift4Smem::W2, a.W2img, w2_bytes, bar_const);
    bulk_g2s(smem + Lift4Smem::F, a.Fimg, kFBytes, bar_const);
  }
  pdl_wait();   // everything below reads what the previous kernel wrote (or overwrites what it read)
  const int stride = gridDim.x * kL4Workers;
  const bool dft = a.Aout != nullptr;

  if (warp >= kL4Fin) {
    // ---------------- finisher lane: forward W-DFT of a handed-over tile, image store ----------------
    const uint32_t leader = elect_one();
    const int w = warp - kL4Fin;
    uint64_t* b = bars + w * L4_PER;
    uint8_t* ws = smem + Lift4Smem::Worker + w * Lift4Smem::WorkerBytes;
    const uint32_t aF = smem_u32(smem + Lift4Smem::F), d3 = tmem_s + 128u * w + 96u;
    const int first = (int)blockIdx.x + w * (int)gridDim.x;
    mbar_wait(bar_const, 0);
    uint32_t u = 0;
    for (int tile = first; tile < g.ntiles; tile += stride, ++u)
      ctrl_finish_tile(leader, dft, aF, ws + Lift4Smem::Y, a.y_img + (size_t)tile * kTileBytes, d3, &b[L4_XTFULL], &b[L4_D3FULL],
                       &b[L4_YFREE], u, TailCfg{1, 1, false}, false, 0, nullptr, g.l2hint ? l2_evict_last() : 0);
    bulk_wait0();
  } else if (warp >= kL4Iss) {
    // ---------------- issuer lane: the MLP's UMMAs, one 32-column chunk per pass ----------------
    // (the first GEMM of pass n + 1 overwrites the buffer the second GEMM of pass n - 1 read as its operand: issued by this
    //  thread after it, and tcgen05.mma instructions of one thread execute in issue order)
    const uint32_t leader = elect_one();
    const int w = warp - kL4Iss;
    uint64_t* b = bars + w * L4_PER;
    const uint32_t tmem_w = tmem_s + 128u * w;
    const uint32_t aA1 = smem_u32(smem + Lift4Smem::Worker + w * Lift4Smem::WorkerBytes + Lift4Smem::A1);
    const uint32_t aW1 = smem_u32(smem + Lift4Smem::W1), aW2 = smem_u32(smem + Lift4Smem::W2);
    const uint32_t idesc = idesc_bf16(128, 32);
    auto mma1 = [&](uint32_t u, int p, uint32_t gp) {
      const uint32_t ub = u & 1u;
      if (p == 0) mbar_wait(&b[L4_A1FULL + ub], full_par(u));
      tc_fence_after();
      const uint64_t dA = chunk_desc(aA1 + ub * 8192u, 128, 0), dW = smem_desc(aW1 + (uint32_t)p * 32 * 16, (uint32_t)hid * 16, 128);
      const uint32_t d = tmem_w + 32u * (gp & 1u);
      umma_bf16_p(leader, d, dA, dW, idesc, 0);                                           // hi * hi
      umma_bf16_p(leader, d, desc_off(dA, 128 * 16 * 2), dW, idesc, 1);                   // lo * hi
      umma_bf16_p(leader, d, dA, desc_off(dW, (uint32_t)hid * 16 * 2), idesc, 1);         // hi * lo
      umma_commit_p(leader, &b[L4_D1FULL + (gp & 1u)]);
      if (p == P - 1) umma_commit_p(leader, &b[L4_A1EMPTY + ub]);
      __syncwarp();
    };
    auto mma2 = [&](uint32_t u, int p, uint32_t gp) {
      mbar_wait(&b[L4_A2FULL + (gp & 1u)], (gp >> 1) & 1u);
      if (p == 0) mbar_wait(&b[L4_D2EMPTY], (u & 1u) ^ 1u);
      tc_fence_after();
      const uint64_t dW = chunk_desc(aW2, 32, p * 32);
      const uint32_t a2 = tmem_w + 32u * (gp & 1u), w_lo = 32u * (uint32_t)hid * 2;      // [K step 0: hi 8 columns | lo 8][K step 1: hi | lo]
#pragma unroll
      for (int term = 0; term < 3; ++term) {      // hi*hi, lo*hi, hi*lo
        const uint32_t ao = term == 1 ? 8u : 0u, wo = term == 2 ? w_lo : 0u;
#pragma unroll
        for (int kk = 0; kk < 2; ++kk)
          umma_ts_bf16_p(leader, tmem_w + 64u, a2 + 16u * kk + ao, desc_off(dW, wo + (uint32_t)kk * 2 * 32 * 16), idesc, (p | term | kk) != 0);
      }
      if (p == P - 1) umma_commit_p(leader, &b[L4_D2FULL]);
      __syncwarp();
    };
    const int first = (int)blockIdx.x + w * (int)gridDim.x;
    mbar_wait(bar_const, 0);
    if (first < g.ntiles) mma1(0, 0, 0);
    uint32_t u = 0;
    for (int tile = first; tile < g.ntiles; tile += stride, ++u) {
      const bool has_next = tile + stride < g.ntiles;
      for (int p = 0; p < P; ++p) {
        const uint32_t gp = u * (uint32_t)P + (uint32_t)p;
        if (p + 1 < P) mma1(u, p + 1, gp + 1);
        else if (has_next) mma1(u + 1, 0, gp + 1);
        mma2(u, p, gp);
      }
    }
  } else {
    // ------------------------------ epilogue warps: thread = pixel of the tile ------------------------------
    const int w = warp >> 2, q4 = warp & 3, row = q4 * 32 + (tid & 31);
    uint64_t* b = bars + w * L4_PER;
    uint8_t* ws = smem + Lift4Smem::Worker + w * Lift4Smem::WorkerBytes;
    const uint32_t tmem_w = tmem_s + 128u * w, tmem_lane = tmem_w + ((uint32_t)(q4 * 32) << 16);
    const int Cin = a.Cin;
    EpiId id(tid);      // for the shared tail helpers: .half selects the 16-channel half a call works on
    id.w = w;
    id.q4 = q4;
    id.row = row;

    // x_t of the thread's pixel (column Cin carries the 1 that multiplies the bias) -> split-bf16 operand rows.  Loaded and
    // used in one go: the other three workers run meanwhile.
    auto load_a1 = [&](int tile, uint32_t u) {
      const long long bb = tile / g.TPS, pix = (long long)(tile - (int)bb * g.TPS) * kTcTile + row;
      float x[16];
#pragma unroll
      for (int k = 0; k < 16; ++k) x[k] = k < Cin ? kptr[k][bb * kbs[k] + pix] : (k == Cin ? 1.0f : 0.0f);
      const int s = u & 1;
      mbar_wait(&b[L4_A1EMPTY + s], empty_par(u));
      uint8_t* d = ws + Lift4Smem::A1 + s * 8192 + row * 16;
#pragma unroll
      for (int c = 0; c < 2; ++c) {      // K chunks of 8 columns
        uint32_t h[4], l[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) split2(x[8 * c + 2 * j], x[8 * c + 2 * j + 1], h[j], l[j]);
        *reinterpret_cast<uint4*>(d + c * 2048) = make_uint4(h[0], h[1], h[2], h[3]);
        *reinterpret_cast<uint4*>(d + 4096 + c * 2048) = make_uint4(l[0], l[1], l[2], l[3]);
      }
      fence_async_smem();
      mbar_arrive(&b[L4_A1FULL + s]);
    };
    // the tile's output: second accumulator + bias -> split-bf16 image for the finisher lane; the W-DFT accumulator of the
    // tile before is drained first (one accumulator per worker)
    int prev_tile = -1;
    auto finish = [&](int tile, uint32_t u) {
      if (dft && prev_tile >= 0) {
#pragma unroll
        for (int hf = 0; hf < 2; ++hf) {
          id.half = hf;
          drain_d3(id, prev_tile, g, a.Aout, tmem_w + 96u, &b[L4_D3FULL], u - 1);
        }
      }
      mbar_wait(&b[L4_D2FULL], u & 1u);
      tc_fence_after();
      uint32_t r[32];
      tmem_ld32(tmem_lane + 64u, r);
      tmem_ld_wait();
      tc_fence_before();
      mbar_arrive(&b[L4_D2EMPTY]);
      if (u > 0) mbar_wait(&b[L4_YFREE], (u - 1) & 1u);      // the bulk store and the W-DFT UMMAs of the tile before have read the image
      uint8_t* dst = ws + Lift4Smem::Y + row * 16;
#pragma unroll
      for (int c = 0; c < 4; ++c) {      // channel groups of 8
        uint32_t hi[4], lo[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int ch = 8 * c + 2 * j;
          split2p(make_float2(__uint_as_float(r[ch]) + sB2[ch], __uint_as_float(r[ch + 1]) + sB2[ch + 1]), hi[j], lo[j]);
        }
        *reinterpret_cast<uint4*>(dst + c * 2048) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
        *reinterpret_cast<uint4*>(dst + kImg + c * 2048) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
      }
      fence_async_smem();
      tc_fence_before();
      mbar_arrive(&b[L4_XTFULL]);
      mbar_arrive(&b[L4_XTFULL]);
      prev_tile = tile;
    };

    const int first = (int)blockIdx.x + w * (int)gridDim.x;
    if (first < g.ntiles) load_a1(first, 0);
    const int a1_pass = P >= 4 ? P - 3 : 0;      // the issuer wants the next tile's operand behind this tile's last pass
    uint32_t u = 0, gp = 0;
    for (int tile = first; tile < g.ntiles; tile += stride, ++u) {
      const bool has_next = tile + stride < g.ntiles;
      for (int p = 0; p < P; ++p, ++gp) {
        const uint32_t buf = gp & 1u, t = tmem_lane + 32u * buf;
        mbar_wait(&b[L4_D1FULL + buf], (gp >> 1) & 1u);
        tc_fence_after();
        uint32_t r[32];
        tmem_ld32(t, r);
        tmem_ld_wait();
#pragma unroll
        for (int hf = 0; hf < 2; ++hf) {      // K steps of the second GEMM: 16 hidden columns -> 8 columns of hi pairs + 8 of lo pairs
          uint32_t h[8], l[8];
          float2 g2[8];             // 2 * gelu: W2's operand image is packed with the factor 1/2
          gelu2x_8(reinterpret_cast<const uint32_t (&)[16]>(r[16 * hf]), g2);
#pragma unroll
          for (int e = 0; e < 8; ++e) split2p(g2[e], h[e], l[e]);
          tmem_st8(t + 16u * hf, h);
          tmem_st8(t + 16u * hf + 8u, l);
        }
        tmem_st_wait();
        tc_fence_before();
        mbar_arrive(&b[L4_A2FULL + buf]);
        if (p == a1_pass && has_next) load_a1(tile + stride, u + 1);
      }
      finish(tile, u);
    }
    if (dft && prev_tile >= 0) {      // the last tile's W-DFT accumulator
#pragma unroll
      for (int hf = 0; hf < 2; ++hf) {
        id.half = hf;
        drain_d3(id, prev_tile, g, a.Aout, tmem_w + 96u, &b[L4_D3FULL], u - 1);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == kL4Iss) tmem_dealloc(tmem_s, 512);
}

int launch_tc_lift(const TcLiftArgs& a, cudaStream_t st) {
  int sms = 0;
  DLWP_TRY(current_sms(&sms));
  DLWP_TRY(ensure_dynamic_smem((const void*)fno_lift_kernel, LiftSmem::Total));
  DLWP_REQUIRE(a.hid == 128 || a.hid == 256, "tensor-core lifting needs width 128 or 256, got %d", a.hid);
  DLWP_REQUIRE(a.Cin >= 1 && a.Cin <= 15 && a.ngrp >= 1, "tensor-core lifting needs 1..15 input channels, got %d", a.Cin);
  // tile t of a CTA's list goes to worker (t / grid) % 2: with fewer tiles than SMs every tile has an SM (and its issue slots) to itself
  const int grid = a.g.ntiles < sms ? a.g.ntiles : sms;
  // four workers per CTA once each of them has a tile (a.workers4: plan option, DLWP_LIFT4=0 keeps the two-worker kernel)
  if (a.workers4 && a.g.ntiles >= 4 * sms) {
    DLWP_TRY(ensure_dynamic_smem((const void*)fno_lift4_kernel, Lift4Smem::Total));
    DLWP_CUDA(launch_pdl(fno_lift4_kernel, dim3(grid), dim3(kL4Cta), (size_t)Lift4Smem::Total, st, a));
    DLWP_LAUNCHED("fno_lift4_kernel");
    return DLWP_OK;
  }
  DLWP_CUDA(launch_pdl(fno_lift_kernel, dim3(grid), dim3(kCtaLayer), (size_t)LiftSmem::Total, st, a));
  DLWP_LAUNCHED("fno_lift_kernel");
  return DLWP_OK;
}

// =============================================================================================
// host-side packing of constant operands into chunk-major bf16 images
// =============================================================================================
namespace {
inline uint16_t f2bf(float f) {
  uint32_t u;
  std::memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return 0x7fc0;
  u += 0x7fffu + ((u >> 16) & 1u);
  return (uint16_t)(u >> 16);
}
inline float bf2f(uint16_t h) {
  const uint32_t u = (uint32_t)h << 16;
  float f;
  std::memcpy(&f, &u, 4);
  return f;
}
inline size_t img_off(int r, int k, int R) { return (size_t)(k >> 3) * R * 16 + (size_t)r * 16 + (size_t)(k & 7) * 2; }
inline void put(uint8_t* img, int r, int k, int R, uint16_t v) { std::memcpy(img + img_off(r, k, R), &v, 2); }
}  // namespace

void tc_pack_images_f64(const double* m, int rows, int cols, int ld, int R, int K, uint8_t* out) {
  const size_t part = (size_t)R * K * 2;
  std::memset(out, 0, 2 * part);
  for (int r = 0; r < rows; ++r)
    for (int k = 0; k < cols; ++k) {
      const double v = m[(size_t)r * ld + k];
      const uint16_t h = f2bf((float)v);
      const uint16_t l = f2bf((float)(v - (double)bf2f(h)));
      put(out, r, k, R, h);
      put(out + part, r, k, R, l);
    }
}

void tc_pack_images(const float* m, int rows, int cols, int ld, int R, int K, uint8_t* out) {
  const size_t part = (size_t)R * K * 2;
  std::memset(out, 0, 2 * part);
  for (int r = 0; r < rows; ++r)
    for (int k = 0; k < cols; ++k) {
      const float v = m[(size_t)r * ld + k];
      const uint16_t h = f2bf(v);
      const uint16_t l = f2bf(v - bf2f(h));
      put(out, r, k, R, h);
      put(out + part, r, k, R, l);
    }
}

void tc_pack_bias_image(const float* bias, int n, int R, uint8_t* out) {
  std::memset(out, 0, (size_t)R * 16 * 2);
  for (int r = 0; r < n; ++r) {
    const uint16_t h = f2bf(bias[r]);
    put(out, r, 0, R, h);
    put(out, r, 1, R, f2bf(bias[r] - bf2f(h)));
  }
}

void tc_pack_stacked_f64(const double* m, int rows, int cols, int ld, int R, int K, uint8_t* out) {
  std::memset(out, 0, (size_t)R * K * 2);
  const int half = R / 2;
  for (int r = 0; r < rows && r < half; ++r)
    for (int k = 0; k < cols; ++k) {
      const double v = m[(size_t)r * ld + k];
      const uint16_t h = f2bf((float)v);
      const uint16_t l = f2bf((float)(v - (double)bf2f(h)));
      put(out, r, k, R, h);
      put(out, half + r, k, R, l);
    }
}

}  // namespace dlwp

#ifdef DLWP_TRACE
extern "C" int dlwp_debug_set_trace_mlp(unsigned long long* p) {
  return cudaMemcpyToSymbol(dlwp::g_trace_mlp, &p, sizeof(p)) == cudaSuccess ? 0 : -2;
}
extern "C" int dlwp_debug_set_trace(unsigned long long* p) {
  return cudaMemcpyToSymbol(dlwp::g_trace, &p, sizeof(p)) == cudaSuccess ? 0 : -2;
}
#endif
