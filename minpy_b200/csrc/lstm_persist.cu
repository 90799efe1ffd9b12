// lstm_temporal (nn/layers.py:383-420) forward as ONE persistent launch for all T time steps.
//
// The launch-per-step path of lstm.cu streams Wh (H x 4H, 16.8 MB at H = 1024) from L2 for every step and pays
// two launches per step (10.5 us per step on C5: 7 us GEMM + 3.5 us gate kernel, profiles/r2m).  Here Wh never
// moves after the first microsecond: H/8 CTAs (128 at H = 1024, one per SM) form H/32 clusters of four; cluster c
// owns hidden units [32c, 32c+32) = 128 gate columns, CTA q of the cluster owns the K slice [q H/4, (q+1) H/4) of
// those columns, H/4 x 128 floats = 128 KB of shared memory at H = 1024, loaded once.  Per time step a CTA
//   1. waits for the H/32 CTAs that produce its K slice of h_{t-1} (one flag word each, release / acquire),
//   2. streams that slice (128 rows x H/4, 128 KB) through a four-stage TMA ring into tcgen05 MMAs
//      (M = 128 batch rows, N = 128 gate columns, accumulator in TMEM),
//   3. parks the partial tile in its own shared memory (the ring is idle by then), and after a cluster barrier
//      sums rows [32q, 32q+32) of the four partials through distributed shared memory in slice order 0..3,
//   4. adds the input projection a_t = x_t . Wx + b (fetched under the MMAs), applies the gate arithmetic of
//      lstm_step (layers.py:353-380: i, f, o sigmoid, g tanh, c' = f c + i g, h' = o tanh(c')), keeps c in
//      registers, writes h_t / c_t / gates for the backward pass and publishes its flag.
// Per step and CTA: 128 KB from L2 instead of 256 KB, no partial tile in HBM, no launch.
#include <initializer_list>
#include <stdlib.h>
#include "tc_common.cuh"

namespace mc {
namespace tc {

constexpr int LP_THREADS = 256;
constexpr int LP_S = 4;                       // CTAs per cluster = K slices
constexpr int LP_STAGES = 6;
constexpr uint32_t LP_TILE = 128 * 32 * 4;    // one k-block of an operand: 128 rows x 128 B
constexpr uint32_t LP_RING = LP_STAGES * LP_TILE;   // 96 KB operand ring; its first 64 KB later hold the parked partial (128 x 128 fp32)
constexpr int LP_MAX_KB = 8;                  // k-blocks per CTA and step (H / 128)

struct LstmPersistParams {
  const float* ax;     // (T, N, 4H) input projections, gate-major columns [i | f | o | g]
  float* hbuf;         // (T + 1, N, H), slot 0 = h0
  float* cbuf;         // (T + 1, N, H), slot 0 = c0
  float* gates;        // (T, N, 4H)
  int* flags;          // one word per CTA: number of hidden states it has published (zero at launch)
  int T, N, H;
};

// Wp[(32c.. packed column) , k] = Wh[k, g H + 32 c + ul], packed column = 128 c + 32 g + ul: the 128 gate columns of
// cluster c side by side, K-major (k contiguous) so that one TMA box (32 k x 128 columns) is an UMMA B tile.
__global__ void __launch_bounds__(256) lstm_pack_wh_kernel(const float* __restrict__ Wh, float* __restrict__ Wp, int H) {
  mc::pdl_sync();
  __shared__ float tile[32][33];
  const int k0 = blockIdx.x * 32;             // 32 k's
  const int strip = blockIdx.y;               // 32 packed columns: strip = 4 c + g
  const int c = strip >> 2, g = strip & 3;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;     // 8 rows a pass
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int k = k0 + ty + 8 * i;
    tile[ty + 8 * i][tx] = Wh[(int64_t)k * 4 * H + (int64_t)g * H + 32 * c + tx];
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int ul = ty + 8 * i;
    Wp[(int64_t)(128 * c + 32 * g + ul) * H + k0 + tx] = tile[tx][ul];
  }
}

__device__ __forceinline__ int ld_acquire_gpu(const int* p) {
  int v;
  asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_gpu(int* p, int v) {
  asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
// Gate non-linearities on the special-function unit: 16 of them per thread sit on the critical path of every time
// step (the accurate expf / tanhf / IEEE division versions cost 1.0 us per step).  Absolute error ~1e-7, three orders
// below the TF32 rounding of the recurrent product they are applied to.
__device__ __forceinline__ float lp_sigmoid(float x) { return __fdividef(1.0f, 1.0f + __expf(-x)); }   // layers.py:243
__device__ __forceinline__ float lp_tanh(float x) {
  const float e = __expf(-2.0f * fabsf(x));                  // in (0, 1]: no overflow
  return copysignf(__fdividef(1.0f - e, 1.0f + e), x);
}

MC_PROBE_DEFINE(lstmp)   // tools/probe.sh: %globaltimer stamps of the second-to-last time step
#define LP_STAMP(i) do { if (t == p.T - 2) { MC_STAMP(lstmp, i); } } while (0)
__global__ void __launch_bounds__(LP_THREADS, 1) lstm_persist_fwd_kernel(const __grid_constant__ CUtensorMap tmH,
                                                                         const __grid_constant__ CUtensorMap tmW,
                                                                         const LstmPersistParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t ring = smem_base;                                   // [0, 64 KB)
  const uint32_t wres = smem_base + LP_RING;                         // resident Wh slice: kb_per tiles
  const int kb_per = p.H >> 7;                                       // (H / 4) / 32
  const uint32_t bar_base = wres + (uint32_t)kb_per * LP_TILE;       // full[4], empty[4], tmem_full, w_full
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (LP_STAGES + s); };
  const uint32_t tmem_full_bar = bar_base + 8u * (2 * LP_STAGES);
  const uint32_t w_full_bar = tmem_full_bar + 8u;
  const uint32_t tmem_slot = w_full_bar + 8u;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint32_t rank;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
  const int cluster = blockIdx.x / LP_S;
  const int k_base = (int)rank * (p.H >> 2);
  const int nprod = p.H >> 5;                                        // CTAs that produce this K slice of h

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmH)) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmW)) : "memory");
    for (int s = 0; s < LP_STAGES; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
    mbar_init(tmem_full_bar, 1);
    mbar_init(w_full_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(128u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  mc::pdl_sync();
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot) : "memory");

  // the resident Wh slice: kb_per boxes of 32 k x 128 packed columns
  if (warp == 0 && elect_one()) {
    mbar_expect_tx(w_full_bar, (uint32_t)kb_per * LP_TILE);
    for (int kb = 0; kb < kb_per; ++kb) tma_load_2d(wres + (uint32_t)kb * LP_TILE, &tmW, w_full_bar, k_base + 32 * kb, 128 * cluster);
  }
  __syncwarp();

  // this thread's cells: batch row 32 rank + r, hidden units unit0 .. unit0 + 3
  const int r = threadIdx.x >> 3, j = threadIdx.x & 7;
  const int trow = (int)rank * 32 + r;                  // row inside the 128-row tile = batch row
  const bool live = trow < p.N;
  const int unit0 = 32 * cluster + 4 * j;
  const int64_t H = p.H, sH = (int64_t)p.N * H, sA = 4 * sH;
  const int64_t cell = (int64_t)trow * H + unit0, gcell = (int64_t)trow * 4 * H + unit0;
  float4 c = make_float4(0.f, 0.f, 0.f, 0.f);
  if (live) c = *reinterpret_cast<const float4*>(p.cbuf + cell);
  // peers in the order rank, rank + 1, ...: at any moment the four CTAs of a cluster read four DIFFERENT shared
  // memories (distributed shared memory is served by the owner's port at ~20 B/clk); the order is fixed per CTA, so
  // the sums stay deterministic
  uint32_t peer[LP_S];
#pragma unroll
  for (int z = 0; z < LP_S; ++z)
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(peer[z]) : "r"(ring), "r"((rank + (uint32_t)z) & (LP_S - 1)));

  const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(128 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
  const uint64_t desc0 = make_desc(0, 16, 1024, 2);
  int s = 0;                 // ring position and phase: the producer thread and the MMA thread each carry their own copy
  uint32_t ph = 0;
  const int quad = warp & 3, chalf = warp >> 2;
  const int prow = quad * 32 + lane;                    // accumulator row this thread parks
  const bool leader = elect_one();                      // the same lane of each warp for the whole sequence

  for (int t = 0; t < p.T; ++t) {
    LP_STAMP(0);
    // input projection of this step: in flight while the recurrent product is computed
    float4 a0 = make_float4(0.f, 0.f, 0.f, 0.f), a1 = a0, a2 = a0, a3 = a0;
    if (live) {
      const float* ap = p.ax + (int64_t)t * sA + gcell;
      a0 = __ldcs(reinterpret_cast<const float4*>(ap));
      a1 = __ldcs(reinterpret_cast<const float4*>(ap + H));
      a2 = __ldcs(reinterpret_cast<const float4*>(ap + 2 * H));
      a3 = __ldcs(reinterpret_cast<const float4*>(ap + 3 * H));
    }
    if (warp == 0) {
      // k-block kb of this CTA's K slice = the 32 hidden units of cluster (rank H/128 + kb): complete once that
      // cluster's four CTAs (flags 4 kb .. 4 kb + 3 of the slice) have published t states.  Every lane polls its own
      // producer (one L2 round trip per poll, all in parallel); the loads go out in order as their four flags arrive.
      // (The proxy fence between the producers' generic-proxy stores and these async-proxy reads is on the WRITER's
      //  side, before its flag: it costs ~0.6 us, and a fence per acquire here put 8 of them on the critical path.)
      bool seen = (t == 0) || (lane >= nprod);
      const int* f = p.flags + (int)rank * nprod + (lane < nprod ? lane : 0);
      int issued = 0;
      long long t0 = 0;
      while (issued < kb_per) {
        if (!seen) seen = ld_acquire_gpu(f) >= t;
        const uint32_t ready = __ballot_sync(0xffffffffu, seen);
        if (issued == 0) LP_STAMP(1);
        while (issued < kb_per && ((ready >> (4 * issued)) & 0xFu) == 0xFu) {
          if (leader) {
            mbar_wait(empty_bar(s), ph ^ 1u);
            mbar_expect_tx(full_bar(s), LP_TILE);
            tma_load_2d(ring + (uint32_t)s * LP_TILE, &tmH, full_bar(s), k_base + 32 * issued, t * p.N);
            if (++s == LP_STAGES) { s = 0; ph ^= 1u; }
          }
          ++issued;
        }
        if (ready != 0xffffffffu) {            // a missing CTA must be an error, never a hung GPU
          if (t0 == 0) t0 = clock64();
          else if (clock64() - t0 > 2000000000LL) __trap();
        }
      }
      __syncwarp();
      LP_STAMP(2);
    } else if (warp == 1 && leader) {
      if (t == 0) mbar_wait(w_full_bar, 0);
      for (int kb = 0; kb < kb_per; ++kb) {
        mbar_wait(full_bar(s), ph);
        tcgen05_fence_after();
        const uint64_t adesc = desc0 + (uint64_t)(((ring + (uint32_t)s * LP_TILE) & 0x3FFFFu) >> 4);
        const uint64_t bdesc = desc0 + (uint64_t)(((wres + (uint32_t)kb * LP_TILE) & 0x3FFFFu) >> 4);
#pragma unroll
        for (int k = 0; k < 4; ++k) umma_tf32(tmem_base, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc, (kb > 0 || k > 0) ? 1u : 0u);
        umma_commit(empty_bar(s));
        if (++s == LP_STAGES) { s = 0; ph ^= 1u; }
      }
      umma_commit(tmem_full_bar);
    }
    __syncwarp();
    mbar_wait_warp(tmem_full_bar, (uint32_t)(t & 1));
    tcgen05_fence_after();
    __syncwarp();
    LP_STAMP(3);
    // park the partial tile: row r at r * 512 B, its 16-byte chunk j at slot j ^ (r & 7)
#pragma unroll
    for (int cc = 0; cc < 2; ++cc) {
      const int c0 = chalf * 64 + cc * 32;
      float v[32];
      tmem_ld32(tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)c0, v);
      const uint32_t rbase = ring + (uint32_t)prow * 512u;
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        const uint32_t slot = (uint32_t)((c0 >> 2) + q) ^ (uint32_t)(prow & 7);
        asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(rbase + slot * 16u), "f"(v[4 * q]), "f"(v[4 * q + 1]),
                     "f"(v[4 * q + 2]), "f"(v[4 * q + 3]) : "memory");
      }
    }
    tcgen05_fence_before();
    LP_STAMP(4);
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
    LP_STAMP(5);
    // recurrent pre-activations of this thread's four units: gate g sits in packed columns 32 g + 4 j .. + 3
    float4 part[4][LP_S];
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      const uint32_t off = (uint32_t)trow * 512u + (uint32_t)((8 * g + j) ^ (trow & 7)) * 16u;
#pragma unroll
      for (int z = 1; z < LP_S; ++z)
        asm volatile("ld.shared::cluster.v4.f32 {%0, %1, %2, %3}, [%4];"
                     : "=f"(part[g][z].x), "=f"(part[g][z].y), "=f"(part[g][z].z), "=f"(part[g][z].w) : "r"(peer[z] + off) : "memory");
    }
    // this CTA's own partial through the local port: a quarter of the bytes off the ~20 B/clk cluster path
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      const uint32_t off = (uint32_t)trow * 512u + (uint32_t)((8 * g + j) ^ (trow & 7)) * 16u;
      asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                   : "=f"(part[g][0].x), "=f"(part[g][0].y), "=f"(part[g][0].z), "=f"(part[g][0].w) : "r"(ring + off) : "memory");
    }
    float4 pre[4] = {a0, a1, a2, a3};
#pragma unroll
    for (int g = 0; g < 4; ++g) {
#pragma unroll
      for (int z = 0; z < LP_S; ++z) { pre[g].x += part[g][z].x; pre[g].y += part[g][z].y; pre[g].z += part[g][z].z; pre[g].w += part[g][z].w; }
    }
    // peers may overwrite their ring (= the parked tiles) only after every CTA of the cluster has read them: the loads
    // above have returned (their values were just added), so the arrival needs no release fence (it cost 1.4 us)
    asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory");
    LP_STAMP(6);
    float4 gi, gf, go, gg, h;
#define LP_CELL(m)                                                                                             \
    gi.m = lp_sigmoid(pre[0].m); gf.m = lp_sigmoid(pre[1].m); go.m = lp_sigmoid(pre[2].m); gg.m = lp_tanh(pre[3].m); \
    c.m = gf.m * c.m + gi.m * gg.m; h.m = go.m * lp_tanh(c.m);
    LP_CELL(x) LP_CELL(y) LP_CELL(z) LP_CELL(w)
#undef LP_CELL
    // h first: it is what the other CTAs wait for; c and the gates (read by the backward pass) follow the flag
    if (live) *reinterpret_cast<float4*>(p.hbuf + (int64_t)(t + 1) * sH + cell) = h;
    LP_STAMP(7);
    __syncthreads();
    LP_STAMP(8);
    if (threadIdx.x == 0) {
      // the block barrier ordered every thread's h stores before this point; the proxy fence orders them against the
      // consumers' TMA (async-proxy) reads; st.release carries the gpu-scope fence (a separate __threadfence cost 0.5 us)
      asm volatile("fence.proxy.async.global;" ::: "memory");
      LP_STAMP(9);
      st_release_gpu(p.flags + blockIdx.x, t + 1);
    }
    LP_STAMP(10);
    if (live) {
      *reinterpret_cast<float4*>(p.cbuf + (int64_t)(t + 1) * sH + cell) = c;
      float* gp = p.gates + (int64_t)t * sA + gcell;
      __stcs(reinterpret_cast<float4*>(gp), gi);
      __stcs(reinterpret_cast<float4*>(gp + H), gf);
      __stcs(reinterpret_cast<float4*>(gp + 2 * H), go);
      __stcs(reinterpret_cast<float4*>(gp + 3 * H), gg);
    }
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
    LP_STAMP(11);
  }
  tcgen05_fence_before();
  __syncthreads();
  if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(128u) : "memory");
}


// ---------------------------------------------------------------------------------------------------------
// Backward (BPTT of lstm_step, numpy_core VJPs of layers.py:353-380) as one persistent launch.
// The recurrent product of the backward pass, dA_t . Wh^T, has K = 4H and only H output columns: four times the
// broadcast volume of the forward pass for a quarter of the output.  Clusters of EIGHT CTAs: cluster c owns the BN
// hidden units [BN c, BN c + BN) (rows of Wh, which as stored IS the K-major B operand -- no packing), CTA q owns
// the K slice [q H/2, (q+1) H/2) of the 4H gate columns (gate q / 2, one half of its units).  Per step: wait for
// the CTAs that produced this slice of dA_{t+1}, stream it (128 rows x H/2) into M = 128 / N = BN MMAs, park the
// partial, cluster barrier, sum rows [16q, 16q+16) of the eight partials through distributed shared memory, add
// the upstream gradient dH[t+1], run the gate backward for those 16 rows x BN units (dc stays in registers),
// write dA_t, publish.  One float4 of units per thread: 4 BN threads.
//   BN = 64: H/64 clusters, 128 KB of Wh resident at H = 1024, six-stage ring.  This pool's B200 holds 15 clusters
//            of eight (cudaOccupancyMaxActiveClusters), so H = 1024 (16 clusters) does not fit.
//   BN = 80: ceil(H/80) clusters (13 at H = 1024; the last one owns 64 live units, the rest of its tile is TMA
//            zero fill), 160 KB resident, four-stage ring, 320 threads.  A k-block of dA may then straddle the
//            units of two clusters: it is loaded once both have published.
constexpr int LB_S = 8;
constexpr int LB_MAX_KB = 16;

struct LstmPersistBwdParams {
  const float* gates;  // (T, N, 4H)
  const float* cbuf;   // (T + 1, N, H)
  float* dH;           // (T + 1, N, H): slot t + 1 = upstream gradient of h_t; slot 0 receives dL/dh0
  float* dA;           // (T, N, 4H) out
  int* flags;
  int T, N, H;
};

__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float* v) {
  uint32_t* r = reinterpret_cast<uint32_t*>(v);
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

MC_PROBE_DEFINE(lstmb)
#define LB_STAMP(i) do { if (it == p.T - 2) { MC_STAMP(lstmb, i); } } while (0)
template <int BN, int STAGES>
__global__ void __launch_bounds__(4 * BN, 1) lstm_persist_bwd_kernel(const __grid_constant__ CUtensorMap tmD,
                                                                     const __grid_constant__ CUtensorMap tmW,
                                                                     const LstmPersistBwdParams p) {
  constexpr uint32_t BTILE = BN * 32 * 4;            // one k-block of the resident operand: BN units x 128 B
  constexpr uint32_t RING = STAGES * LP_TILE;
  constexpr int CH = BN / 4;                         // 16-byte chunks per parked row
  constexpr uint32_t ROWB = BN * 4;                  // bytes per parked row
  constexpr uint32_t TMEM_COLS = BN <= 64 ? 64 : 128;
  static_assert(128 * ROWB <= RING, "the parked partial lives in the idle ring");
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t ring = smem_base;
  const uint32_t wres = smem_base + RING;
  const int kb_per = p.H >> 6;                                       // (H / 2) / 32
  const uint32_t bar_base = wres + (uint32_t)kb_per * BTILE;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (STAGES + s); };
  const uint32_t tmem_full_bar = bar_base + 8u * (2 * STAGES);
  const uint32_t w_full_bar = tmem_full_bar + 8u;
  const uint32_t tmem_slot = w_full_bar + 8u;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint32_t rank;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
  const int cluster = blockIdx.x / LB_S;
  const int k_base = (int)rank * (p.H >> 1);
  const int slice_u0 = (int)(rank & 1u) * (p.H >> 1);                // first hidden unit of this K slice
  // producers of the slice: clusters c_first .. c_last, flag words 8 c .. 8 c + 7 (= blockIdx.x of their CTAs)
  const int c_first = slice_u0 / BN, c_last = (slice_u0 + (p.H >> 1) - 1) / BN;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmD)) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmW)) : "memory");
    for (int s = 0; s < STAGES; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
    mbar_init(tmem_full_bar, 1);
    mbar_init(w_full_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(TMEM_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  mc::pdl_sync();
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot) : "memory");

  if (warp == 0 && elect_one()) {      // the resident slice of Wh: kb_per boxes of 32 k x BN units (rows >= H: zero fill)
    mbar_expect_tx(w_full_bar, (uint32_t)kb_per * BTILE);
    for (int kb = 0; kb < kb_per; ++kb) tma_load_2d(wres + (uint32_t)kb * BTILE, &tmW, w_full_bar, k_base + 32 * kb, BN * cluster);
  }
  __syncwarp();

  // this thread's cells: batch row 16 rank + r, hidden units unit0 .. unit0 + 3
  const int r = threadIdx.x / CH, j = threadIdx.x % CH;
  const int trow = (int)rank * 16 + r;
  const int unit0 = BN * cluster + 4 * j;
  const bool live = trow < p.N && unit0 < p.H;
  const int64_t H = p.H, sH = (int64_t)p.N * H, sA = 4 * sH;
  const int64_t cell = (int64_t)trow * H + unit0, gcell = (int64_t)trow * 4 * H + unit0;
  float4 dcn = make_float4(0.f, 0.f, 0.f, 0.f);          // dc flowing in from the later time step
  uint32_t peer[LB_S];
#pragma unroll
  for (int z = 0; z < LB_S; ++z)
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(peer[z]) : "r"(ring), "r"((rank + (uint32_t)z) & (LB_S - 1)));

  const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
  const uint64_t desc0 = make_desc(0, 16, 1024, 2);
  int s = 0;
  uint32_t ph = 0;
  const int quad = warp & 3, chalf = warp >> 2;
  const int prow = quad * 32 + lane;
  const bool leader = elect_one();
  const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);

  // iteration `it` runs the recurrent product of the step after t = T - 1 - it (none at it = 0) and the gate
  // backward of step t; the last iteration (it = T) only has the product: dH[0] += dA_0 . Wh^T
  for (int it = 0; it <= p.T; ++it) {
    const int t = p.T - 1 - it;
    LB_STAMP(0);
    float4 gi = zero4, gf = zero4, go = zero4, gg = zero4, cn = zero4, cp = zero4, dh = zero4;
    if (live) {
      if (it < p.T) {
        const float* gp = p.gates + (int64_t)t * sA + gcell;
        gi = __ldcs(reinterpret_cast<const float4*>(gp));
        gf = __ldcs(reinterpret_cast<const float4*>(gp + H));
        go = __ldcs(reinterpret_cast<const float4*>(gp + 2 * H));
        gg = __ldcs(reinterpret_cast<const float4*>(gp + 3 * H));
        cn = __ldcs(reinterpret_cast<const float4*>(p.cbuf + (int64_t)(t + 1) * sH + cell));
        cp = __ldcs(reinterpret_cast<const float4*>(p.cbuf + (int64_t)t * sH + cell));
        dh = __ldcs(reinterpret_cast<const float4*>(p.dH + (int64_t)(t + 1) * sH + cell));
      } else {
        dh = *reinterpret_cast<const float4*>(p.dH + cell);
      }
    }
    if (it > 0) {
      if (warp == 0) {
        // Lane l polls the flags l, l + 32, l + 64, l + 96 that belong to this slice's producers; a k-block goes
        // out once the one or two clusters whose units it covers have published `it` gradients (all eight CTAs).
        bool seen[4];
        const int* fp[4];
#pragma unroll
        for (int w = 0; w < 4; ++w) {
          const int f = 32 * w + lane;
          const bool relevant = f >= 8 * c_first && f < 8 * (c_last + 1);
          seen[w] = !relevant;
          fp[w] = p.flags + (relevant ? f : 0);
        }
        int issued = 0;
        long long t0 = 0;
        while (issued < kb_per) {
          uint32_t rd[4];
#pragma unroll
          for (int w = 0; w < 4; ++w) {
            if (!seen[w]) seen[w] = ld_acquire_gpu(fp[w]) >= it;
            rd[w] = __ballot_sync(0xffffffffu, seen[w]);
          }
          if (issued == 0) LB_STAMP(1);
          while (issued < kb_per) {
            const int u_lo = slice_u0 + 32 * issued;
            const int c_lo = u_lo / BN, c_hi = (u_lo + 31) / BN;
            bool ok = true;
            for (int c = c_lo; c <= c_hi; ++c) {
              const int w = c >> 2;
              const uint32_t word = w == 0 ? rd[0] : (w == 1 ? rd[1] : (w == 2 ? rd[2] : rd[3]));
              ok = ok && (((word >> (8 * (c & 3))) & 0xFFu) == 0xFFu);
            }
            if (!ok) break;
            if (leader) {
              mbar_wait(empty_bar(s), ph ^ 1u);
              mbar_expect_tx(full_bar(s), LP_TILE);
              tma_load_2d(ring + (uint32_t)s * LP_TILE, &tmD, full_bar(s), k_base + 32 * issued, (t + 1) * p.N);
              if (++s == STAGES) { s = 0; ph ^= 1u; }
            }
            ++issued;
          }
          if ((rd[0] & rd[1] & rd[2] & rd[3]) != 0xffffffffu) {          // a missing CTA must be an error, never a hung GPU
            if (t0 == 0) t0 = clock64();
            else if (clock64() - t0 > 2000000000LL) __trap();
          }
        }
        __syncwarp();
        LB_STAMP(2);
      } else if (warp == 1 && leader) {
        if (it == 1) mbar_wait(w_full_bar, 0);
        for (int kb = 0; kb < kb_per; ++kb) {
          mbar_wait(full_bar(s), ph);
          tcgen05_fence_after();
          const uint64_t adesc = desc0 + (uint64_t)(((ring + (uint32_t)s * LP_TILE) & 0x3FFFFu) >> 4);
          const uint64_t bdesc = desc0 + (uint64_t)(((wres + (uint32_t)kb * BTILE) & 0x3FFFFu) >> 4);
#pragma unroll
          for (int k = 0; k < 4; ++k) umma_tf32(tmem_base, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc, (kb > 0 || k > 0) ? 1u : 0u);
          umma_commit(empty_bar(s));
          if (++s == STAGES) { s = 0; ph ^= 1u; }
        }
        umma_commit(tmem_full_bar);
      }
      __syncwarp();
      mbar_wait_warp(tmem_full_bar, (uint32_t)((it - 1) & 1));
      tcgen05_fence_after();
      __syncwarp();
      LB_STAMP(3);
      if (warp < 8) {   // park the partial tile (128 x BN): row r at r * ROWB, its 16-byte chunk j at slot (j + r) % CH
        const uint32_t rbase = ring + (uint32_t)prow * ROWB;
        {
          const int c0 = chalf * 32;
          float v[32];
          tmem_ld32(tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)c0, v);
#pragma unroll
          for (int q = 0; q < 8; ++q) {
            const uint32_t slot = (uint32_t)((c0 >> 2) + q + prow) % (uint32_t)CH;
            asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(rbase + slot * 16u), "f"(v[4 * q]), "f"(v[4 * q + 1]),
                         "f"(v[4 * q + 2]), "f"(v[4 * q + 3]) : "memory");
          }
        }
        if (BN > 64 && chalf == 0) {       // columns 64 .. BN - 1
          float v[16];
          tmem_ld16(tmem_base + ((uint32_t)(quad * 32) << 16) + 64u, v);
#pragma unroll
          for (int q = 0; q < (BN - 64) / 4; ++q) {
            const uint32_t slot = (uint32_t)(16 + q + prow) % (uint32_t)CH;
            asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(rbase + slot * 16u), "f"(v[4 * q]), "f"(v[4 * q + 1]),
                         "f"(v[4 * q + 2]), "f"(v[4 * q + 3]) : "memory");
          }
        }
      }
      tcgen05_fence_before();
      LB_STAMP(4);
      asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
      LB_STAMP(5);
      const uint32_t off = (uint32_t)trow * ROWB + ((uint32_t)(j + trow) % (uint32_t)CH) * 16u;
      float4 part[LB_S];
#pragma unroll
      for (int z = 1; z < LB_S; ++z)
        asm volatile("ld.shared::cluster.v4.f32 {%0, %1, %2, %3}, [%4];"
                     : "=f"(part[z].x), "=f"(part[z].y), "=f"(part[z].z), "=f"(part[z].w) : "r"(peer[z] + off) : "memory");
      asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                   : "=f"(part[0].x), "=f"(part[0].y), "=f"(part[0].z), "=f"(part[0].w) : "r"(ring + off) : "memory");
#pragma unroll
      for (int z = 0; z < LB_S; ++z) { dh.x += part[z].x; dh.y += part[z].y; dh.z += part[z].z; dh.w += part[z].w; }
      asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory");
      LB_STAMP(6);
    }
    if (it < p.T) {
      float4 d0, d1, d2, d3;
#define LB_CELL(m)                                                          \
      {                                                                     \
        const float tc = lp_tanh(cn.m);                                     \
        const float dc = dh.m * go.m * (1.f - tc * tc) + dcn.m;             \
        d0.m = dc * gg.m * gi.m * (1.f - gi.m);                             \
        d1.m = dc * cp.m * gf.m * (1.f - gf.m);                             \
        d2.m = dh.m * tc * go.m * (1.f - go.m);                             \
        d3.m = dc * gi.m * (1.f - gg.m * gg.m);                             \
        dcn.m = dc * gf.m;                                                  \
      }
      LB_CELL(x) LB_CELL(y) LB_CELL(z) LB_CELL(w)
#undef LB_CELL
      if (live) {
        float* dp = p.dA + (int64_t)t * sA + gcell;
        *reinterpret_cast<float4*>(dp) = d0;
        *reinterpret_cast<float4*>(dp + H) = d1;
        *reinterpret_cast<float4*>(dp + 2 * H) = d2;
        *reinterpret_cast<float4*>(dp + 3 * H) = d3;
      }
      LB_STAMP(7);
      __syncthreads();
      if (threadIdx.x == 0) {
        asm volatile("fence.proxy.async.global;" ::: "memory");
        st_release_gpu(p.flags + blockIdx.x, it + 1);
      }
      LB_STAMP(8);
    } else if (live) {
      *reinterpret_cast<float4*>(p.dH + cell) = dh;          // dL/dh0
    }
    if (it > 0) asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
    LB_STAMP(9);
  }
  tcgen05_fence_before();
  __syncthreads();
  if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS) : "memory");
}


// ---------------------------------------------------------------------------------------------------------
// Backward recurrence without clusters (H = 1024, where 16 eight-CTA clusters do not fit the device): one persistent
// launch of H/8 independent CTAs that STREAM Wh from L2 (16.8 MB, resident there) instead of keeping it in shared
// memory -- what is saved against the launch-per-step path is the two launches per step, the launch ramps and the
// round trip of the split-K partials through a separate gate kernel.
//   CTA (j, s): output units [128 j, 128 j + 128), K slice s of 16 = gate columns [s H/4, (s+1) H/4).
//   per step   wait for the CTAs that produced this slice of dA_{t+1} (flags, as above) -> TMA ring of
//              {dA tile, Wh tile} pairs (32 KB a stage, six stages) -> M128 x N128 MMAs -> the partial tile (64 KB)
//              goes to a workspace in L2, flag -> wait for the 16 partials of unit tile j -> sum rows [8 s, 8 s + 8)
//              in slice order -> + dH[t + 1], gate backward (dc in registers) -> dA_t, flag.
// The workspace is double-buffered by step parity: a partial of step i is overwritten at step i + 2, and no CTA can
// reach step i + 2 before every CTA has published dA of step i, i.e. finished reading the partials of step i.
constexpr int LS2_S = 16;
constexpr int LS2_STAGES = 6;
constexpr uint32_t LS2_STAGE = 2 * LP_TILE;        // dA tile + Wh tile

struct LstmStreamBwdParams {
  const float* gates;
  const float* cbuf;
  float* dH;
  float* dA;
  float* ws;           // [2][CTAs][128][128]
  int* flag_da;        // per CTA: number of dA steps published
  int* flag_part;      // per CTA: number of partial tiles published
  int T, N, H;
};

MC_PROBE_DEFINE(lstms)
#define LS2_STAMP(i) do { if (it == p.T - 2) { MC_STAMP(lstms, i); } } while (0)
__global__ void __launch_bounds__(LP_THREADS, 1) lstm_stream_bwd_kernel(const __grid_constant__ CUtensorMap tmD,
                                                                        const __grid_constant__ CUtensorMap tmW,
                                                                        const LstmStreamBwdParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t ring = smem_base;
  const uint32_t bar_base = ring + LS2_STAGES * LS2_STAGE;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (LS2_STAGES + s); };
  const uint32_t tmem_full_bar = bar_base + 8u * (2 * LS2_STAGES);
  const uint32_t tmem_slot = tmem_full_bar + 8u;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int jt = blockIdx.x / LS2_S, sl = blockIdx.x % LS2_S;      // unit tile, K slice
  const int kb_per = p.H >> 7;                                       // (H / 4) / 32
  const int k_base = sl * (p.H >> 2);
  const int slice_u0 = (sl & 3) * (p.H >> 2);                        // first hidden unit of the slice
  const int jt_first = slice_u0 >> 7, jt_last = (slice_u0 + (p.H >> 2) - 1) >> 7;   // producer unit tiles (<= 2)

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmD)) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmW)) : "memory");
    for (int s = 0; s < LS2_STAGES; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
    mbar_init(tmem_full_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(128u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  mc::pdl_sync();
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot) : "memory");

  // this thread's cells: batch row 8 sl + r, hidden units unit0 .. unit0 + 3
  const int r = threadIdx.x >> 5;
  const int trow = sl * 8 + r;
  const int unit0 = 128 * jt + 4 * lane;
  const bool live = trow < p.N;
  const int64_t H = p.H, sH = (int64_t)p.N * H, sA = 4 * sH;
  const int64_t cell = (int64_t)trow * H + unit0, gcell = (int64_t)trow * 4 * H + unit0;
  float4 dcn = make_float4(0.f, 0.f, 0.f, 0.f);
  const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(128 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
  const uint64_t desc0 = make_desc(0, 16, 1024, 2);
  int s = 0;
  uint32_t ph = 0;
  const int quad = warp & 3, chalf = warp >> 2;
  const int prow = quad * 32 + lane;
  const bool leader = elect_one();
  const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
  const int64_t tile_elems = 128 * 128;
  const int ctas = gridDim.x;

  for (int it = 0; it <= p.T; ++it) {
    const int t = p.T - 1 - it;
    LS2_STAMP(0);
    float4 gi = zero4, gf = zero4, go = zero4, gg = zero4, cn = zero4, cp = zero4, dh = zero4;
    if (live) {
      if (it < p.T) {
        const float* gp = p.gates + (int64_t)t * sA + gcell;
        gi = __ldcs(reinterpret_cast<const float4*>(gp));
        gf = __ldcs(reinterpret_cast<const float4*>(gp + H));
        go = __ldcs(reinterpret_cast<const float4*>(gp + 2 * H));
        gg = __ldcs(reinterpret_cast<const float4*>(gp + 3 * H));
        cn = __ldcs(reinterpret_cast<const float4*>(p.cbuf + (int64_t)(t + 1) * sH + cell));
        cp = __ldcs(reinterpret_cast<const float4*>(p.cbuf + (int64_t)t * sH + cell));
        dh = __ldcs(reinterpret_cast<const float4*>(p.dH + (int64_t)(t + 1) * sH + cell));
      } else {
        dh = *reinterpret_cast<const float4*>(p.dH + cell);
      }
    }
    if (it > 0) {
      float* wsp = p.ws + ((int64_t)((it & 1) * ctas) + blockIdx.x) * tile_elems;
      if (warp == 0) {
        // lane l polls CTA (jt_first + l / 16, slice l % 16): the producers of this K slice of dA
        const int pj = jt_first + (lane >> 4);
        bool seen = pj > jt_last;
        const int* f = p.flag_da + (pj <= jt_last ? pj * LS2_S + (lane & 15) : 0);
        int issued = 0;
        long long t0 = 0;
        while (issued < kb_per) {
          if (!seen) seen = ld_acquire_gpu(f) >= it;
          const uint32_t ready = __ballot_sync(0xffffffffu, seen);
          if (issued == 0) LS2_STAMP(1);
          while (issued < kb_per) {
            const int pt = ((slice_u0 + 32 * issued) >> 7) - jt_first;          // producer tile of this k-block: 0 or 1
            if (((ready >> (16 * pt)) & 0xFFFFu) != 0xFFFFu) break;
            if (leader) {
              mbar_wait(empty_bar(s), ph ^ 1u);
              mbar_expect_tx(full_bar(s), LS2_STAGE);
              const uint32_t st = ring + (uint32_t)s * LS2_STAGE;
              tma_load_2d(st, &tmD, full_bar(s), k_base + 32 * issued, (t + 1) * p.N);
              tma_load_2d(st + LP_TILE, &tmW, full_bar(s), k_base + 32 * issued, 128 * jt);
              if (++s == LS2_STAGES) { s = 0; ph ^= 1u; }
            }
            ++issued;
          }
          if (ready != 0xffffffffu) {
            if (t0 == 0) t0 = clock64();
            else if (clock64() - t0 > 2000000000LL) __trap();
          }
        }
        __syncwarp();
        LS2_STAMP(2);
      } else if (warp == 1 && leader) {
        for (int kb = 0; kb < kb_per; ++kb) {
          mbar_wait(full_bar(s), ph);
          tcgen05_fence_after();
          const uint32_t st = ring + (uint32_t)s * LS2_STAGE;
          const uint64_t adesc = desc0 + (uint64_t)((st & 0x3FFFFu) >> 4);
          const uint64_t bdesc = desc0 + (uint64_t)(((st + LP_TILE) & 0x3FFFFu) >> 4);
#pragma unroll
          for (int k = 0; k < 4; ++k) umma_tf32(tmem_base, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc, (kb > 0 || k > 0) ? 1u : 0u);
          umma_commit(empty_bar(s));
          if (++s == LS2_STAGES) { s = 0; ph ^= 1u; }
        }
        umma_commit(tmem_full_bar);
      }
      __syncwarp();
      mbar_wait_warp(tmem_full_bar, (uint32_t)((it - 1) & 1));
      tcgen05_fence_after();
      __syncwarp();
      LS2_STAMP(3);
      // the partial tile to the workspace: this thread's row, 64 columns
#pragma unroll
      for (int cc = 0; cc < 2; ++cc) {
        const int c0 = chalf * 64 + cc * 32;
        float v[32];
        tmem_ld32(tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)c0, v);
        float4* dst = reinterpret_cast<float4*>(wsp + (int64_t)prow * 128 + c0);
#pragma unroll
        for (int q = 0; q < 8; ++q) __stcg(dst + q, make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]));
      }
      tcgen05_fence_before();
      __syncthreads();
      if (threadIdx.x == 0) st_release_gpu(p.flag_part + blockIdx.x, it);
      LS2_STAMP(4);
      // the 16 partials of unit tile jt: lanes 0..15 of every warp poll one each (every warp reads them)
      {
        const int* f = p.flag_part + jt * LS2_S + (lane & 15);
        long long t0 = 0;
        while (true) {
          const bool ok = ld_acquire_gpu(f) >= it;
          if (__all_sync(0xffffffffu, ok)) break;
          if (t0 == 0) t0 = clock64();
          else if (clock64() - t0 > 2000000000LL) __trap();
        }
      }
      LS2_STAMP(5);
      const float* src = p.ws + ((int64_t)((it & 1) * ctas) + jt * LS2_S) * tile_elems + (int64_t)trow * 128 + 4 * lane;
      float4 part[LS2_S];
#pragma unroll
      for (int z = 0; z < LS2_S; ++z) part[z] = __ldcg(reinterpret_cast<const float4*>(src + (int64_t)z * tile_elems));
#pragma unroll
      for (int z = 0; z < LS2_S; ++z) { dh.x += part[z].x; dh.y += part[z].y; dh.z += part[z].z; dh.w += part[z].w; }
      LS2_STAMP(6);
    }
    if (it < p.T) {
      float4 d0, d1, d2, d3;
#define LS2_CELL(m)                                                         \
      {                                                                     \
        const float tc = lp_tanh(cn.m);                                     \
        const float dc = dh.m * go.m * (1.f - tc * tc) + dcn.m;             \
        d0.m = dc * gg.m * gi.m * (1.f - gi.m);                             \
        d1.m = dc * cp.m * gf.m * (1.f - gf.m);                             \
        d2.m = dh.m * tc * go.m * (1.f - go.m);                             \
        d3.m = dc * gi.m * (1.f - gg.m * gg.m);                             \
        dcn.m = dc * gf.m;                                                  \
      }
      LS2_CELL(x) LS2_CELL(y) LS2_CELL(z) LS2_CELL(w)
#undef LS2_CELL
      if (live) {
        float* dp = p.dA + (int64_t)t * sA + gcell;
        *reinterpret_cast<float4*>(dp) = d0;
        *reinterpret_cast<float4*>(dp + H) = d1;
        *reinterpret_cast<float4*>(dp + 2 * H) = d2;
        *reinterpret_cast<float4*>(dp + 3 * H) = d3;
      }
      __syncthreads();
      if (threadIdx.x == 0) {
        asm volatile("fence.proxy.async.global;" ::: "memory");
        st_release_gpu(p.flag_da + blockIdx.x, it + 1);
      }
      LS2_STAMP(7);
    } else if (live) {
      *reinterpret_cast<float4*>(p.dH + cell) = dh;          // dL/dh0
    }
  }
  tcgen05_fence_before();
  __syncthreads();
  if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(128u) : "memory");
}

}  // namespace tc

// The whole forward recurrence as one launch.  *taken = 0 (nothing launched) when the shape is outside the
// kernel's range (N <= 128, H a multiple of 128 up to 1024, 16-byte aligned buffers), when the device cannot
// hold all H/32 clusters at once (the CTAs wait for each other), or with MC_LSTM_PERSIST=0.
int lstm_persist_fwd(const float* A, const float* Wh, float* hbuf, float* cbuf, float* gates, int64_t T, int64_t N, int64_t H,
                     int* taken) {
  using namespace tc;
  *taken = 0;
  static int enabled = -1;
  if (enabled < 0) { const char* e = getenv("MC_LSTM_PERSIST"); enabled = (e && e[0] == '0') ? 0 : 1; }
  if (!enabled || T <= 0 || N <= 0 || N > 128 || H < 128 || H > 1024 || (H & 127)) return 0;
  if (T * N >= (int64_t)1 << 30) return 0;
  for (const void* q : {(const void*)A, (const void*)Wh, (const void*)hbuf, (const void*)cbuf, (const void*)gates})
    if (!aligned16(q)) return 0;
  const int kb_per = (int)(H >> 7), ctas = (int)(H >> 3);
  const size_t smem = (size_t)LP_RING + (size_t)kb_per * LP_TILE + 128 + 1024;
  static uint32_t configured = 0;
  if (first_use_on_device(&configured))
    MC_CHECK(cudaFuncSetAttribute(lstm_persist_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)(LP_RING + LP_MAX_KB * LP_TILE + 128 + 1024)));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)ctas); cfg.blockDim = dim3(LP_THREADS); cfg.dynamicSmemBytes = smem; cfg.stream = stream();
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = LP_S; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr; cfg.numAttrs = 1;
  // every CTA waits for others: all clusters must be resident at once
  static int fits[16][LP_MAX_KB + 1] = {};          // [device][kb_per]: 0 unknown, 1 yes, -1 no
  int& fit = fits[device() & 15][kb_per];
  if (fit == 0) {
    int nclusters = 0;
    cudaError_t e = cudaOccupancyMaxActiveClusters(&nclusters, lstm_persist_fwd_kernel, &cfg);
    if (getenv("MC_LSTM_DEBUG")) fprintf(stderr, "lstm_persist_fwd: H=%d needs %d clusters of %d, device holds %d (%s)\n", (int)H, ctas / LP_S, LP_S, nclusters, cudaGetErrorString(e));
    if (e != cudaSuccess) { cudaGetLastError(); nclusters = 0; }
    fit = (nclusters >= ctas / LP_S) ? 1 : -1;
  }
  if (fit < 0) return 0;

  ScratchFrame scratch;
  float* wp = scratch.floats((size_t)(4 * H * H));
  int* flags = reinterpret_cast<int*>(scratch.floats((size_t)ctas));
  if (!wp || !flags) return 1;
  mc::launch(lstm_pack_wh_kernel, dim3((unsigned)(H / 32), (unsigned)(4 * H / 32)), 256, 0, Wh, wp, (int)H);
  MC_LAUNCHED();
  MC_CHECK(cudaMemsetAsync(flags, 0, sizeof(int) * (size_t)ctas, stream()));
  CUtensorMap tmH, tmW;
  {
    const uint64_t dims[2] = {(uint64_t)H, (uint64_t)((T + 1) * N)};
    const uint64_t strides[1] = {(uint64_t)H * 4};
    const uint32_t box[2] = {32, 128};
    if (make_tmap(&tmH, hbuf, 2, dims, strides, box, false)) return 1;
  }
  {
    const uint64_t dims[2] = {(uint64_t)H, (uint64_t)(4 * H)};
    const uint64_t strides[1] = {(uint64_t)H * 4};
    const uint32_t box[2] = {32, 128};
    if (make_tmap(&tmW, wp, 2, dims, strides, box, false)) return 1;
  }
  LstmPersistParams p;
  p.ax = A; p.hbuf = hbuf; p.cbuf = cbuf; p.gates = gates; p.flags = flags; p.T = (int)T; p.N = (int)N; p.H = (int)H;
  MC_CHECK(cudaLaunchKernelEx(&cfg, lstm_persist_fwd_kernel, tmH, tmW, p));
  count_launch();
  *taken = 1;
  return 0;
}

// One configuration of the backward kernel: *taken = 1 when it was launched, 0 when the device cannot hold its
// clusters at once (or its shared memory does not fit).
template <int BN, int STAGES>
static int lstm_persist_bwd_launch(const float* gates, const float* cbuf, const float* Wh, float* dH, float* dA, int64_t T,
                                   int64_t N, int64_t H, int* taken) {
  using namespace tc;
  *taken = 0;
  const int kb_per = (int)(H >> 6), clusters = (int)((H + BN - 1) / BN), ctas = clusters * LB_S;
  const size_t smem = (size_t)STAGES * LP_TILE + (size_t)kb_per * (BN * 128) + 128 + 1024;
  if (smem > 227 * 1024 || ctas > 128) return 0;
  static uint32_t configured = 0;
  if (first_use_on_device(&configured))
    MC_CHECK(cudaFuncSetAttribute(lstm_persist_bwd_kernel<BN, STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)ctas); cfg.blockDim = dim3(4 * BN); cfg.dynamicSmemBytes = smem; cfg.stream = stream();
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = LB_S; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr; cfg.numAttrs = 1;
  static int fits[16][LB_MAX_KB + 1] = {};          // [device][kb_per]: 0 unknown, 1 yes, -1 no
  int& fit = fits[device() & 15][kb_per];
  if (fit == 0) {
    int nclusters = 0;
    cudaError_t e = cudaOccupancyMaxActiveClusters(&nclusters, lstm_persist_bwd_kernel<BN, STAGES>, &cfg);
    if (getenv("MC_LSTM_DEBUG"))
      fprintf(stderr, "lstm_persist_bwd<%d>: H=%d needs %d clusters of %d, device holds %d (%s)\n", BN, (int)H, clusters, LB_S,
              nclusters, cudaGetErrorString(e));
    if (e != cudaSuccess) { cudaGetLastError(); nclusters = 0; }
    fit = (nclusters >= clusters) ? 1 : -1;
  }
  if (fit < 0) return 0;

  ScratchFrame scratch;
  int* flags = reinterpret_cast<int*>(scratch.floats((size_t)ctas));
  if (!flags) return 1;
  MC_CHECK(cudaMemsetAsync(flags, 0, sizeof(int) * (size_t)ctas, stream()));
  CUtensorMap tmD, tmW;
  {
    const uint64_t dims[2] = {(uint64_t)(4 * H), (uint64_t)(T * N)};
    const uint64_t strides[1] = {(uint64_t)(4 * H) * 4};
    const uint32_t box[2] = {32, 128};
    if (make_tmap(&tmD, dA, 2, dims, strides, box, false)) return 1;
  }
  {
    const uint64_t dims[2] = {(uint64_t)(4 * H), (uint64_t)H};
    const uint64_t strides[1] = {(uint64_t)(4 * H) * 4};
    const uint32_t box[2] = {32, (uint32_t)BN};
    if (make_tmap(&tmW, Wh, 2, dims, strides, box, false)) return 1;
  }
  LstmPersistBwdParams p;
  p.gates = gates; p.cbuf = cbuf; p.dH = dH; p.dA = dA; p.flags = flags; p.T = (int)T; p.N = (int)N; p.H = (int)H;
  MC_CHECK(cudaLaunchKernelEx(&cfg, lstm_persist_bwd_kernel<BN, STAGES>, tmD, tmW, p));
  count_launch();
  *taken = 1;
  return 0;
}

// The cluster-free backward kernel (lstm_stream_bwd_kernel): *taken = 1 when launched.
static int lstm_stream_bwd_launch(const float* gates, const float* cbuf, const float* Wh, float* dH, float* dA, int64_t T,
                                  int64_t N, int64_t H, int* taken) {
  using namespace tc;
  *taken = 0;
  const int ctas = (int)(H >> 3), slice = (int)(H >> 2);
  for (int s = 0; s < 4; ++s)                        // a K slice's units must come from at most two 128-unit tiles
    if (((s * slice + slice - 1) >> 7) - ((s * slice) >> 7) > 1) return 0;
  const size_t smem = (size_t)LS2_STAGES * LS2_STAGE + 128 + 1024;
  static uint32_t configured = 0;
  if (first_use_on_device(&configured))
    MC_CHECK(cudaFuncSetAttribute(lstm_stream_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  static int per_sm[16] = {};
  int& fit = per_sm[device() & 15];
  if (fit == 0) {
    int nb = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, lstm_stream_bwd_kernel, LP_THREADS, smem) != cudaSuccess) { cudaGetLastError(); nb = 0; }
    fit = nb > 0 ? nb : -1;
  }
  if (fit < 0 || (int64_t)fit * sm_count() < ctas) return 0;     // every CTA waits for others: all must be resident

  ScratchFrame scratch;
  float* ws = scratch.floats((size_t)2 * ctas * 128 * 128);
  int* flags = reinterpret_cast<int*>(scratch.floats((size_t)2 * ctas));
  if (!ws || !flags) return 1;
  MC_CHECK(cudaMemsetAsync(flags, 0, sizeof(int) * 2 * (size_t)ctas, stream()));
  CUtensorMap tmD, tmW;
  {
    const uint64_t dims[2] = {(uint64_t)(4 * H), (uint64_t)(T * N)};
    const uint64_t strides[1] = {(uint64_t)(4 * H) * 4};
    const uint32_t box[2] = {32, 128};
    if (make_tmap(&tmD, dA, 2, dims, strides, box, false)) return 1;
  }
  {
    const uint64_t dims[2] = {(uint64_t)(4 * H), (uint64_t)H};
    const uint64_t strides[1] = {(uint64_t)(4 * H) * 4};
    const uint32_t box[2] = {32, 128};
    if (make_tmap(&tmW, Wh, 2, dims, strides, box, false)) return 1;
  }
  LstmStreamBwdParams p;
  p.gates = gates; p.cbuf = cbuf; p.dH = dH; p.dA = dA; p.ws = ws; p.flag_da = flags; p.flag_part = flags + ctas;
  p.T = (int)T; p.N = (int)N; p.H = (int)H;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)ctas); cfg.blockDim = dim3(LP_THREADS); cfg.dynamicSmemBytes = smem; cfg.stream = stream();
  MC_CHECK(cudaLaunchKernelEx(&cfg, lstm_stream_bwd_kernel, tmD, tmW, p));
  count_launch();
  *taken = 1;
  return 0;
}

// The whole backward recurrence as one launch (same range and conditions as lstm_persist_fwd): 64-unit clusters
// when the device holds H/64 of them (H <= 896 on this pool's B200), else nothing is launched.
int lstm_persist_bwd(const float* gates, const float* cbuf, const float* Wh, float* dH, float* dA, int64_t T, int64_t N, int64_t H,
                     int* taken) {
  *taken = 0;
  static int enabled = -1, force80 = 0;
  if (enabled < 0) {
    const char* e = getenv("MC_LSTM_PERSIST");
    const char* b = getenv("MC_LSTM_PERSIST_BWD");
    enabled = ((e && e[0] == '0') || (b && b[0] == '0')) ? 0 : 1;
    force80 = (b && b[0] == '8') ? 1 : ((b && b[0] == 's') ? 2 : 0);   // =80: the 80-unit clusters; =stream: the cluster-free kernel (tests)
  }
  if (!enabled || T <= 0 || N <= 0 || N > 128 || H < 128 || H > 1024 || (H & 127)) return 0;
  if (T * N >= (int64_t)1 << 30) return 0;
  for (const void* q : {(const void*)gates, (const void*)cbuf, (const void*)Wh, (const void*)dH, (const void*)dA})
    if (!aligned16(q)) return 0;
  // MEASURED (profiles/r2r, C5: T = 100, N = 128, H = 1024): the 80-unit configuration is correct but takes 13.4 us per
  // step against 10.0 us for the launch-per-step path: 160 KB of resident Wh leave room for four 16 KB stages only, and
  // 64 KB in flight against ~1.2 us of TMA latency is 53 GB/s per SM -- 4.9 us for the 256 KB of dA a CTA needs per
  // step.  It therefore runs only on request (MC_LSTM_PERSIST_BWD=80); the 64-unit configuration (six stages,
  // 8.7 us per step at H = 896) is the default wherever its clusters fit.
  if (force80 == 1) return lstm_persist_bwd_launch<80, 4>(gates, cbuf, Wh, dH, dA, T, N, H, taken);
  if (force80 == 2) return lstm_stream_bwd_launch(gates, cbuf, Wh, dH, dA, T, N, H, taken);
  // (lstm_stream_bwd_kernel, the cluster-free variant for H = 1024, is correct but MEASURED at 16.9 us per step against
  //  10.0 us for the launch-per-step path (profiles/r2w): its two flag hops per step -- partial tiles through L2, then
  //  dA -- cost 3.1 + 4.2 us with the skew of 128 CTAs, more than the two kernel boundaries they replace.  On request
  //  only: MC_LSTM_PERSIST_BWD=stream.)
  return lstm_persist_bwd_launch<64, 6>(gates, cbuf, Wh, dH, dA, T, N, H, taken);
}

}  // namespace mc
