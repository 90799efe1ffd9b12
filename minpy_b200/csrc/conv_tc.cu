// Implicit-GEMM convolution on the sm_100a tensor cores: the `Convolution` symbol the reference
// reaches through core.Function (minpy/core.py:148-190; examples/nn/cnn.py:32-40), which MXNet
// ran through cuDNN.  No im2col buffer ever exists in HBM: TMA gathers the shifted input windows
// straight into the UMMA operand layout, tcgen05.mma (kind::tf32) accumulates in TMEM and the
// epilogue writes the activation with bias / relu fused.
//
// Layout.  Activations are channels-last (N, H, W, Cp), Cp = channels rounded up to 4 (16 B):
// TMA requires the innermost start coordinate of a box to be 16-byte aligned (measured on this
// part: a +-1 element shift along W of an NCHW tensor raises "illegal instruction"), so the
// filter-tap shifts must land on OUTER dimensions.  The symbol executor keeps whole conv trunks
// in this layout and converts NCHW <-> NHWC only at their boundaries.
//
//   forward / data-gradient (stride 1):  D[pixel, n] = sum_{tap, k} A[pixel, (tap,k)] * B[n, (tap,k)]
//       A: for one tap and a block of 32 channels, 32 consecutive output pixels (a "chunk") are
//          ONE 4-D TMA box {32 ch, bw, bh, 1} at (c0, ow0+kw-pad, oh0+kh-pad, n) -> 32 rows of
//          128 B = a K-major SWIZZLE_128B operand block; halo / channel overrun is zero-filled by
//          TMA.  A CTA tile is 4 chunks = 128 pixels = the 128 TMEM lanes.
//       B: weights repacked per call to [tap][n][k] (k contiguous): K-major too.
//       "merged" mode for few channels (KW * Cp <= 32, e.g. RGB): the input is physically padded
//          along W and one box row is the contiguous run of KW pixels x Cp channels, so a tap is a
//          filter ROW and the first layer costs KH k-blocks instead of KH*KW.
//       dgrad is the same kernel on dy with the filter flipped and k/n swapped (pad' = K-1-pad).
//   weight-gradient:  D[f, n] = sum_{pixel} dy[pixel, f] * x[pixel + tap, n]   per tap
//       both operands MN-major (f / channels contiguous), SWIZZLE_128B_ATOM_32B blocks of
//       [32 pixels][32 f|c]; the pixel range is split over CTAs and reduced deterministically.
//
// Output pixels are enumerated over a "virtual" image whose row width is a divisor or multiple of
// 32 so that a chunk is a rectangle; virtual pixels outside the real output are computed on zeros
// and masked in the epilogue.
#include "tc_common.cuh"

namespace mc {

int reduce_sum_internal(const float* x, float* out, int64_t outer, int64_t red, int64_t inner);   // reduce.cu

namespace tc {

constexpr int CONV_THREADS = 128;
constexpr int BKC = 32;        // channels (fp32) per k-block = one 128-byte swizzle row

struct ConvGeom {
  int N, OH, OW;         // real output extent (pixels the epilogue keeps)
  int OHv, OWv;          // virtual extent enumerated by chunks
  int bw, bh;            // chunk rectangle: bw * bh == 32
  int cpr, cpi;          // chunks per virtual row (OWv >= 32) / per image
  int total_chunks;
  int dh0, dw0;          // input coordinate = output coordinate + tap + d?0
  int TH, TW;            // taps along h / w enumerated by the kernel
  int one_box;           // the 4 chunks of a tile form ONE TMA box {32, bw, box_h, box_n}
  int box_h, box_n;
  int cpi_shift;         // log2(cpi) when cpi is a power of two, else -1
  int owv_shift;         // log2(OWv) when OWv < 32 (8 or 16)
};

__device__ __forceinline__ void chunk_origin(const ConvGeom& g, int chunk, int& n, int& oh0, int& ow0) {
  n = g.cpi_shift >= 0 ? (chunk >> g.cpi_shift) : chunk / g.cpi;
  const int r = chunk - n * g.cpi;
  if (g.OWv >= 32) { oh0 = r / g.cpr; ow0 = (r - oh0 * g.cpr) * 32; }
  else { oh0 = r * g.bh; ow0 = 0; }
  if (chunk >= g.total_chunks) n = g.N;   // fully out of bounds: TMA fills zeros, epilogue masks
}

struct ConvFwdParams {
  ConvGeom g;
  float* y;              // (N, OH, OW, ldy)
  const float* bias;     // (Nout,) or nullptr
  int Nout, ldy;         // real output channels, channel pitch of y
  int BN, ncb, stages, relu;
  int nacc;              // accumulator tiles in the TMEM ring (nacc * BN <= 512 columns)
  int wres;              // all weight blocks stay resident in shared memory (one n-tile, <= 96 KB)
  long long* dbg;        // MC_CONV_DBG: clock64 stamps of CTA 0 [role][tile][slot] (nullptr = off)
  int halo;              // stage = one channel block of a whole tile WITH its filter halo (see below)
  uint32_t copy_bytes;   // halo mode: bytes of one column-shifted copy, (box_h + TH - 1) * bw * 128
  // pool != 0: the epilogue applies bias (+ relu) and a 2x2 / stride-2 max pooling and stores ONLY the pooled
  // tensor y (N, OH/2, OW/2, ldy) and the window index (0..3 = kh*2 + kw of the first maximum, the format of
  // maxpool_nhwc_fwd) -- the full-resolution activation never goes to memory.
  int pool;
  uint8_t* pidx;         // (N, OH/2, OW/2, ldy) bytes, or nullptr (inference)
  uint32_t pool_smem;    // byte offset (from the 1024-aligned base) of the cross-warp exchange area (OWv == 32 only)
};

// Persistent: grid = min(#tiles, #SMs); a CTA walks tiles t = blockIdx.x, += gridDim.x, where a
// tile is (4 chunks of pixels) x (BN output channels).  Six warps: 0-3 epilogue (TMEM lanes
// 32w..32w+31), 4 TMA producer, 5 MMA issuer.  The shared-memory ring runs continuously across
// tiles and the accumulator is double-buffered in TMEM (2 x BN columns), so the epilogue of tile
// i overlaps the loads and MMAs of tile i+1 and the per-CTA set-up cost is paid once per SM.
constexpr int IGEMM_THREADS = 320;   // warps 0-7: two epilogue groups (alternate tiles), 8: TMA producer, 9: MMA issuer

__global__ void __launch_bounds__(IGEMM_THREADS, 1) conv_igemm_kernel(const __grid_constant__ CUtensorMap tmX,
                                                                       const __grid_constant__ CUtensorMap tmW,
                                                                       const ConvFwdParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  constexpr uint32_t chunk_bytes = 32u * 128u;                    // [32 pixels][32 ch]
  constexpr uint32_t a_bytes = 4u * chunk_bytes;
  const uint32_t b_bytes = (uint32_t)p.BN * 128u;                 // [BN][32 ch]
  const uint32_t w_bytes = p.wres ? (uint32_t)(p.g.TH * p.g.TW * p.ncb) * b_bytes : 0u;   // resident weights first
  const uint32_t stage_bytes = p.halo ? (uint32_t)p.g.TW * p.copy_bytes : a_bytes + (p.wres ? 0u : b_bytes);
  const int STAGES = p.stages;
  const int NACC = p.nacc;
  const uint32_t ring_base = smem_base + w_bytes;
  const uint32_t bar_base = ring_base + STAGES * stage_bytes;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (STAGES + s); };
  auto tmem_full_bar = [&](int a) { return bar_base + 8u * (2 * STAGES + a); };
  auto tmem_empty_bar = [&](int a) { return bar_base + 8u * (2 * STAGES + NACC + a); };
  const uint32_t w_bar = bar_base + 8u * (2 * STAGES + 2 * NACC);
  const uint32_t tmem_slot = w_bar + 8u;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int taps = p.g.TH * p.g.TW;
  const int num_it = p.halo ? p.ncb : taps * p.ncb;       // pipeline stages per tile
  const int ptiles = (p.g.total_chunks + 3) / 4;
  const int ntiles = (p.Nout + p.BN - 1) / p.BN;
  const int total_tiles = ptiles * ntiles;
  const uint32_t tcols = (uint32_t)(p.nacc * p.BN);               // BN, nacc powers of two; nacc * BN in [64, 512]

  if (warp == 8 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
    for (int a = 0; a < NACC; ++a) { mbar_init(tmem_full_bar(a), 1); mbar_init(tmem_empty_bar(a), 4); }
    mbar_init(w_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  if (warp == 9) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(tcols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  mc::pdl_sync();      // prologue done (on-chip state only); global memory from here on
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot) : "memory");

  if (warp == 8) {
    if (elect_one()) {
      // ------------------------------ TMA producer ------------------------------------
      // One thread: every instruction here is serial latency (an EMPTY mainloop still cost 0.2 us
      // per k-block, profiles/r1p), so everything loop-invariant lives in registers, the stage
      // index / phase are carried incrementally and nothing is divided inside the tap loop.
      int s = 0;
      uint32_t ph = 0;
      const int TH = p.g.TH, TW = p.g.TW, ncb = p.ncb, dh0 = p.g.dh0, dw0 = p.g.dw0, BN = p.BN;
      const bool one_box = p.g.one_box != 0, wres = p.wres != 0;
      int pt = blockIdx.x % ptiles, nt = blockIdx.x / ptiles;
      const int dpt = gridDim.x % ptiles, dnt = gridDim.x / ptiles;
      if (wres && (int)blockIdx.x < total_tiles) {
        // every CTA needs every weight block for every tile: fetch them once per CTA instead of
        // once per tile (148 SMs re-reading the same few L2 lines per k-block was the bottleneck
        // of the first layer: 50 of 120 us, profiles/r1p)
        mbar_expect_tx(w_bar, w_bytes);
        for (int t = 0; t < taps; ++t)
          for (int cb = 0; cb < ncb; ++cb)
            tma_load_3d(smem_base + (uint32_t)(t * ncb + cb) * b_bytes, &tmW, w_bar, cb * BKC, 0, t);
      }
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        const int chunk0 = pt * 4;
        const int n0 = nt * BN;
        pt += dpt; nt += dnt;
        if (pt >= ptiles) { pt -= ptiles; ++nt; }
        int cn[4], coh[4], cow[4];
        if (one_box) {
          chunk_origin(p.g, chunk0, cn[0], coh[0], cow[0]);
        } else {
#pragma unroll
          for (int j = 0; j < 4; ++j) chunk_origin(p.g, chunk0 + j, cn[j], coh[j], cow[j]);
        }
        const int dti = (tile - (int)blockIdx.x) / (int)gridDim.x;
        if (p.dbg && blockIdx.x == 0 && dti < 32) p.dbg[(0 * 32 + dti) * 4 + 0] = clock64();
        if (p.halo) {
          // Halo mode.  The TMA unit, not the tensor pipe, bounds this kernel (~5 cycles per 128-byte
          // pixel row, MC_CONV_DBG timeline), and loading one box per filter tap fetches every input
          // row TH*TW times.  Here a stage holds, per column shift tw, ONE box of the tile's rows plus
          // the TH-1 halo rows; the row shifts th are then just +th*bw*128 bytes on the shared-memory
          // descriptor (a multiple of the 1024-byte swizzle atom since bw >= 8): TW boxes of
          // (box_h + TH - 1) * bw rows instead of TH*TW boxes of box_h * bw rows.
          for (int cb = 0; cb < ncb; ++cb) {
            mbar_wait(empty_bar(s), ph ^ 1u);
            if (p.dbg && blockIdx.x == 0 && dti < 32 && cb == 0) p.dbg[(0 * 32 + dti) * 4 + 1] = clock64();
            const uint32_t fb = full_bar(s);
            mbar_expect_tx(fb, stage_bytes);
            uint32_t sa = ring_base + s * stage_bytes;
            for (int tw = 0; tw < TW; ++tw, sa += p.copy_bytes)
              tma_load_4d(sa, &tmX, fb, cb * BKC, cow[0] + tw + dw0, coh[0] + dh0, cn[0]);
            if (++s == STAGES) { s = 0; ph ^= 1u; }
          }
          if (p.dbg && blockIdx.x == 0 && dti < 32) p.dbg[(0 * 32 + dti) * 4 + 2] = clock64();
          continue;
        }
        int tap = 0;
        for (int th = 0; th < TH; ++th) {
          for (int tw = 0; tw < TW; ++tw, ++tap) {
            for (int cb = 0; cb < ncb; ++cb) {
              mbar_wait(empty_bar(s), ph ^ 1u);
              if (p.dbg && blockIdx.x == 0 && dti < 32 && tap == 0 && cb == 0) p.dbg[(0 * 32 + dti) * 4 + 1] = clock64();
              const uint32_t fb = full_bar(s);
              mbar_expect_tx(fb, stage_bytes);
              const uint32_t sa = ring_base + s * stage_bytes;
              if (one_box) {
                tma_load_4d(sa, &tmX, fb, cb * BKC, cow[0] + tw + dw0, coh[0] + th + dh0, cn[0]);
              } else {
#pragma unroll
                for (int j = 0; j < 4; ++j)
                  tma_load_4d(sa + j * chunk_bytes, &tmX, fb, cb * BKC, cow[j] + tw + dw0, coh[j] + th + dh0, cn[j]);
              }
              if (!wres) tma_load_3d(sa + a_bytes, &tmW, fb, cb * BKC, n0, tap);
              if (++s == STAGES) { s = 0; ph ^= 1u; }
            }
          }
        }
        if (p.dbg && blockIdx.x == 0 && dti < 32) p.dbg[(0 * 32 + dti) * 4 + 2] = clock64();
      }
    }
  } else if (warp == 9) {
    if (elect_one()) {
      // ------------------------------ MMA issuer --------------------------------------
      // c_format F32, a/b TF32, both operands K-major, N>>3 at [17,23), M>>4 at [24,29)
      const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(p.BN >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
      int s = 0, acc = 0;
      uint32_t ph = 0, aph = 0;
      const bool wres = p.wres != 0;
      const int BN = p.BN;
      // descriptor of address A = constant bits | (A >> 4): stepping 32 bytes along K adds 2
      const uint64_t desc0 = make_desc(0, 16, 1024, 2);
      if (wres && (int)blockIdx.x < total_tiles) mbar_wait(w_bar, 0);
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        const int dti = (tile - (int)blockIdx.x) / (int)gridDim.x;
        if (p.dbg && blockIdx.x == 0 && dti < 32) p.dbg[(1 * 32 + dti) * 4 + 0] = clock64();
        mbar_wait(tmem_empty_bar(acc), aph ^ 1u);                          // epilogue drained this accumulator
        tcgen05_fence_after();
        if (p.dbg && blockIdx.x == 0 && dti < 32) p.dbg[(1 * 32 + dti) * 4 + 1] = clock64();
        const uint32_t tacc = tmem_base + (uint32_t)(acc * BN);
        if (p.halo) {
          const int TH = p.g.TH, TW = p.g.TW, ncb = p.ncb;
          const uint32_t row_bytes = (uint32_t)p.g.bw * 128u;
          for (int cb = 0; cb < ncb; ++cb) {
            mbar_wait(full_bar(s), ph);
            tcgen05_fence_after();
            if (p.dbg && blockIdx.x == 0 && dti < 32 && cb == 0) p.dbg[(1 * 32 + dti) * 4 + 2] = clock64();
            const uint32_t sa = ring_base + s * stage_bytes;
            uint32_t wb = smem_base + (uint32_t)cb * b_bytes;              // weight block (tap 0, cb); taps are ncb blocks apart
            if (TH == 3 && TW == 3) {
              // 3x3: the 36 MMAs of a stage as straight-line code -- every descriptor is a constant
              // offset from two bases, so the issuing thread spends ~30 instead of ~100 cycles per MMA
              const uint64_t a0 = desc0 + (uint64_t)((sa & 0x3FFFFu) >> 4);
              const uint64_t b0 = desc0 + (uint64_t)((wb & 0x3FFFFu) >> 4);
              const uint32_t rstep = row_bytes >> 4, cstep = p.copy_bytes >> 4, wstep = ((uint32_t)ncb * b_bytes) >> 4;
#pragma unroll
              for (int th = 0; th < 3; ++th)
#pragma unroll
                for (int tw = 0; tw < 3; ++tw)
#pragma unroll
                  for (int k = 0; k < BKC / UMMA_K; ++k)
                    umma_tf32(tacc, a0 + (uint64_t)(th * rstep + tw * cstep + 2u * k), b0 + (uint64_t)((th * 3 + tw) * wstep + 2u * k), idesc,
                              (cb > 0 || th > 0 || tw > 0 || k > 0) ? 1u : 0u);
            } else
            for (int th = 0; th < TH; ++th) {
              uint32_t at = sa + (uint32_t)th * row_bytes;
              for (int tw = 0; tw < TW; ++tw, at += p.copy_bytes, wb += (uint32_t)ncb * b_bytes) {
                const uint64_t adesc = desc0 + (uint64_t)((at & 0x3FFFFu) >> 4);
                const uint64_t bdesc = desc0 + (uint64_t)((wb & 0x3FFFFu) >> 4);
#pragma unroll
                for (int k = 0; k < BKC / UMMA_K; ++k)
                  umma_tf32(tacc, adesc + 2u * k, bdesc + 2u * k, idesc, (cb > 0 || th > 0 || tw > 0 || k > 0) ? 1u : 0u);
              }
            }
            umma_commit(empty_bar(s));
            if (++s == STAGES) { s = 0; ph ^= 1u; }
          }
          umma_commit(tmem_full_bar(acc));
          if (p.dbg && blockIdx.x == 0 && dti < 32) p.dbg[(1 * 32 + dti) * 4 + 3] = clock64();
          if (++acc == NACC) { acc = 0; aph ^= 1u; }
          continue;
        }
        uint32_t wb = smem_base;                                           // resident weight block of iteration i
        for (int i = 0; i < num_it; ++i, wb += b_bytes) {
          mbar_wait(full_bar(s), ph);
          tcgen05_fence_after();
          if (p.dbg && blockIdx.x == 0 && dti < 32 && i == 0) p.dbg[(1 * 32 + dti) * 4 + 2] = clock64();
          const uint32_t sa = ring_base + s * stage_bytes;
          const uint32_t sb = wres ? wb : sa + a_bytes;
          const uint64_t adesc = desc0 + (uint64_t)((sa & 0x3FFFFu) >> 4);
          const uint64_t bdesc = desc0 + (uint64_t)((sb & 0x3FFFFu) >> 4);
#pragma unroll
          for (int k = 0; k < BKC / UMMA_K; ++k)
            umma_tf32(tacc, adesc + 2u * k, bdesc + 2u * k, idesc, (i > 0 || k > 0) ? 1u : 0u);
          umma_commit(empty_bar(s));
          if (++s == STAGES) { s = 0; ph ^= 1u; }
        }
        umma_commit(tmem_full_bar(acc));
        if (p.dbg && blockIdx.x == 0 && dti < 32) p.dbg[(1 * 32 + dti) * 4 + 3] = clock64();
        if (++acc == NACC) { acc = 0; aph ^= 1u; }
      }
    }
  } else {
    // ------------------------- epilogue (warps 0-3: even tiles, 4-7: odd tiles) -----------------
    // One tile's epilogue (barrier wait, tcgen05.ld, 8 scattered 16-byte stores per 32 columns,
    // release) is ~1 us of serial latency per warp; with a few k-blocks per tile that, not the
    // loads or the MMAs, set the tile rate.  Two groups of four warps drain alternate accumulators.
    const bool vec_ok = ((p.ldy & 3) == 0) && ((reinterpret_cast<uintptr_t>(p.y) & 15u) == 0);
    const int grp = warp >> 2, quarter = warp & 3;
    int acc = grp % NACC;
    uint32_t aph = 0, xpar = 0;                       // xpar: which half of the pooled epilogue's exchange buffer
    const int first = blockIdx.x + grp * gridDim.x, stride = 2 * gridDim.x;
    int pt = first % ptiles, nt = first / ptiles;
    const int dpt = stride % ptiles, dnt = stride / ptiles;
    for (int tile = first; tile < total_tiles; tile += stride) {
      const int chunk0 = pt * 4;
      const int n0 = nt * p.BN;
      pt += dpt; nt += dnt;
      if (pt >= ptiles) { pt -= ptiles; ++nt; }
      int n, oh0, ow0;
      chunk_origin(p.g, chunk0 + quarter, n, oh0, ow0);
      int oh, ow;
      if (p.g.OWv >= 32) { oh = oh0; ow = ow0 + lane; }
      else { oh = oh0 + (lane >> p.g.owv_shift); ow = lane & (p.g.OWv - 1); }
      const bool live = (n < p.g.N) && (oh < p.g.OH) && (ow < p.g.OW);
      float* __restrict__ yrow = p.y + (((int64_t)n * p.g.OH + oh) * p.g.OW + ow) * p.ldy;
      const int dti = (tile - (int)blockIdx.x) / (int)gridDim.x;
      const bool dbgw = p.dbg && blockIdx.x == 0 && dti < 32 && quarter == 0 && lane == 0;
      if (dbgw) p.dbg[(2 * 32 + dti) * 4 + 0] = clock64();
      mbar_wait_warp(tmem_full_bar(acc), aph);
      tcgen05_fence_after();
      if (dbgw) p.dbg[(2 * 32 + dti) * 4 + 1] = clock64();
      const uint32_t tacc = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(acc * p.BN);
#pragma unroll 1
      for (int c0 = 0; c0 < p.BN; c0 += 32) {
        float v[32];
        tmem_ld32(tacc + (uint32_t)c0, v);
        if (p.pool) {
          // ---- bias, relu, 2x2 max pooling.  Candidates in window order (0,0) (0,1) (1,0) (1,1), the first maximum
          // wins, a NaN beats numbers (maxpool_nhwc_fwd's rule).  The four lanes of a window split the 32 channels:
          // a lane trades the half it does not keep with its horizontal neighbour, then the quarter it does not
          // keep with its vertical neighbour, and finishes 8 channels -- 24 shuffles, all lanes storing.  (The first
          // version shuffled every channel three times and left the result in one lane of four: ~1100 instructions
          // and ~5000 cycles per 32 channels, which made all three C4 layers epilogue-bound: profiles/r2l.)
          // relu commutes with the maximum; where it flattens the whole window to zero the first candidate wins.
          const int f0 = n0 + c0;
          if (p.bias) {
            if ((reinterpret_cast<uintptr_t>(p.bias) & 15u) == 0) {
#pragma unroll
              for (int j = 0; j < 32; j += 4) {
                const float4 b = __ldg(reinterpret_cast<const float4*>(p.bias + f0 + j));
                v[j] += b.x; v[j + 1] += b.y; v[j + 2] += b.z; v[j + 3] += b.w;
              }
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j) v[j] += __ldg(p.bias + f0 + j);
            }
          }
          const bool hp = (lane & 1) != 0;      // odd column of the window
          float m1[16];
          uint32_t kb = 0;                      // bit s: the right-hand candidate won channel s of the kept half
#pragma unroll
          for (int s2 = 0; s2 < 16; ++s2) {
            const float mine = hp ? v[16 + s2] : v[s2];
            const float send = hp ? v[s2] : v[16 + s2];
            const float recv = __shfl_xor_sync(0xffffffffu, send, 1);
            const float left = hp ? recv : mine, right = hp ? mine : recv;
            const bool tr = !(left >= right) && (left == left);
            m1[s2] = tr ? right : left;
            kb |= (tr ? 1u : 0u) << s2;
          }
          const bool within_warp = p.g.OWv < 32;
          const bool vp = within_warp ? (((lane >> p.g.owv_shift) & 1) != 0) : ((quarter & 1) != 0);   // bottom row
          float own[8], rcv[8];
          uint32_t okb;
          if (within_warp) {
            const int rowbit = 1 << p.g.owv_shift;
#pragma unroll
            for (int s2 = 0; s2 < 8; ++s2) {
              own[s2] = vp ? m1[8 + s2] : m1[s2];
              rcv[s2] = __shfl_xor_sync(0xffffffffu, vp ? m1[s2] : m1[8 + s2], rowbit);
            }
            okb = __shfl_xor_sync(0xffffffffu, kb, rowbit);
          } else {
            // the rows of a window live in two warps (quarter q and q ^ 1): 48 bytes per lane through shared memory,
            // double-buffered by pass so that one named barrier per pass is enough
            const int pair = grp * 2 + (quarter >> 1);
            const uint32_t xbase = smem_base + p.pool_smem + (uint32_t)((pair * 2 + (int)xpar) * 3072);
            const uint32_t mine_at = xbase + (uint32_t)(((quarter & 1) * 32 + lane) * 48);
            const uint32_t peer_at = xbase + (uint32_t)((((quarter & 1) ^ 1) * 32 + lane) * 48);
            xpar ^= 1u;
#pragma unroll
            for (int s2 = 0; s2 < 8; ++s2) own[s2] = vp ? m1[8 + s2] : m1[s2];
            asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(mine_at), "f"(vp ? m1[0] : m1[8]), "f"(vp ? m1[1] : m1[9]),
                         "f"(vp ? m1[2] : m1[10]), "f"(vp ? m1[3] : m1[11]) : "memory");
            asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(mine_at + 16u), "f"(vp ? m1[4] : m1[12]), "f"(vp ? m1[5] : m1[13]),
                         "f"(vp ? m1[6] : m1[14]), "f"(vp ? m1[7] : m1[15]) : "memory");
            asm volatile("st.shared.b32 [%0], %1;" ::"r"(mine_at + 32u), "r"(kb) : "memory");
            asm volatile("bar.sync %0, 64;" ::"r"(pair + 1) : "memory");
            asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(rcv[0]), "=f"(rcv[1]), "=f"(rcv[2]), "=f"(rcv[3]) : "r"(peer_at) : "memory");
            asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(rcv[4]), "=f"(rcv[5]), "=f"(rcv[6]), "=f"(rcv[7]) : "r"(peer_at + 16u) : "memory");
            asm volatile("ld.shared.b32 %0, [%1];" : "=r"(okb) : "r"(peer_at + 32u) : "memory");
          }
          const uint32_t mybits = (vp ? (kb >> 8) : kb) & 0xffu, otherbits = (vp ? (okb >> 8) : okb) & 0xffu;
          const uint32_t topbits = vp ? otherbits : mybits, botbits = vp ? mybits : otherbits;
          float outv[8];
          uint32_t kk0 = 0, kk1 = 0;
#pragma unroll
          for (int s2 = 0; s2 < 8; ++s2) {
            const float top = vp ? rcv[s2] : own[s2], bot = vp ? own[s2] : rcv[s2];
            const bool tb = !(top >= bot) && (top == top);
            float val = tb ? bot : top;
            uint32_t idx = tb ? (2u + ((botbits >> s2) & 1u)) : ((topbits >> s2) & 1u);
            if (p.relu && !(val > 0.f) && val == val) { val = 0.f; idx = 0u; }
            outv[s2] = val;
            if (s2 < 4) kk0 |= idx << (8 * s2);
            else kk1 |= idx << (8 * (s2 - 4));
          }
          if (live) {
            const int POH = p.g.OH >> 1, POW = p.g.OW >> 1;
            const int64_t at = (((int64_t)n * POH + (oh >> 1)) * POW + (ow >> 1)) * p.ldy + f0 + (hp ? 16 : 0) + (vp ? 8 : 0);
            float* __restrict__ dst = p.y + at;
            *reinterpret_cast<float4*>(dst) = make_float4(outv[0], outv[1], outv[2], outv[3]);
            *reinterpret_cast<float4*>(dst + 4) = make_float4(outv[4], outv[5], outv[6], outv[7]);
            if (p.pidx) *reinterpret_cast<uint2*>(p.pidx + at) = make_uint2(kk0, kk1);
          }
          continue;
        }
        if (live) {
          const int f0 = n0 + c0;
          if (vec_ok && f0 + 32 <= p.Nout) {
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
              float4 o = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
              if (p.bias) {
                const float4 b = __ldg(reinterpret_cast<const float4*>(p.bias + f0 + j));
                o.x += b.x; o.y += b.y; o.z += b.z; o.w += b.w;
              }
              if (p.relu) {
                o.x = (o.x != o.x) ? o.x : fmaxf(o.x, 0.f); o.y = (o.y != o.y) ? o.y : fmaxf(o.y, 0.f);
                o.z = (o.z != o.z) ? o.z : fmaxf(o.z, 0.f); o.w = (o.w != o.w) ? o.w : fmaxf(o.w, 0.f);
              }
              *reinterpret_cast<float4*>(yrow + f0 + j) = o;
            }
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              const int f = f0 + j;
              if (f < p.ldy) {                       // pitch padding is written as zeros
                float o = 0.f;
                if (f < p.Nout) {
                  o = v[j];
                  if (p.bias) o += __ldg(p.bias + f);
                  if (p.relu) o = (o != o) ? o : fmaxf(o, 0.f);
                }
                yrow[f] = o;
              }
            }
          }
        }
      }
      tcgen05_fence_before();
      __syncwarp();
      if (dbgw) p.dbg[(2 * 32 + dti) * 4 + 2] = clock64();
      if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(tmem_empty_bar(acc)) : "memory");
      acc += 2;                                        // NACC is an even power of two
      if (acc >= NACC) { acc -= NACC; aph ^= 1u; }
    }
  }
  tcgen05_fence_before();
  __syncthreads();
  if (warp == 9) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(tcols) : "memory");
  }
}

// wp[tap][n][k] (k pitch Kp, zero padded) from w (F, C, KH, KW).
//   mode 0 (forward)       : tap = kh*KW + kw, n = f, k = c
//   mode 1 (forward merged): tap = kh,         n = f, k = kw*Cp + c
//   mode 2 (dgrad)         : tap = (KH-1-kh)*KW + (KW-1-kw), n = c, k = f
__global__ void __launch_bounds__(256) conv_pack_weights_kernel(const float* __restrict__ w, float* __restrict__ wp, int F,
                                                                int C, int KH, int KW, int Cp, int taps, int Nn, int Kp,
                                                                int mode) {
  mc::pdl_sync();
  const int64_t total = (int64_t)taps * Nn * Kp;
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < total; i += (int64_t)gridDim.x * 256) {
    const int k = (int)(i % Kp);
    const int n = (int)((i / Kp) % Nn);
    const int tap = (int)(i / ((int64_t)Kp * Nn));
    float v = 0.f;
    if (mode == 0) {
      if (k < C) v = w[((int64_t)n * C + k) * KH * KW + tap];
    } else if (mode == 1) {
      const int kw = k / Cp, c = k - kw * Cp;
      if (kw < KW && c < C) v = w[(((int64_t)n * C + c) * KH + tap) * KW + kw];
    } else {
      const int kh = KH - 1 - tap / KW, kw = KW - 1 - tap % KW;
      if (k < F) v = w[(((int64_t)k * C + n) * KH + kh) * KW + kw];
    }
    wp[i] = v;
  }
}

// ------------------------------------ weight gradient --------------------------------------
struct ConvWgradParams {
  ConvGeom g;            // geometry of dy (the output pixels)
  float* ws;             // [splits][taps][F][Nn]
  float* ws_bias;        // [splits][F] or nullptr: column sums of dy (the bias gradient), fused as one more "tap"
  int F, Nn;             // filters; columns per tap (channel pitch, or KW*Cp merged)
  int BNc, stages;       // BNc: Nn rounded up to 32
  int tpc;               // taps per CTA: their accumulators sit side by side in TMEM
  int chunks_per_split;
};

// grid.x = groups of `tpc` taps, grid.y = splits of the pixel range, grid.z = tiles of 128 filters.
// Per 32-pixel chunk a CTA loads dy ONCE (only the 32-filter blocks that exist) and one shifted
// window of x per tap; tap t accumulates into TMEM columns [t*BNc, (t+1)*BNc).  The CTA of tap
// group 0 also multiplies dy by a block of ones: that is the bias gradient, for free.
// A warp's 32 (rows = filters, one per lane) x 32 (columns) block of partial sums to  dst_row0 + r * ld + c :
// turned through shared memory (4 KB at `stage`, 16-byte chunks XOR-swizzled by row) so that eight lanes write one
// 128-byte row segment.  Storing straight from the row-per-lane TMEM layout was 32 scalar stores per lane and tap,
// every warp store touching 32 lines: 13 of the 31 us of the weight-gradient kernel (tools/probe.py wgh, r2l).
__device__ __forceinline__ void store_block_staged(uint32_t stage, const float (&v)[32], float* __restrict__ dst_row0, int64_t ld,
                                                    int rows_left, int cols_left, bool vec) {
  const int lane = threadIdx.x & 31;
#pragma unroll
  for (int j = 0; j < 8; ++j)
    asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(stage + (uint32_t)lane * 128u + (uint32_t)((j ^ (lane & 7)) * 16)),
                 "f"(v[4 * j]), "f"(v[4 * j + 1]), "f"(v[4 * j + 2]), "f"(v[4 * j + 3]) : "memory");
  __syncwarp();
  const int j4 = (lane & 7) * 4, rsub = lane >> 3;
#pragma unroll
  for (int it = 0; it < 8; ++it) {
    const int rr = it * 4 + rsub;
    float4 o;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(o.x), "=f"(o.y), "=f"(o.z), "=f"(o.w)
                 : "r"(stage + (uint32_t)rr * 128u + (uint32_t)(((lane & 7) ^ (rr & 7)) * 16)) : "memory");
    if (rr >= rows_left || j4 >= cols_left) continue;
    float* d = dst_row0 + (int64_t)rr * ld + j4;
    if (vec && j4 + 4 <= cols_left) *reinterpret_cast<float4*>(d) = o;
    else {
      d[0] = o.x;
      if (j4 + 1 < cols_left) d[1] = o.y;
      if (j4 + 2 < cols_left) d[2] = o.z;
      if (j4 + 3 < cols_left) d[3] = o.w;
    }
  }
  __syncwarp();
}

__global__ void __launch_bounds__(CONV_THREADS, 1) conv_wgrad_kernel(const __grid_constant__ CUtensorMap tmDy,
                                                                      const __grid_constant__ CUtensorMap tmX,
                                                                      const ConvWgradParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  constexpr uint32_t blk = 32u * 128u;                          // [32 pixels][32 f|c]
  const int taps = p.g.TH * p.g.TW;
  const int tap0 = blockIdx.x * p.tpc;
  const int ntap = min(p.tpc, taps - tap0);
  const int f0 = blockIdx.z * 128;
  const int fblocks = min(4, (p.F - f0 + 31) / 32);
  const int nb = p.BNc / 32;
  const bool do_bias = (p.ws_bias != nullptr) && (blockIdx.x == 0);
  constexpr uint32_t a_bytes = 4u * blk;                        // room for 128 filters
  // B region of a stage: the shifted x windows of the CTA's taps side by side, then one
  // [32 pixels][32] block of 1.0f (written once, never touched by TMA).  Side by side in shared
  // memory = side by side along N of ONE tcgen05.mma: all taps (and the bias column sums, dy x
  // ones) accumulate with 4 instructions per 256 columns instead of 4 per tap -- the MMA issuer is
  // a single thread and its instruction latency, not the tensor pipe, bounded this kernel.
  const uint32_t b_bytes = (uint32_t)(p.tpc * nb + 1) * blk;
  const uint32_t stage_bytes = a_bytes + b_bytes;
  const uint32_t tx_bytes = (uint32_t)(fblocks + ntap * nb) * blk;
  const int STAGES = p.stages;
  const uint32_t bar_base = smem_base + STAGES * stage_bytes;
  const uint32_t tmem_slot = bar_base + (2 * STAGES + 1) * 8;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (STAGES + s); };
  const uint32_t tmem_full_bar = bar_base + 8u * (2 * STAGES);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int cbeg = blockIdx.y * p.chunks_per_split;
  const int cend = min(p.g.total_chunks, cbeg + p.chunks_per_split);
  const int num_it = max(0, cend - cbeg);
  const uint32_t bias_col = (uint32_t)(p.tpc * p.BNc);
  uint32_t tcols = 32;
  while (tcols < bias_col + 32u) tcols <<= 1;

  if (warp == 0 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
    mbar_init(tmem_full_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(tcols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  for (int s = 0; s < STAGES; ++s) {
    const uint32_t ones_base = smem_base + s * stage_bytes + a_bytes + (uint32_t)(ntap * nb) * blk;
    for (uint32_t i = threadIdx.x; i < blk / 4; i += CONV_THREADS)   // all ones: invariant under the swizzle
      asm volatile("st.shared.f32 [%0], %1;" ::"r"(ones_base + 4u * i), "f"(1.0f) : "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");      // generic-proxy writes -> visible to the MMA
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  mc::pdl_sync();      // prologue done (on-chip state only); global memory from here on
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot) : "memory");

  if (warp == 0 && elect_one()) {
    // single thread: loop-invariant values in registers, chunk coordinates carried incrementally
    const int TW = p.g.TW, dh0 = p.g.dh0, dw0 = p.g.dw0, cpi = p.g.cpi, cpr = p.g.cpr, bh = p.g.bh;
    const bool wide = p.g.OWv >= 32;
    const int th0 = tap0 / TW, tw0 = tap0 - th0 * TW;
    int n = cbeg / cpi, r = cbeg - n * cpi;          // image, chunk within the image
    int row = wide ? r / cpr : r, col = wide ? r - row * cpr : 0;
    int s = 0;
    uint32_t ph = 0;
    for (int it = 0; it < num_it; ++it) {
      const int oh0 = wide ? row : row * bh, ow0 = col * 32;
      mbar_wait(empty_bar(s), ph ^ 1u);
      const uint32_t fb = full_bar(s);
      mbar_expect_tx(fb, tx_bytes);
      const uint32_t sa = smem_base + s * stage_bytes;
      uint32_t sb = sa + a_bytes;
      for (int j = 0; j < fblocks; ++j) tma_load_4d(sa + j * blk, &tmDy, fb, f0 + 32 * j, ow0, oh0, n);
      int th = th0, tw = tw0;
      for (int t = 0; t < ntap; ++t) {
        for (int j = 0; j < nb; ++j, sb += blk) tma_load_4d(sb, &tmX, fb, 32 * j, ow0 + tw + dw0, oh0 + th + dh0, n);
        if (++tw == TW) { tw = 0; ++th; }
      }
      if (++s == STAGES) { s = 0; ph ^= 1u; }
      if (wide) { if (++col == cpr) { col = 0; ++row; } }
      else ++row;
      if (++r == cpi) { r = 0; row = 0; col = 0; ++n; }
    }
  } else if (warp == 1 && elect_one()) {
    // both operands MN-major (f / channels contiguous, pixels = K): major bits 15, 16 set.
    // Filter blocks that were not loaded hold stale shared memory: they only reach accumulator
    // rows >= F, which the epilogue never reads (row m of D depends on row m of A alone).
    const uint32_t idesc0 = (1u << 4) | (2u << 7) | (2u << 10) | (1u << 15) | (1u << 16) | ((uint32_t)(128 >> 4) << 24);
    const int ntot = ntap * p.BNc + (do_bias ? 32 : 0);             // accumulator columns of this CTA
    const int ngroups = (ntot + 255) / 256;
    const uint64_t desc0 = make_desc(0, blk, 512, 1);
    int s = 0;
    uint32_t ph = 0;
    for (int it = 0; it < num_it; ++it) {
      mbar_wait(full_bar(s), ph);
      tcgen05_fence_after();
      const uint32_t sa = smem_base + s * stage_bytes, sb = sa + a_bytes;
      const uint64_t adesc = desc0 + (uint64_t)((sa & 0x3FFFFu) >> 4);
      for (int g = 0; g < ngroups; ++g) {
        const int width = min(256, ntot - g * 256);
        const uint32_t idesc = idesc0 | ((uint32_t)(width >> 3) << 17);
        const uint64_t bdesc = desc0 + (uint64_t)(((sb + (uint32_t)g * 8u * blk) & 0x3FFFFu) >> 4);
#pragma unroll
        for (int k = 0; k < 32 / UMMA_K; ++k)                        // 8 pixel rows = 1024 bytes per K step
          umma_tf32(tmem_base + (uint32_t)(g * 256), adesc + 64u * k, bdesc + 64u * k, idesc, (it > 0 || k > 0) ? 1u : 0u);
      }
      umma_commit(empty_bar(s));
      if (++s == STAGES) { s = 0; ph ^= 1u; }
    }
    umma_commit(tmem_full_bar);
  }
  __syncwarp();

  mbar_wait_warp(tmem_full_bar, 0);
  tcgen05_fence_after();
  __syncwarp();
  const int f = f0 + warp * 32 + lane;
  const uint32_t lane_addr = tmem_base + ((uint32_t)(warp * 32) << 16);
  const uint32_t stage = smem_base + (uint32_t)warp * 4096u;      // the operand ring is idle now
  const bool ws_vec = (p.Nn & 3) == 0 && (reinterpret_cast<uintptr_t>(p.ws) & 15u) == 0;
  for (int t = 0; t < ntap; ++t) {
    float* __restrict__ dst0 = p.ws + (((int64_t)blockIdx.y * taps + tap0 + t) * p.F + f0 + warp * 32) * p.Nn;
#pragma unroll 1
    for (int c0 = 0; c0 < p.BNc; c0 += 32) {
      float v[32];
      tmem_ld32(lane_addr + (uint32_t)(t * p.BNc + c0), v);
      if (num_it <= 0) {
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = 0.f;
      }
      if (f0 + warp * 32 < p.F && c0 < p.Nn)
        store_block_staged(stage, v, dst0 + c0, p.Nn, p.F - (f0 + warp * 32), p.Nn - c0, ws_vec);
    }
  }
  if (do_bias) {
    float v[32];
    tmem_ld32(lane_addr + bias_col, v);
    if (f < p.F) p.ws_bias[(int64_t)blockIdx.y * p.F + f] = (num_it > 0) ? v[0] : 0.f;
  }
  tcgen05_fence_before();
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(tcols) : "memory");
  }
}

// ------------------------------ weight gradient, halo mode ---------------------------------
// Same contraction as conv_wgrad_kernel, for layers with at most 32 contraction columns per tap
// (RGB merged rows, or <= 32 input channels) whose 128-pixel tiles are one rectangle of one image.
// A stage is a whole TILE (box_h rows x bw pixels = 128 pixels = 16 K-steps):
//   dy : one box per 32-filter block, [128 pixels][32 f]                         (MN-major A, LBO 16 KB)
//   x  : one box per column shift tw, [(box_h + TH - 1) * bw pixels][32 c]      (MN-major B)
// The row shift th of a tap is an offset of th*bw pixel rows inside the x box, so the TH taps of one
// column shift are TH overlapping N-blocks at a uniform stride of bw*128 bytes (the descriptor's
// LBO): ONE tcgen05.mma per (K-step, tw) covers them, and every input row is fetched TW times
// instead of TH*TW times in 32-row boxes (the TMA unit retires ~5 cycles per row and was the bound).
// TMEM columns: tap (th, tw) at (tw*TH + th)*32, bias (dy x ones) at TW*TH*32.
struct ConvWgradHaloParams {
  ConvGeom g;
  float* ws;             // [splits][taps][F][Nn]
  float* ws_bias;        // [splits][F] or nullptr
  int F, Nn, stages, fbmax;
  int tiles_per_split, total_tiles, tiles_per_image;
  uint32_t copy_bytes;   // box_n * (box_h + TH - 1) * bw * 128
  int kpi;               // K-steps (8 pixels) per image inside a tile: 16 / box_n
  uint32_t img_skip16;   // (TH - 1) * bw * 128 / 16: halo rows between two images of one x box (descriptor units)
  long long* dbg;        // MC_CONV_DBG: clock64 stamps of CTA 0 [role][tile][slot] (nullptr = off)
  uint32_t ring_bytes;   // stages * stage_bytes, padded so that the 4-block A descriptor of the last stage stays inside
};

MC_PROBE_DEFINE(wgh)
template <int TWC>   // column shifts known at compile time (1 = merged rows, 3 = 3x3 filter), 0 = run time
__global__ void __launch_bounds__(CONV_THREADS, 1) conv_wgrad_halo_kernel(const __grid_constant__ CUtensorMap tmDy,
                                                                           const __grid_constant__ CUtensorMap tmX,
                                                                           const ConvWgradHaloParams p) {
  extern __shared__ uint8_t smem_raw[];
  MC_STAMP(wgh, 0);
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  constexpr uint32_t fblk = 128u * 128u;                        // [128 pixels][32 f]
  const int TH = p.g.TH, TW = TWC ? TWC : p.g.TW, taps = TH * TW;
  const int f0 = blockIdx.z * (p.fbmax * 32);
  const int fblocks = min(p.fbmax, (p.F - f0 + 31) / 32);
  const int c0 = blockIdx.x * 32;                               // this CTA's 32 contraction columns of every tap
  const uint32_t a_bytes = (uint32_t)p.fbmax * fblk;
  const uint32_t stage_bytes = a_bytes + (uint32_t)TW * p.copy_bytes;
  const uint32_t tx_bytes = (uint32_t)fblocks * fblk + (uint32_t)TW * p.copy_bytes;
  const int STAGES = p.stages;
  const uint32_t ones_base = smem_base + p.ring_bytes;          // [32 pixels][32] of 1.0f
  const uint32_t bar_base = ones_base + 4096u;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (STAGES + s); };
  const uint32_t tmem_full_bar = bar_base + 8u * (2 * STAGES);
  const uint32_t tmem_slot = tmem_full_bar + 8u;
  const bool do_bias = p.ws_bias != nullptr && blockIdx.x == 0;
  const uint32_t bias_col = (uint32_t)(taps * 32);
  uint32_t tcols = 32;
  while (tcols < bias_col + 32u) tcols <<= 1;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tbeg = blockIdx.y * p.tiles_per_split;
  const int tend = min(p.total_tiles, tbeg + p.tiles_per_split);
  const int num_it = max(0, tend - tbeg);

  if (warp == 0 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
    mbar_init(tmem_full_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(tcols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  for (uint32_t i = threadIdx.x; i < 1024u; i += CONV_THREADS)
    asm volatile("st.shared.f32 [%0], %1;" ::"r"(ones_base + 4u * i), "f"(1.0f) : "memory");
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  MC_STAMP(wgh, 1);
  mc::pdl_sync();      // prologue done (on-chip state only); global memory from here on
  MC_STAMP(wgh, 2);
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot) : "memory");

  if (warp == 0 && elect_one()) {
    const int dh0 = p.g.dh0, dw0 = p.g.dw0, BH = p.g.box_h, tpi = p.tiles_per_image, bn = p.g.box_n;
    int n = tbeg / tpi, r = tbeg - n * tpi;
    n *= bn;                                                     // bn > 1: whole images, tpi == 1
    int s = 0;
    uint32_t ph = 0;
    for (int it = 0; it < num_it; ++it) {
      const int oh0 = r * BH;
      const bool dbgw = p.dbg && blockIdx.x + blockIdx.y + blockIdx.z == 0 && it < 32;
      if (dbgw) p.dbg[(0 * 32 + it) * 4 + 0] = clock64();
      mbar_wait(empty_bar(s), ph ^ 1u);
      if (dbgw) p.dbg[(0 * 32 + it) * 4 + 1] = clock64();
      const uint32_t fb = full_bar(s);
      mbar_expect_tx(fb, tx_bytes);
      const uint32_t sa = smem_base + s * stage_bytes;
      for (int j = 0; j < fblocks; ++j) tma_load_4d(sa + j * fblk, &tmDy, fb, f0 + 32 * j, 0, oh0, n);
      uint32_t sb = sa + a_bytes;
      for (int tw = 0; tw < TW; ++tw, sb += p.copy_bytes) tma_load_4d(sb, &tmX, fb, c0, tw + dw0, oh0 + dh0, n);
      if (dbgw) p.dbg[(0 * 32 + it) * 4 + 2] = clock64();
      if (++s == STAGES) { s = 0; ph ^= 1u; }
      if (++r == tpi) { r = 0; n += bn; }
    }
  } else if (warp == 1 && elect_one()) {
    const uint32_t idesc0 = (1u << 4) | (2u << 7) | (2u << 10) | (1u << 15) | (1u << 16) | ((uint32_t)(128 >> 4) << 24);
    const uint32_t idesc_t = idesc0 | ((uint32_t)((TH * 32) >> 3) << 17);       // TH taps of one column shift
    const uint32_t idesc_b = idesc0 | ((uint32_t)(32 >> 3) << 17);
    const uint64_t adesc0 = make_desc(0, fblk, 512, 1);                         // filter blocks 16 KB apart
    const uint64_t bdesc0 = make_desc(0, (uint32_t)p.g.bw * 128u, 512, 1);      // taps bw pixel rows apart
    const uint64_t odesc = make_desc(ones_base, 4096, 512, 1);
    const uint32_t tw_cols = (uint32_t)(TH * 32);
    int s = 0;
    uint32_t ph = 0;
    for (int it = 0; it < num_it; ++it) {
      const bool dbgw = p.dbg && blockIdx.x + blockIdx.y + blockIdx.z == 0 && it < 32;
      if (dbgw) p.dbg[(1 * 32 + it) * 4 + 0] = clock64();
      mbar_wait(full_bar(s), ph);
      tcgen05_fence_after();
      if (dbgw) p.dbg[(1 * 32 + it) * 4 + 1] = clock64();
      const uint32_t sa = smem_base + s * stage_bytes, sb = sa + a_bytes;
      uint64_t adesc = adesc0 + (uint64_t)((sa & 0x3FFFFu) >> 4);
      uint64_t bdesc = bdesc0 + (uint64_t)((sb & 0x3FFFFu) >> 4);
      const uint64_t cstep = (uint64_t)(p.copy_bytes >> 4);
      const int kmask = p.kpi - 1;                                              // kpi: 4, 8 or 16
      const uint64_t skip = (uint64_t)p.img_skip16;
#pragma unroll 4
      for (int kk = 0; kk < 16; ++kk, adesc += 64u, bdesc += 64u) {             // 8 pixel rows = 1024 bytes per K step
        const uint32_t accf = (it > 0 || kk > 0) ? 1u : 0u;
        if (kk > 0 && (kk & kmask) == 0) bdesc += skip;                         // next image of the box: over its halo rows
        if (TWC) {
#pragma unroll
          for (int tw = 0; tw < TWC; ++tw) umma_tf32(tmem_base + tw * tw_cols, adesc, bdesc + tw * cstep, idesc_t, accf);
        } else {
          uint64_t bd = bdesc;
          uint32_t col = tmem_base;
          for (int tw = 0; tw < TW; ++tw, bd += cstep, col += tw_cols) umma_tf32(col, adesc, bd, idesc_t, accf);
        }
        if (do_bias) umma_tf32(tmem_base + bias_col, adesc, odesc, idesc_b, accf);
      }
      umma_commit(empty_bar(s));
      if (dbgw) p.dbg[(1 * 32 + it) * 4 + 2] = clock64();
      if (++s == STAGES) { s = 0; ph ^= 1u; }
    }
    umma_commit(tmem_full_bar);
  }
  __syncwarp();

  if (p.dbg && blockIdx.x + blockIdx.y + blockIdx.z == 0 && threadIdx.x == 64) p.dbg[(2 * 32 + 0) * 4 + 0] = clock64();
  mbar_wait_warp(tmem_full_bar, 0);
  tcgen05_fence_after();
  __syncwarp();
  MC_STAMP(wgh, 3);
  if (p.dbg && blockIdx.x + blockIdx.y + blockIdx.z == 0 && threadIdx.x == 64) p.dbg[(2 * 32 + 0) * 4 + 1] = clock64();
  const int f = f0 + warp * 32 + lane;
  const uint32_t lane_addr = tmem_base + ((uint32_t)(warp * 32) << 16);
  const uint32_t stage = smem_base + (uint32_t)warp * 4096u;      // the operand ring is idle now
  const bool ws_vec = (p.Nn & 3) == 0 && (reinterpret_cast<uintptr_t>(p.ws) & 15u) == 0;
  for (int tw = 0; tw < TW; ++tw) {
    for (int th = 0; th < TH; ++th) {
      float v[32];
      tmem_ld32(lane_addr + (uint32_t)((tw * TH + th) * 32), v);
      if (num_it <= 0) {
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = 0.f;
      }
      if (warp < fblocks && f0 + warp * 32 < p.F && c0 < p.Nn) {
        float* __restrict__ dst0 = p.ws + (((int64_t)blockIdx.y * taps + th * TW + tw) * p.F + f0 + warp * 32) * p.Nn + c0;
        store_block_staged(stage, v, dst0, p.Nn, p.F - (f0 + warp * 32), p.Nn - c0, ws_vec);
      }
    }
  }
  if (do_bias) {
    float v[32];
    tmem_ld32(lane_addr + bias_col, v);
    if (f < p.F && warp < fblocks) p.ws_bias[(int64_t)blockIdx.y * p.F + f] = (num_it > 0) ? v[0] : 0.f;
  }
  if (p.dbg && blockIdx.x + blockIdx.y + blockIdx.z == 0 && threadIdx.x == 64) p.dbg[(2 * 32 + 0) * 4 + 2] = clock64();
  MC_STAMP(wgh, 4);
  tcgen05_fence_before();
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(tcols) : "memory");
  }
  MC_STAMP(wgh, 5);
}

// dw (F, C, KH, KW) = sum_z ws[z][tap][f][n] (merged: tap = kh, n = kw*Cp + c; else tap = kh*KW+kw, n = c)
// and, in the trailing blocks of the same launch, dbias[f] = sum_z ws_bias[z][f].
// A block is 32 consecutive workspace columns x 8 slices of the split range: the reads are
// coalesced along the workspace row and 8-way parallel along z (the first version summed up to
// 148 splits serially per thread: 16-26 us for a few thousand outputs); the slices are combined
// in a fixed order, so the result is deterministic.
__global__ void __launch_bounds__(256) conv_wgrad_reduce_kernel(const float* __restrict__ ws, float* __restrict__ dw,
                                                                const float* __restrict__ wsb, float* __restrict__ dbias,
                                                                int splits, int taps, int F, int Nn, int C, int KH, int KW,
                                                                int Cp, int merged, int wblocks) {
  mc::pdl_sync();
  __shared__ float part[8][33];
  const int lane = threadIdx.x & 31, slice = threadIdx.x >> 5;
  const bool bias_block = (int)blockIdx.x >= wblocks;
  const int64_t total = bias_block ? F : (int64_t)taps * F * Nn;
  const float* __restrict__ src = bias_block ? wsb : ws;
  const int64_t j = (int64_t)(bias_block ? blockIdx.x - wblocks : blockIdx.x) * 32 + lane;
  float acc = 0.f;
  if (j < total) {
    int z = slice;
    for (; z + 24 < splits; z += 32) {     // 4 independent loads in flight
      const float a0 = src[(int64_t)z * total + j], a1 = src[(int64_t)(z + 8) * total + j];
      const float a2 = src[(int64_t)(z + 16) * total + j], a3 = src[(int64_t)(z + 24) * total + j];
      acc += a0; acc += a1; acc += a2; acc += a3;
    }
    for (; z < splits; z += 8) acc += src[(int64_t)z * total + j];
  }
  part[slice][lane] = acc;
  __syncthreads();
  if (slice != 0 || j >= total) return;
  float sum = part[0][lane];
#pragma unroll
  for (int q = 1; q < 8; ++q) sum += part[q][lane];
  if (bias_block) { dbias[j] = sum; return; }
  const int n = (int)(j % Nn), f = (int)((j / Nn) % F), tap = (int)(j / ((int64_t)Nn * F));
  int kh, kw, c;
  if (merged) { kh = tap; kw = n / Cp; c = n - kw * Cp; }
  else { kh = tap / KW; kw = tap - kh * KW; c = n; }
  if (c < C && kw < KW) dw[(((int64_t)f * C + c) * KH + kh) * KW + kw] = sum;
}

// ----------------------------------- layout conversion -------------------------------------
// y (N, H, W + 2*wpad, Cp) <- x (N, C, H, W); pad pixels and pad channels are zero.
// One CTA: 32 channels x 32 (padded) pixels of one (n, h) row, transposed through shared memory
// so that both the NCHW reads (along w) and the NHWC writes (along c) are coalesced.
__global__ void __launch_bounds__(256) nchw_to_nhwc_kernel(const float* __restrict__ x, float* __restrict__ y, int N, int C,
                                                           int H, int W, int Cp, int wpad) {
  mc::pdl_sync();
  __shared__ float tile[32][33];
  const int Wp = W + 2 * wpad;
  const int n = blockIdx.z / H, h = blockIdx.z % H;
  const int c0 = blockIdx.y * 32, w0 = blockIdx.x * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;   // 32 x 8
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int c = c0 + ty + 8 * j, w = w0 + tx - wpad;      // w0 indexes the padded row
    float v = 0.f;
    if (c < C && w >= 0 && w < W) v = x[(((int64_t)n * C + c) * H + h) * W + w];
    tile[ty + 8 * j][tx] = v;
  }
  __syncthreads();
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int wp = w0 + ty + 8 * j, c = c0 + tx;
    if (wp < Wp && c < Cp) y[(((int64_t)n * H + h) * Wp + wp) * Cp + c] = tile[tx][ty + 8 * j];
  }
}

// y (N, C, H, W) <- x (N, H, W, Cp)
__global__ void __launch_bounds__(256) nhwc_to_nchw_kernel(const float* __restrict__ x, float* __restrict__ y, int N, int C,
                                                           int H, int W, int Cp) {
  mc::pdl_sync();
  __shared__ float tile[32][33];
  const int n = blockIdx.z / H, h = blockIdx.z % H;
  const int c0 = blockIdx.y * 32, w0 = blockIdx.x * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int w = w0 + ty + 8 * j, c = c0 + tx;
    float v = 0.f;
    if (w < W && c < C) v = x[(((int64_t)n * H + h) * W + w) * Cp + c];
    tile[ty + 8 * j][tx] = v;
  }
  __syncthreads();
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int c = c0 + ty + 8 * j, w = w0 + tx;
    if (c < C && w < W) y[(((int64_t)n * C + c) * H + h) * W + w] = tile[tx][ty + 8 * j];
  }
}

// channels-last max pooling (window KH x KW, stride s, 'valid'); idx = kh*KW + kw of the FIRST
// maximal element in row-major window order (the arg-max the NCHW kernel and the oracle pick).
__global__ void __launch_bounds__(256) maxpool_nhwc_fwd_kernel(const float* __restrict__ x, float* __restrict__ y,
                                                               uint8_t* __restrict__ idx, int N, int H, int W, int Cp, int KH,
                                                               int KW, int stride, int OH, int OW) {
  mc::pdl_sync();
  const int64_t total = (int64_t)N * OH * OW * Cp;
  for (int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x; t < total; t += (int64_t)gridDim.x * 256) {
    const int c = (int)(t % Cp);
    const int ow = (int)((t / Cp) % OW), oh = (int)((t / ((int64_t)Cp * OW)) % OH);
    const int n = (int)(t / ((int64_t)Cp * OW * OH));
    float best = 0.f;
    int bi = -1;
    for (int kh = 0; kh < KH; ++kh)
      for (int kw = 0; kw < KW; ++kw) {
        const float v = x[(((int64_t)n * H + oh * stride + kh) * W + ow * stride + kw) * Cp + c];
        if (bi < 0 || v > best || (v != v && best == best)) { best = v; bi = kh * KW + kw; }
      }
    y[t] = best;
    if (idx) idx[t] = (uint8_t)bi;
  }
}

// dx (N, H, W, Cp): non-overlapping windows (stride == kernel) -> every input element belongs to
// at most one window: a pure gather, no atomics, no memset.  If `relu_y` is given (the pooled
// activation of a relu'd tensor), the relu mask of the producer is applied here as well:
// dx = dy * (y > 0) at the arg-max position (MXNet Activation backward: out_grad * (out > 0)).
__global__ void __launch_bounds__(256) maxpool_nhwc_bwd_kernel(const float* __restrict__ dy, const uint8_t* __restrict__ idx,
                                                               const float* __restrict__ relu_y, float* __restrict__ dx, int N,
                                                               int H, int W, int Cp, int K, int OH, int OW) {
  mc::pdl_sync();
  const int64_t total = (int64_t)N * H * W * Cp;
  for (int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x; t < total; t += (int64_t)gridDim.x * 256) {
    const int c = (int)(t % Cp);
    const int w = (int)((t / Cp) % W), h = (int)((t / ((int64_t)Cp * W)) % H);
    const int n = (int)(t / ((int64_t)Cp * W * H));
    const int oh = h / K, ow = w / K;
    float g = 0.f;
    if (oh < OH && ow < OW) {
      const int64_t o = (((int64_t)n * OH + oh) * OW + ow) * Cp + c;
      if ((int)idx[o] == (h - oh * K) * K + (w - ow * K)) {
        g = dy[o];
        if (relu_y && !(relu_y[o] > 0.f)) g = 0.f;
      }
    }
    dx[t] = g;
  }
}

// 128-bit versions: one thread = 4 channels of one pooled pixel (Cp is a multiple of 4).
__global__ void __launch_bounds__(256) maxpool_nhwc_fwd_vec_kernel(const float4* __restrict__ x, float4* __restrict__ y,
                                                                   uchar4* __restrict__ idx, int N, int H, int W, int C4,
                                                                   int KH, int KW, int stride, int OH, int OW) {
  mc::pdl_sync();
  const int64_t total = (int64_t)N * OH * OW * C4;
  for (int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x; t < total; t += (int64_t)gridDim.x * 256) {
    const int c = (int)(t % C4);
    const int ow = (int)((t / C4) % OW), oh = (int)((t / ((int64_t)C4 * OW)) % OH);
    const int n = (int)(t / ((int64_t)C4 * OW * OH));
    float4 best = make_float4(0.f, 0.f, 0.f, 0.f);
    uchar4 bi = make_uchar4(255, 255, 255, 255);
    for (int kh = 0; kh < KH; ++kh)
      for (int kw = 0; kw < KW; ++kw) {
        const float4 v = __ldg(x + (((int64_t)n * H + oh * stride + kh) * W + ow * stride + kw) * C4 + c);
        const unsigned char k = (unsigned char)(kh * KW + kw);
#define MC_POOL_PICK(f)                                                                   \
        if (bi.f == 255 || v.f > best.f || (v.f != v.f && best.f == best.f)) { best.f = v.f; bi.f = k; }
        MC_POOL_PICK(x) MC_POOL_PICK(y) MC_POOL_PICK(z) MC_POOL_PICK(w)
#undef MC_POOL_PICK
      }
    y[t] = best;
    if (idx) idx[t] = bi;
  }
}

// requires H == OH*K and W == OW*K: every input element is written exactly once
__global__ void __launch_bounds__(256) maxpool_nhwc_bwd_vec_kernel(const float4* __restrict__ dy, const uchar4* __restrict__ idx,
                                                                   const float4* __restrict__ relu_y, float4* __restrict__ dx,
                                                                   int N, int H, int W, int C4, int K, int OH, int OW) {
  mc::pdl_sync();
  const int64_t total = (int64_t)N * OH * OW * C4;
  for (int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x; t < total; t += (int64_t)gridDim.x * 256) {
    const int c = (int)(t % C4);
    const int ow = (int)((t / C4) % OW), oh = (int)((t / ((int64_t)C4 * OW)) % OH);
    const int n = (int)(t / ((int64_t)C4 * OW * OH));
    float4 g = __ldg(dy + t);
    const uchar4 bi = idx[t];
    if (relu_y) {
      const float4 r = __ldg(relu_y + t);
      if (!(r.x > 0.f)) g.x = 0.f;
      if (!(r.y > 0.f)) g.y = 0.f;
      if (!(r.z > 0.f)) g.z = 0.f;
      if (!(r.w > 0.f)) g.w = 0.f;
    }
    for (int kh = 0; kh < K; ++kh)
      for (int kw = 0; kw < K; ++kw) {
        const unsigned char k = (unsigned char)(kh * K + kw);
        const float4 o = make_float4(bi.x == k ? g.x : 0.f, bi.y == k ? g.y : 0.f, bi.z == k ? g.z : 0.f, bi.w == k ? g.w : 0.f);
        dx[(((int64_t)n * H + oh * K + kh) * W + ow * K + kw) * C4 + c] = o;
      }
  }
}

// few channels (C <= 4 -> Cp == 4): one thread = one (padded) pixel, one 16-byte store
__global__ void __launch_bounds__(256) nchw_to_nhwc4_kernel(const float* __restrict__ x, float4* __restrict__ y, int N, int C,
                                                            int H, int W, int wpad) {
  mc::pdl_sync();
  const int Wp = W + 2 * wpad;
  const int64_t total = (int64_t)N * H * Wp;
  const int64_t plane = (int64_t)H * W;
  for (int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x; t < total; t += (int64_t)gridDim.x * 256) {
    const int wp = (int)(t % Wp);
    const int h = (int)((t / Wp) % H);
    const int64_t n = t / ((int64_t)Wp * H);
    const int w = wp - wpad;
    float v[4] = {0.f, 0.f, 0.f, 0.f};
    if (w >= 0 && w < W) {
      const float* __restrict__ src = x + n * C * plane + (int64_t)h * W + w;
      for (int c = 0; c < C; ++c) v[c] = __ldg(src + c * plane);
    }
    y[t] = make_float4(v[0], v[1], v[2], v[3]);
  }
}

// ------------------------------------- host side --------------------------------------------
// tile_box: round the virtual height so that the 4 chunks of every 128-pixel tile are one
// rectangle {bw, box_h rows, box_n images} -- one TMA box per operand tile instead of four (the
// producer thread is issue-bound: ~130 ns per UTMALDG.4D, profiles/r1i).
static ConvGeom make_geom(int N, int OH, int OW, bool tile_box = false) {
  ConvGeom g;
  g.N = N; g.OH = OH; g.OW = OW;
  if (OW <= 8) g.OWv = 8;
  else if (OW <= 16) g.OWv = 16;
  else g.OWv = (OW + 31) / 32 * 32;
  g.bw = g.OWv < 32 ? g.OWv : 32;
  g.bh = 32 / g.bw;
  g.OHv = (OH + g.bh - 1) / g.bh * g.bh;
  g.one_box = 0; g.box_h = g.bh; g.box_n = 1;
  if (tile_box && g.OWv <= 32) {
    const int BH = 128 / g.bw;                 // rows of a whole tile
    if (OH <= BH) {
      int ohv = g.bh;
      while (ohv < OH) ohv <<= 1;              // power of two: divides BH
      g.OHv = ohv; g.box_h = ohv; g.box_n = BH / ohv;
    } else {
      g.OHv = (OH + BH - 1) / BH * BH; g.box_h = BH; g.box_n = 1;
    }
    g.one_box = 1;
  }
  g.cpr = g.OWv >= 32 ? g.OWv / 32 : 1;
  g.cpi = (g.OHv / g.bh) * g.cpr;
  g.total_chunks = N * g.cpi;
  g.dh0 = g.dw0 = 0; g.TH = g.TW = 1;
  g.cpi_shift = -1;
  for (int b = 0; b < 30; ++b) if ((1 << b) == g.cpi) g.cpi_shift = b;
  g.owv_shift = g.OWv == 8 ? 3 : (g.OWv == 16 ? 4 : 5);
  return g;
}

// Channels-last activation (N, H, Wt, Cp) as the 4-D map (K, Wdim, H, N): dim 0 is a contiguous
// run of K floats starting at a pixel (K == Cp: that pixel's channels; merged: KW pixels' worth,
// overlapping the next start pixel), dim 1 walks Wdim start pixels.  Box {32, bw, bh, 1}.
static int make_act_tmap(CUtensorMap* map, const float* ptr, int N, int H, int Wt, int Cp, int K, int Wdim, int bw, int bh,
                         bool atom32, int bn = 1) {
  const uint64_t dims[4] = {(uint64_t)K, (uint64_t)Wdim, (uint64_t)H, (uint64_t)N};
  const uint64_t strides[3] = {(uint64_t)Cp * 4, (uint64_t)Wt * Cp * 4, (uint64_t)H * Wt * Cp * 4};
  const uint32_t box[4] = {32, (uint32_t)bw, (uint32_t)bh, (uint32_t)bn};
  return make_tmap(map, ptr, 4, dims, strides, box, atom32);
}

static int set_max_smem(const void* fn, const char* what) {
  cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(220 * 1024));
  if (e != cudaSuccess) { set_error("cudaFuncSetAttribute(%s): %s", what, cudaGetErrorString(e)); return 1; }
  return 0;
}

// y (N, OH, OW, ldy) = conv over x (N, H, Wt, Cp) with `Kreal` contraction channels per pixel.
// mode 0 forward, 1 forward merged (x physically W-padded by `pad`), 2 dgrad (x := dy, w flipped,
// `pad` already K-1-pad).  Nout = output channels (F for forward, C for dgrad).
// 2x2 / stride-2 max pooling in the epilogue is possible when a pooling window never straddles two tiles
// and the channel pitch is a whole number of 32-channel epilogue chunks
static bool conv_pool_ok(int OH, int OW, int Nout, int ldy) {
  if ((OH & 1) || (OW & 1) || OW > 32 || (Nout & 31) || ldy != Nout) return false;
  const ConvGeom g = make_geom(1, OH, OW, true);
  return g.OWv < 32 ? (g.bh % 2 == 0) : (g.OWv == 32 && g.cpi % 4 == 0);
}

int conv_igemm(const float* x, const float* w, const float* bias, float* y, int N, int H, int Wt, int Cp, int Kreal,
               int Nout, int ldy, int F, int C, int KH, int KW, int pad, int OH, int OW, int relu, int mode,
               int pool = 0, uint8_t* pidx = nullptr) {
  ConvFwdParams p;
  p.g = make_geom(N, OH, OW, true);
  p.pool = pool; p.pidx = pidx; p.pool_smem = 0;
  if (pool && !conv_pool_ok(OH, OW, Nout, ldy)) { set_error("conv_igemm: shape outside the pooled epilogue"); return 2; }
  int K, Wdim;                     // contraction length per tap; number of start pixels along W
  if (mode == 1) { p.g.TH = KH; p.g.TW = 1; p.g.dh0 = -pad; p.g.dw0 = 0; K = KW * Cp; Wdim = Wt - KW + 1; }
  else { p.g.TH = KH; p.g.TW = KW; p.g.dh0 = -pad; p.g.dw0 = -pad; K = Kreal; Wdim = Wt; }
  const int taps = p.g.TH * p.g.TW;
  const int Kp = (K + 3) / 4 * 4;
  ScratchFrame scratch;
  float* wp = scratch.floats((size_t)taps * Nout * Kp);
  if (!wp) return 1;
  mc::launch(conv_pack_weights_kernel, grid_for((int64_t)taps * Nout * Kp, 256), 256, 0, w, wp, F, C, KH, KW, Cp, taps, Nout,
                                                                                          Kp, mode);
  MC_LAUNCHED();

  p.y = y; p.bias = bias; p.Nout = Nout; p.ldy = ldy; p.relu = relu;
  p.ncb = (K + BKC - 1) / BKC;
  int bn = 32;
  while (bn < Nout && bn < 256) bn <<= 1;
  p.BN = bn;
  // TMEM ring: as many accumulator tiles as fit in the 512 columns (power of two, at most 8): the
  // tcgen05.commit -> epilogue -> release chain of one tile is ~1 us, which two buffers cannot hide
  // when a tile is only a few k-blocks long
  int nacc = 2;
  while (nacc * 2 * p.BN <= 512 && nacc < 8) nacc <<= 1;
  p.nacc = nacc;
  const int ntiles_n = (Nout + p.BN - 1) / p.BN;
  const size_t w_bytes = (size_t)taps * p.ncb * p.BN * 128;
  p.wres = (ntiles_n == 1 && w_bytes <= 96 * 1024) ? 1 : 0;
  // halo mode: needs resident weights, one image per tile box, and room for >= 2 stages
  p.halo = 0;
  p.copy_bytes = (uint32_t)((p.g.box_h + p.g.TH - 1) * p.g.bw * 128);
  if (p.wres && p.g.one_box && p.g.box_n == 1 && taps > 1 && getenv("MC_CONV_NO_HALO") == nullptr &&
      2 * (size_t)p.g.TW * p.copy_bytes + w_bytes <= 200 * 1024)
    p.halo = 1;
  const size_t stage_bytes = p.halo ? (size_t)p.g.TW * p.copy_bytes : 4 * 4096 + (p.wres ? 0 : (size_t)p.BN * 128);
  int stages = (int)((200 * 1024 - (p.wres ? w_bytes : 0)) / stage_bytes);
  if (stages > 8) stages = 8;
  if (stages < 1) stages = 1;
  p.stages = stages;
  size_t smem = (p.wres ? w_bytes : 0) + stages * stage_bytes + (2 * stages + 2 * nacc + 1) * 8 + 16 + 1024;
  if (pool && p.g.OWv >= 32) {          // cross-warp exchange area of the pooled epilogue: 4 warp pairs
    const size_t used = (p.wres ? w_bytes : 0) + stages * stage_bytes + (2 * stages + 2 * nacc + 1) * 8 + 16;
    p.pool_smem = (uint32_t)((used + 15) & ~(size_t)15);
    smem = p.pool_smem + 4 * 2 * 3072 + 1024;        // 4 warp pairs x 2 buffers x (64 lanes x 48 bytes)
  }

  CUtensorMap tmX, tmW;
  int rc = make_act_tmap(&tmX, x, N, H, Wt, Cp, K, Wdim, p.g.bw,
                         p.halo ? p.g.box_h + p.g.TH - 1 : (p.g.one_box ? p.g.box_h : p.g.bh), false,
                         p.g.one_box ? p.g.box_n : 1);
  if (rc == 0) {
    const uint64_t dims[3] = {(uint64_t)Kp, (uint64_t)Nout, (uint64_t)taps};
    const uint64_t strides[2] = {(uint64_t)Kp * 4, (uint64_t)Kp * Nout * 4};
    const uint32_t box[3] = {32, (uint32_t)p.BN, 1};
    rc = make_tmap(&tmW, wp, 3, dims, strides, box, false);
  }
  static uint32_t configured = 0;      // function attributes are per device
  if (rc == 0 && first_use_on_device(&configured)) rc = set_max_smem((const void*)conv_igemm_kernel, "conv_igemm_kernel");
  if (rc == 0) {
    const int64_t tiles = (int64_t)((p.g.total_chunks + 3) / 4) * ((Nout + p.BN - 1) / p.BN);
    const unsigned grid = (unsigned)(tiles < sm_count() ? tiles : sm_count());
    p.dbg = nullptr;
    static long long* dbg_buf = nullptr;
    const bool dbg = getenv("MC_CONV_DBG") != nullptr;
    if (dbg) {
      if (!dbg_buf) cudaMalloc((void**)&dbg_buf, sizeof(long long) * 3 * 32 * 4);
      cudaMemset(dbg_buf, 0, sizeof(long long) * 3 * 32 * 4);
      p.dbg = dbg_buf;
    }
    mc::launch_pdl(conv_igemm_kernel, grid, IGEMM_THREADS, smem, tmX, tmW, p);
    if (dbg) {
      static long long host[3 * 32 * 4];
      cudaDeviceSynchronize();
      cudaMemcpy(host, dbg_buf, sizeof(host), cudaMemcpyDeviceToHost);
      const long long t0 = host[0];
      fprintf(stderr, "conv dbg: tiles/CTA~%d num_it=%d stages=%d nacc=%d wres=%d\n", (int)(tiles / grid), taps * p.ncb, p.stages, p.nacc, p.wres);
      for (int t = 0; t < 12; ++t)
        fprintf(stderr, "  tile %2d  prod start %6lld first-wait-done %6lld issued %6lld | mma start %6lld acc-free %6lld first-full %6lld committed %6lld | epi start %6lld full %6lld done %6lld\n", t,
                host[(0 * 32 + t) * 4 + 0] - t0, host[(0 * 32 + t) * 4 + 1] - t0, host[(0 * 32 + t) * 4 + 2] - t0,
                host[(1 * 32 + t) * 4 + 0] - t0, host[(1 * 32 + t) * 4 + 1] - t0, host[(1 * 32 + t) * 4 + 2] - t0, host[(1 * 32 + t) * 4 + 3] - t0,
                host[(2 * 32 + t) * 4 + 0] - t0, host[(2 * 32 + t) * 4 + 1] - t0, host[(2 * 32 + t) * 4 + 2] - t0);
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { set_error("conv_igemm_kernel launch failed: %s", cudaGetErrorString(e)); rc = 1; }
    else count_launch();
  }
  return rc;
}

// dw (F, C, KH, KW) [and dbias (F,) when non-null] from x (N, H, Wt, Cp) and dy (N, OH, OW, Fp);
// merged as in conv_igemm mode 1.
int conv_wgrad_tc(const float* x, const float* dy, float* dw, float* dbias, int N, int H, int Wt, int Cp, int C, int F, int Fp,
                  int KH, int KW, int pad, int OH, int OW, int merged) {
  ConvWgradParams p;
  p.g = make_geom(N, OH, OW);
  int Wdim;
  if (merged) { p.g.TH = KH; p.g.TW = 1; p.g.dh0 = -pad; p.g.dw0 = 0; p.Nn = KW * Cp; Wdim = Wt - KW + 1; }
  else { p.g.TH = KH; p.g.TW = KW; p.g.dh0 = -pad; p.g.dw0 = -pad; p.Nn = Cp; Wdim = Wt; }
  const int taps = p.g.TH * p.g.TW;
  p.F = F;
  p.BNc = (p.Nn + 31) / 32 * 32;
  if (p.BNc > 256) { set_error("conv_wgrad_tc: more than 256 input channels is not supported"); return 2; }
  {
    // halo mode (conv_wgrad_halo_kernel): <= 32 columns per tap, rectangular one-image tiles, room for >= 2 stages
    ConvGeom hg = make_geom(N, OH, OW, true);
    hg.TH = p.g.TH; hg.TW = p.g.TW; hg.dh0 = p.g.dh0; hg.dw0 = p.g.dw0;
    int fbmax = F >= 128 ? 4 : (F + 31) / 32;                    // 32-filter blocks per CTA: fewer when the x boxes are large
    const size_t copy_bytes = (size_t)hg.box_n * (hg.box_h + hg.TH - 1) * hg.bw * 128;
    size_t stage_bytes = 0, ring = 0;
    int stages = 0;
    for (;; fbmax = (fbmax + 1) / 2) {
      stage_bytes = (size_t)fbmax * 16384 + (size_t)hg.TW * copy_bytes;
      stages = (int)((196 * 1024) / stage_bytes);
      if (stages > 4) stages = 4;
      ring = (size_t)stages * stage_bytes;
      if (stages >= 1 && (size_t)(stages - 1) * stage_bytes + 65536 > ring) ring = (size_t)(stages - 1) * stage_bytes + 65536;
      if ((stages >= 2 && ring <= 200 * 1024) || fbmax == 1) break;
    }
    const int cblocks = p.BNc / 32;
    if (hg.one_box && taps * 32 + 32 <= 512 && stages >= 2 && ring <= 200 * 1024 && getenv("MC_CONV_NO_HALO") == nullptr) {
      ConvWgradHaloParams h;
      h.g = hg; h.F = F; h.Nn = p.Nn; h.stages = stages; h.fbmax = fbmax;
      h.copy_bytes = (uint32_t)copy_bytes; h.ring_bytes = (uint32_t)ring;
      h.kpi = 16 / hg.box_n;
      h.img_skip16 = (uint32_t)((hg.TH - 1) * hg.bw * 128) >> 4;
      h.tiles_per_image = hg.OHv / hg.box_h;
      h.total_tiles = (N + hg.box_n - 1) / hg.box_n * h.tiles_per_image;
      const int ftiles = (F + fbmax * 32 - 1) / (fbmax * 32);
      int splits = sm_count() / (ftiles * cblocks);
      if (splits > h.total_tiles) splits = h.total_tiles;
      if (splits < 1) splits = 1;
      h.tiles_per_split = (h.total_tiles + splits - 1) / splits;
      splits = (h.total_tiles + h.tiles_per_split - 1) / h.tiles_per_split;
      const size_t ws_elems = (size_t)splits * taps * F * p.Nn;
      ScratchFrame scratch;
      float* ws = scratch.floats(ws_elems + (dbias ? (size_t)splits * F : 0));
      if (!ws) return 1;
      h.ws = ws;
      h.ws_bias = dbias ? ws + ws_elems : nullptr;
      CUtensorMap tmDy, tmX;
      int rc = make_act_tmap(&tmDy, dy, N, OH, OW, Fp, Fp, OW, hg.bw, hg.box_h, true, hg.box_n);
      if (rc == 0) rc = make_act_tmap(&tmX, x, N, H, Wt, Cp, p.Nn, Wdim, hg.bw, hg.box_h + hg.TH - 1, true, hg.box_n);
      static uint32_t configured = 0;
      if (rc == 0 && first_use_on_device(&configured)) {
        rc = set_max_smem((const void*)conv_wgrad_halo_kernel<0>, "conv_wgrad_halo_kernel") ||
             set_max_smem((const void*)conv_wgrad_halo_kernel<1>, "conv_wgrad_halo_kernel") ||
             set_max_smem((const void*)conv_wgrad_halo_kernel<3>, "conv_wgrad_halo_kernel");
      }
      if (rc == 0) {
        const size_t smem = ring + 4096 + (2 * stages + 1) * 8 + 16 + 1024;
        dim3 grid((unsigned)cblocks, (unsigned)splits, (unsigned)ftiles);
        h.dbg = nullptr;
        static long long* dbg_buf = nullptr;
        const bool dbg = getenv("MC_CONV_DBG") != nullptr;
        if (dbg) {
          if (!dbg_buf) cudaMalloc((void**)&dbg_buf, sizeof(long long) * 3 * 32 * 4);
          cudaMemset(dbg_buf, 0, sizeof(long long) * 3 * 32 * 4);
          h.dbg = dbg_buf;
        }
        if (hg.TW == 1) mc::launch_pdl(conv_wgrad_halo_kernel<1>, grid, CONV_THREADS, smem, tmDy, tmX, h);
        else if (hg.TW == 3) mc::launch_pdl(conv_wgrad_halo_kernel<3>, grid, CONV_THREADS, smem, tmDy, tmX, h);
        else mc::launch_pdl(conv_wgrad_halo_kernel<0>, grid, CONV_THREADS, smem, tmDy, tmX, h);
        if (dbg) {
          static long long host[3 * 32 * 4];
          cudaDeviceSynchronize();
          cudaMemcpy(host, dbg_buf, sizeof(host), cudaMemcpyDeviceToHost);
          const long long t0 = host[0];
          fprintf(stderr, "wgrad halo dbg: tiles/CTA=%d stages=%d fbmax=%d grid=(%d,%d,%d) stage_bytes=%d\n", h.tiles_per_split, stages, fbmax, cblocks, splits, ftiles, (int)stage_bytes);
          for (int t = 0; t < 10 && t < h.tiles_per_split; ++t)
            fprintf(stderr, "  tile %2d  prod start %6lld slot-free %6lld issued %6lld | mma start %6lld full %6lld committed %6lld\n", t,
                    host[(0 * 32 + t) * 4 + 0] - t0, host[(0 * 32 + t) * 4 + 1] - t0, host[(0 * 32 + t) * 4 + 2] - t0,
                    host[(1 * 32 + t) * 4 + 0] - t0, host[(1 * 32 + t) * 4 + 1] - t0, host[(1 * 32 + t) * 4 + 2] - t0);
          fprintf(stderr, "  epilogue wait %6lld accumulators complete %6lld stored %6lld\n", host[(2 * 32) * 4 + 0] - t0, host[(2 * 32) * 4 + 1] - t0, host[(2 * 32) * 4 + 2] - t0);
        }
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) { set_error("conv_wgrad_halo_kernel launch failed: %s", cudaGetErrorString(e)); rc = 1; }
        else count_launch();
      }
      if (rc == 0) {
        const int wblocks = (int)(((int64_t)taps * F * p.Nn + 31) / 32);
        const int bblocks = dbias ? (F + 31) / 32 : 0;
        mc::launch(conv_wgrad_reduce_kernel, wblocks + bblocks, 256, 0, ws, dw, h.ws_bias, dbias, splits, taps, F, p.Nn, C, KH,
                                                                         KW, Cp, merged, wblocks);
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) { set_error("conv_wgrad_reduce_kernel launch failed: %s", cudaGetErrorString(e)); rc = 1; }
        else count_launch();
      }
      return rc;
    }
  }
  // taps per CTA: as many accumulators as TMEM (512 columns, 32 kept for the bias) and a >= 2-stage
  // shared-memory ring allow, balanced over the tap groups
  int tpc = (512 - 32) / p.BNc;
  const int smem_cap = (int)((100 * 1024 - 5 * 4096) / ((p.BNc / 32) * 4096));
  if (tpc > smem_cap) tpc = smem_cap;
  if (tpc > taps) tpc = taps;
  if (tpc < 1) tpc = 1;
  const int groups = (taps + tpc - 1) / tpc;
  tpc = (taps + groups - 1) / groups;
  p.tpc = tpc;
  const int ftiles = (F + 127) / 128;
  int splits = (sm_count() + groups * ftiles - 1) / (groups * ftiles);   // ~1 CTA per SM (each owns ~100-200 KB smem)
  const int max_splits = (p.g.total_chunks + 7) / 8;                     // >= 8 chunks (256 pixels) per CTA
  if (splits > max_splits) splits = max_splits;
  if (splits < 1) splits = 1;
  p.chunks_per_split = (p.g.total_chunks + splits - 1) / splits;
  splits = (p.g.total_chunks + p.chunks_per_split - 1) / p.chunks_per_split;
  const size_t stage_bytes = 4 * 4096 + ((size_t)tpc * (p.BNc / 32) + 1) * 4096;
  int stages = (int)((200 * 1024) / stage_bytes);
  if (stages > 8) stages = 8;
  if (stages < 1) stages = 1;
  p.stages = stages;
  const size_t smem = stages * stage_bytes + (2 * stages + 1) * 8 + 16 + 1024;

  const size_t ws_elems = (size_t)splits * taps * F * p.Nn;
  ScratchFrame scratch;
  float* ws = scratch.floats(ws_elems + (dbias ? (size_t)splits * F : 0));
  if (!ws) return 1;
  p.ws = ws;
  p.ws_bias = dbias ? ws + ws_elems : nullptr;
  CUtensorMap tmDy, tmX;
  int rc = make_act_tmap(&tmDy, dy, N, OH, OW, Fp, Fp, OW, p.g.bw, p.g.bh, true);
  if (rc == 0) rc = make_act_tmap(&tmX, x, N, H, Wt, Cp, p.Nn, Wdim, p.g.bw, p.g.bh, true);
  static uint32_t configured = 0;
  if (rc == 0 && first_use_on_device(&configured)) rc = set_max_smem((const void*)conv_wgrad_kernel, "conv_wgrad_kernel");
  if (rc == 0) {
    dim3 grid((unsigned)groups, (unsigned)splits, (unsigned)ftiles);
    mc::launch_pdl(conv_wgrad_kernel, grid, CONV_THREADS, smem, tmDy, tmX, p);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { set_error("conv_wgrad_kernel launch failed: %s", cudaGetErrorString(e)); rc = 1; }
    else count_launch();
  }
  if (rc == 0) {
    const int wblocks = (int)(((int64_t)taps * F * p.Nn + 31) / 32);
    const int bblocks = dbias ? (F + 31) / 32 : 0;
    mc::launch(conv_wgrad_reduce_kernel, wblocks + bblocks, 256, 0, ws, dw, p.ws_bias, dbias, splits, taps, F, p.Nn, C, KH, KW,
                                                                     Cp, merged, wblocks);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { set_error("conv_wgrad_reduce_kernel launch failed: %s", cudaGetErrorString(e)); rc = 1; }
    else count_launch();
  }
  return rc;
}

int nchw_to_nhwc(const float* x, float* y, int N, int C, int H, int W, int Cp, int wpad) {
  if (Cp == 4 && aligned16(y)) {
    const int64_t total = (int64_t)N * H * (W + 2 * wpad);
    mc::launch(nchw_to_nhwc4_kernel, grid_for(total, 256), 256, 0, x, reinterpret_cast<float4*>(y), N, C, H, W, wpad);
    MC_LAUNCHED();
    return 0;
  }
  if (wpad == 0) { W *= H; H = 1; }            // no pad columns: (h, w) is one flat pixel axis, full 32-pixel tiles for small images
  dim3 grid((unsigned)((W + 2 * wpad + 31) / 32), (unsigned)((Cp + 31) / 32), (unsigned)(N * H));
  mc::launch(nchw_to_nhwc_kernel, grid, 256, 0, x, y, N, C, H, W, Cp, wpad);
  MC_LAUNCHED();
  return 0;
}

int nhwc_to_nchw(const float* x, float* y, int N, int C, int H, int W, int Cp) {
  W *= H; H = 1;                               // (h, w) as one flat pixel axis
  dim3 grid((unsigned)((W + 31) / 32), (unsigned)((C + 31) / 32), (unsigned)(N * H));
  mc::launch(nhwc_to_nchw_kernel, grid, 256, 0, x, y, N, C, H, W, Cp);
  MC_LAUNCHED();
  return 0;
}

}  // namespace tc
}  // namespace mc

using namespace mc;

namespace {

bool merged_mode(int64_t KW, int64_t Cp, int64_t wpad, int64_t pad) { return KW * Cp <= 32 && wpad == pad; }

int check_nhwc(const void* a, const void* b, const void* c, int64_t Cp, int64_t Fp, const char* who) {
  if ((Cp & 3) || (Fp & 3)) { set_error("%s: channel pitches must be multiples of 4 (got %lld, %lld)", who, (long long)Cp, (long long)Fp); return 2; }
  if (!aligned16(a) || !aligned16(b) || !aligned16(c)) { set_error("%s: tensors must be 16-byte aligned", who); return 2; }
  return 0;
}

}  // namespace

extern "C" {

int mc_nchw_to_nhwc(const float* x, float* y, int64_t N, int64_t C, int64_t H, int64_t W, int64_t Cp, int64_t wpad) {
  MC_REQUIRE(Cp >= C && wpad >= 0, "Cp must be >= C and wpad >= 0");
  if (N * H * W == 0) return 0;
  MC_REQUIRE(N * H <= 2147483647LL, "N*H exceeds the grid limit");
  return tc::nchw_to_nhwc(x, y, (int)N, (int)C, (int)H, (int)W, (int)Cp, (int)wpad);
}

int mc_nhwc_to_nchw(const float* x, float* y, int64_t N, int64_t C, int64_t H, int64_t W, int64_t Cp) {
  MC_REQUIRE(Cp >= C, "Cp must be >= C");
  if (N * H * W * C == 0) return 0;
  MC_REQUIRE(N * H <= 2147483647LL, "N*H exceeds the grid limit");
  return tc::nhwc_to_nchw(x, y, (int)N, (int)C, (int)H, (int)W, (int)Cp);
}

int mc_conv2d_nhwc_pool_supported(int64_t OH, int64_t OW, int64_t F, int64_t Fp, int* ok) {
  *ok = tc::conv_pool_ok((int)OH, (int)OW, (int)F, (int)Fp) ? 1 : 0;
  return 0;
}

int mc_conv2d_nhwc_fwd_pool(const float* x, const float* w, const float* bias, float* y_pooled, uint8_t* idx_u8, int64_t N,
                            int64_t C, int64_t H, int64_t W, int64_t Cp, int64_t wpad, int64_t F, int64_t Fp, int64_t KH,
                            int64_t KW, int64_t pad, int relu) {
  if (int rc = check_nhwc(x, y_pooled, y_pooled, Cp, Fp, "mc_conv2d_nhwc_fwd_pool")) return rc;
  const int64_t OH = H + 2 * pad - KH + 1, OW = W + 2 * pad - KW + 1;
  MC_REQUIRE(OH > 0 && OW > 0, "kernel larger than padded input");
  MC_REQUIRE(wpad == 0 || merged_mode(KW, Cp, wpad, pad), "a W-padded input needs wpad == pad and KW*Cp <= 32");
  MC_REQUIRE(tc::conv_pool_ok((int)OH, (int)OW, (int)F, (int)Fp), "shape outside the pooled epilogue (mc_conv2d_nhwc_pool_supported)");
  MC_REQUIRE(idx_u8 == nullptr || aligned16(idx_u8), "idx must be 16-byte aligned");
  if (N == 0) return 0;
  const int mode = merged_mode(KW, Cp, wpad, pad) ? 1 : 0;
  return tc::conv_igemm(x, w, bias, y_pooled, (int)N, (int)H, (int)(W + 2 * wpad), (int)Cp, (int)C, (int)F, (int)Fp, (int)F,
                        (int)C, (int)KH, (int)KW, (int)pad, (int)OH, (int)OW, relu, mode, 1, idx_u8);
}

int mc_conv2d_nhwc_fwd(const float* x, const float* w, const float* bias, float* y, int64_t N, int64_t C, int64_t H,
                       int64_t W, int64_t Cp, int64_t wpad, int64_t F, int64_t Fp, int64_t KH, int64_t KW, int64_t pad,
                       int relu) {
  if (int rc = check_nhwc(x, y, y, Cp, Fp, "mc_conv2d_nhwc_fwd")) return rc;
  const int64_t OH = H + 2 * pad - KH + 1, OW = W + 2 * pad - KW + 1;
  MC_REQUIRE(OH > 0 && OW > 0, "kernel larger than padded input");
  MC_REQUIRE(wpad == 0 || merged_mode(KW, Cp, wpad, pad), "a W-padded input needs wpad == pad and KW*Cp <= 32");
  if (N == 0) return 0;
  const int mode = merged_mode(KW, Cp, wpad, pad) ? 1 : 0;
  return tc::conv_igemm(x, w, bias, y, (int)N, (int)H, (int)(W + 2 * wpad), (int)Cp, (int)C, (int)F, (int)Fp, (int)F, (int)C,
                        (int)KH, (int)KW, (int)pad, (int)OH, (int)OW, relu, mode);
}

int mc_conv2d_nhwc_dgrad(const float* dy, const float* w, float* dx, int64_t N, int64_t C, int64_t H, int64_t W, int64_t Cp,
                         int64_t F, int64_t Fp, int64_t KH, int64_t KW, int64_t pad) {
  if (int rc = check_nhwc(dy, dx, dx, Cp, Fp, "mc_conv2d_nhwc_dgrad")) return rc;
  MC_REQUIRE(KH == KW && KH - 1 - pad >= 0, "dgrad needs a square filter with pad <= K-1");
  const int64_t OH = H + 2 * pad - KH + 1, OW = W + 2 * pad - KW + 1;
  MC_REQUIRE(OH > 0 && OW > 0, "kernel larger than padded input");
  if (N == 0) return 0;
  // dx[pixel, c] = sum_{tap', f} dy[pixel + tap' - (K-1-pad), f] * w[f, c, flipped tap']
  return tc::conv_igemm(dy, w, nullptr, dx, (int)N, (int)OH, (int)OW, (int)Fp, (int)F, (int)C, (int)Cp, (int)F, (int)C, (int)KH,
                        (int)KW, (int)(KH - 1 - pad), (int)H, (int)W, 0, 2);
}

int mc_conv2d_nhwc_wgrad(const float* x, const float* dy, float* dw, float* dbias, int64_t N, int64_t C, int64_t H, int64_t W,
                         int64_t Cp, int64_t wpad, int64_t F, int64_t Fp, int64_t KH, int64_t KW, int64_t pad) {
  if (int rc = check_nhwc(x, dy, dy, Cp, Fp, "mc_conv2d_nhwc_wgrad")) return rc;
  const int64_t OH = H + 2 * pad - KH + 1, OW = W + 2 * pad - KW + 1;
  MC_REQUIRE(OH > 0 && OW > 0, "kernel larger than padded input");
  MC_REQUIRE(wpad == 0 || merged_mode(KW, Cp, wpad, pad), "a W-padded input needs wpad == pad and KW*Cp <= 32");
  if (dw) {
    if (N == 0) { if (int rc = mc_fill(dw, 0.f, F * C * KH * KW)) return rc; }
    else {
      // the bias gradient rides along as one more accumulator of the same kernel
      const int merged = merged_mode(KW, Cp, wpad, pad) ? 1 : 0;
      return tc::conv_wgrad_tc(x, dy, dw, dbias, (int)N, (int)H, (int)(W + 2 * wpad), (int)Cp, (int)C, (int)F, (int)Fp, (int)KH,
                               (int)KW, (int)pad, (int)OH, (int)OW, merged);
    }
  }
  if (dbias) {
    // column sums of dy viewed as (pixels, Fp); only the first F entries are real channels
    if (Fp == F) return reduce_sum_internal(dy, dbias, 1, N * OH * OW, F);
    ScratchFrame scratch;
    float* tmp = scratch.floats(Fp);
    if (!tmp) return 1;
    int rc = reduce_sum_internal(dy, tmp, 1, N * OH * OW, Fp);
    if (rc == 0) rc = mc_memcpy_d2d(dbias, tmp, sizeof(float) * F);
    return rc;
  }
  return 0;
}

int mc_maxpool2d_nhwc_fwd(const float* x, float* y, uint8_t* idx_u8, int64_t N, int64_t H, int64_t W, int64_t Cp, int64_t KH,
                          int64_t KW, int64_t stride) {
  const int64_t OH = (H - KH) / stride + 1, OW = (W - KW) / stride + 1;
  MC_REQUIRE(OH > 0 && OW > 0 && KH * KW <= 255, "bad pooling window");
  const int64_t n = N * OH * OW * Cp;
  if (n <= 0) return 0;
  if ((Cp & 3) == 0 && aligned16(x) && aligned16(y) && (!idx_u8 || (reinterpret_cast<uintptr_t>(idx_u8) & 3u) == 0)) {
    mc::launch(tc::maxpool_nhwc_fwd_vec_kernel, grid_for(n / 4, 256), 256, 0, 
        reinterpret_cast<const float4*>(x), reinterpret_cast<float4*>(y), reinterpret_cast<uchar4*>(idx_u8), (int)N, (int)H,
        (int)W, (int)(Cp / 4), (int)KH, (int)KW, (int)stride, (int)OH, (int)OW);
  } else {
    mc::launch(tc::maxpool_nhwc_fwd_kernel, grid_for(n, 256), 256, 0, x, y, idx_u8, (int)N, (int)H, (int)W, (int)Cp, (int)KH,
                                                                       (int)KW, (int)stride, (int)OH, (int)OW);
  }
  MC_LAUNCHED();
  return 0;
}

int mc_maxpool2d_nhwc_bwd(const float* dy, const uint8_t* idx_u8, const float* relu_y, float* dx, int64_t N, int64_t H,
                          int64_t W, int64_t Cp, int64_t K, int64_t OH, int64_t OW) {
  const int64_t n = N * H * W * Cp;
  if (n <= 0) return 0;
  if ((Cp & 3) == 0 && H == OH * K && W == OW * K && aligned16(dy) && aligned16(dx) && (!relu_y || aligned16(relu_y)) &&
      (reinterpret_cast<uintptr_t>(idx_u8) & 3u) == 0) {
    mc::launch(tc::maxpool_nhwc_bwd_vec_kernel, grid_for(N * OH * OW * (Cp / 4), 256), 256, 0, 
        reinterpret_cast<const float4*>(dy), reinterpret_cast<const uchar4*>(idx_u8), reinterpret_cast<const float4*>(relu_y),
        reinterpret_cast<float4*>(dx), (int)N, (int)H, (int)W, (int)(Cp / 4), (int)K, (int)OH, (int)OW);
  } else {
    mc::launch(tc::maxpool_nhwc_bwd_kernel, grid_for(n, 256), 256, 0, dy, idx_u8, relu_y, dx, (int)N, (int)H, (int)W, (int)Cp,
                                                                       (int)K, (int)OH, (int)OW);
  }
  MC_LAUNCHED();
  return 0;
}

}  // extern "C"
