// Dense contractions: np.dot and its two VJPs (numpy_core.py:224-225).
//
// Two paths behind one entry point (mc_gemm):
//   * TF32 tensor-core path (sm_100a): tcgen05.mma kind::tf32 issued by one thread,
//     operands staged in shared memory by TMA (SWIZZLE_128B), FP32 accumulator in TMEM,
//     read back with tcgen05.ld for a fused bias / relu / accumulate epilogue.  The FP32
//     operands are consumed in place -- no conversion pass over HBM.  All four
//     transpose combinations are expressed through the UMMA "major" bits and the TMA
//     box shape (K-major vs MN-major operand tiles), never through a transposed copy.
//     Skinny problems are split along K across CTAs (workspace + deterministic reduce)
//     so that every SM streams operand bytes: these GEMMs sit below the B200 ridge
//      and are bound by how many SMs pull from HBM/L2.
//   * FP32 CUDA-core path: exact FP32 FMA accumulation; serves shapes TMA cannot
//     describe (row pitch not a multiple of 16 B, e.g. the 10-class output layer) and is
//     the FP32-accumulate cross-check the parity tests compare the TF32 path with.
#include "tc_common.cuh"

namespace mc {

// =====================================================================================
//                                FP32 CUDA-core GEMM
// =====================================================================================
constexpr int SG_BM = 64, SG_BN = 64, SG_BK = 16, SG_THREADS = 256;

template <bool TA, bool TB>
__global__ void __launch_bounds__(SG_THREADS) sgemm_kernel(int M, int N, int K, const float* __restrict__ A, int64_t lda,
                                                           const float* __restrict__ B, int64_t ldb,
                                                           float* __restrict__ C, int64_t ldc,
                                                           const float* __restrict__ bias, int relu, int accumulate,
                                                           int k_per_split, float* __restrict__ ws) {
  mc::pdl_sync();
  __shared__ float As[SG_BK][SG_BM + 4];
  __shared__ float Bs[SG_BK][SG_BN + 4];
  const int t = threadIdx.x;
  const int tx = t & 15, ty = t >> 4;
  const int m0 = blockIdx.y * SG_BM, n0 = blockIdx.x * SG_BN;
  const int kbeg = blockIdx.z * k_per_split;
  const int kend = min(K, kbeg + k_per_split);
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  for (int k0 = kbeg; k0 < kend; k0 += SG_BK) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int m, k;
      if (TA) { m = t & 63; k = (t >> 6) + 4 * j; }
      else { k = t & 15; m = (t >> 4) + 16 * j; }
      const int gm = m0 + m, gk = k0 + k;
      float v = 0.f;
      if (gm < M && gk < kend) v = TA ? A[(int64_t)gk * lda + gm] : A[(int64_t)gm * lda + gk];
      As[k][m] = v;
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int n, k;
      if (TB) { k = t & 15; n = (t >> 4) + 16 * j; }
      else { n = t & 63; k = (t >> 6) + 4 * j; }
      const int gn = n0 + n, gk = k0 + k;
      float v = 0.f;
      if (gn < N && gk < kend) v = TB ? B[(int64_t)gn * ldb + gk] : B[(int64_t)gk * ldb + gn];
      Bs[k][n] = v;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < SG_BK; ++k) {
      const float4 a = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
      const float4 b = *reinterpret_cast<const float4*>(&Bs[k][tx * 4]);
      const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int gm = m0 + ty * 4 + i;
    if (gm >= M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int gn = n0 + tx * 4 + j;
      if (gn >= N) continue;
      float v = acc[i][j];
      if (ws) {
        ws[((int64_t)blockIdx.z * M + gm) * N + gn] = v;
      } else {
        if (bias) v += bias[gn];
        if (accumulate) v += C[(int64_t)gm * ldc + gn];
        if (relu) v = relu_apply(v, relu);
        C[(int64_t)gm * ldc + gn] = v;
      }
    }
  }
}

// Deterministic split-K combine with the fused epilogue.
__global__ void __launch_bounds__(256) splitk_reduce_kernel(const float* __restrict__ ws, int splits, int64_t M, int64_t N,
                                                            float* __restrict__ C, int64_t ldc,
                                                            const float* __restrict__ bias, int relu, int accumulate) {
  mc::pdl_sync();
  const int64_t total = M * N;
  const int64_t stride = (int64_t)gridDim.x * 256;
  if ((N & 3) == 0 && (ldc & 3) == 0 && ((reinterpret_cast<uintptr_t>(C) & 15u) == 0)) {
    const int64_t total4 = total >> 2;
    for (int64_t i4 = (int64_t)blockIdx.x * 256 + threadIdx.x; i4 < total4; i4 += stride) {
      const int64_t i = i4 << 2;
      float4 s = *reinterpret_cast<const float4*>(ws + i);
      for (int z = 1; z < splits; ++z) {
        const float4 p = *reinterpret_cast<const float4*>(ws + (int64_t)z * total + i);
        s.x += p.x; s.y += p.y; s.z += p.z; s.w += p.w;
      }
      const int64_t m = i / N, n = i % N;
      if (bias) { s.x += bias[n]; s.y += bias[n + 1]; s.z += bias[n + 2]; s.w += bias[n + 3]; }
      float4* dst = reinterpret_cast<float4*>(C + m * ldc + n);
      if (accumulate) { const float4 c = *dst; s.x += c.x; s.y += c.y; s.z += c.z; s.w += c.w; }
      if (relu) {
        s.x = relu_apply(s.x, relu); s.y = relu_apply(s.y, relu);
        s.z = relu_apply(s.z, relu); s.w = relu_apply(s.w, relu);
      }
      *dst = s;
    }
  } else {
    for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < total; i += stride) {
      float s = ws[i];
      for (int z = 1; z < splits; ++z) s += ws[(int64_t)z * total + i];
      const int64_t m = i / N, n = i % N;
      if (bias) s += bias[n];
      if (accumulate) s += C[m * ldc + n];
      if (relu) s = relu_apply(s, relu);
      C[m * ldc + n] = s;
    }
  }
}

// =====================================================================================
//                 skinny contractions: the 10-class output layer and its two VJPs
// =====================================================================================
// logits = h . W (N = classes <= 16), dW = h^T . g (N <= 16, long K = batch), dh = g . W^T (K <= 16).
// TMA cannot describe a 40-byte row pitch and a 64x64 SIMT tile wastes 54 of 64 columns plus a
// split-K pass (10 + 10 us per call measured); these three are bandwidth-trivial, so each gets the
// access pattern that makes its one big operand a coalesced stream.  FP32 FMA accumulation.
constexpr int SK_MAXN = 16;

__device__ __forceinline__ float sk_epilogue(float v, const float* bias, int n, const float* cptr, int relu, int accumulate) {
  if (bias) v += bias[n];
  if (accumulate) v += *cptr;
  if (relu) v = relu_apply(v, relu);
  return v;
}

// C(M,N) = A(M,K) . B(K,N), N <= 16: a block of 8 warps = (8 / S) rows x S slices of K; lanes stride K
// inside a slice (A row segments are coalesced streams), 4 independent loads in flight per lane.
template <int S>
__global__ void __launch_bounds__(256) skinny_nn_kernel(int M, int N, int K, const float* __restrict__ A, int64_t lda,
                                                        const float* __restrict__ B, int64_t ldb, float* __restrict__ C,
                                                        int64_t ldc, const float* __restrict__ bias, int relu, int accumulate) {
  mc::pdl_sync();
  constexpr int ROWS = 8 / S;
  __shared__ float red[8][SK_MAXN];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int slice = warp % S, rloc = warp / S;
  const int m = blockIdx.x * ROWS + rloc;
  float acc[SK_MAXN];
#pragma unroll
  for (int n = 0; n < SK_MAXN; ++n) acc[n] = 0.f;
  if (m < M) {
    const float* __restrict__ a = A + (int64_t)m * lda;
    int k = slice * 32 + lane;
    for (; k + 3 * 32 * S < K; k += 4 * 32 * S) {
      const float a0 = a[k], a1 = a[k + 32 * S], a2 = a[k + 64 * S], a3 = a[k + 96 * S];
      const float* __restrict__ b0 = B + (int64_t)k * ldb;
      const float* __restrict__ b1 = b0 + (int64_t)32 * S * ldb;
      const float* __restrict__ b2 = b1 + (int64_t)32 * S * ldb;
      const float* __restrict__ b3 = b2 + (int64_t)32 * S * ldb;
#pragma unroll
      for (int n = 0; n < SK_MAXN; ++n)
        if (n < N) {
          acc[n] = fmaf(a0, __ldg(b0 + n), acc[n]); acc[n] = fmaf(a1, __ldg(b1 + n), acc[n]);
          acc[n] = fmaf(a2, __ldg(b2 + n), acc[n]); acc[n] = fmaf(a3, __ldg(b3 + n), acc[n]);
        }
    }
    for (; k < K; k += 32 * S) {
      const float av = a[k];
      const float* __restrict__ b = B + (int64_t)k * ldb;
#pragma unroll
      for (int n = 0; n < SK_MAXN; ++n) if (n < N) acc[n] = fmaf(av, __ldg(b + n), acc[n]);
    }
  }
#pragma unroll
  for (int n = 0; n < SK_MAXN; ++n) {
    if (n < N) {
      float v = acc[n];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if (lane == 0) red[warp][n] = v;
    }
  }
  __syncthreads();
  if (slice == 0 && lane < N && m < M) {
    float v = 0.f;
#pragma unroll
    for (int q = 0; q < S; ++q) v += red[rloc * S + q][lane];
    float* cptr = C + (int64_t)m * ldc + lane;
    *cptr = sk_epilogue(v, bias, lane, cptr, relu, accumulate);
  }
}

// C(M,N) = A^T . B with A stored (K,M), B (K,N), N <= 16.  A block owns 8 columns m (one 32-byte
// sector of every A row) and splits K 64 ways (4 lane groups x 16 warps), four rows in flight per
// thread: M = 512 gives 64 blocks with 4 dependent load rounds each (the first version -- 32 columns
// per block, 16 K slices -- ran 16 blocks x 16 rounds of DRAM latency: 14 us for 2 MB).  The 64 partial
// sums are combined in a fixed order (lane groups by shuffle, warps through shared memory).
constexpr int SK_TN_WARPS = 16;
__global__ void __launch_bounds__(32 * SK_TN_WARPS) skinny_tn_kernel(int M, int N, int K, const float* __restrict__ A,
                                                                     int64_t lda, const float* __restrict__ B, int64_t ldb,
                                                                     float* __restrict__ C, int64_t ldc,
                                                                     const float* __restrict__ bias, int relu, int accumulate) {
  mc::pdl_sync();
  __shared__ float red[SK_TN_WARPS][SK_MAXN][8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int mi = lane & 7, ks = lane >> 3;
  const int m = blockIdx.x * 8 + mi;
  constexpr int ST = 4 * SK_TN_WARPS;               // K slices
  float acc[SK_MAXN];
#pragma unroll
  for (int n = 0; n < SK_MAXN; ++n) acc[n] = 0.f;
  if (m < M) {
    const float* __restrict__ a = A + m;
    int k = warp * 4 + ks;
    for (; k + 3 * ST < K; k += 4 * ST) {
      const float a0 = a[(int64_t)k * lda], a1 = a[(int64_t)(k + ST) * lda];
      const float a2 = a[(int64_t)(k + 2 * ST) * lda], a3 = a[(int64_t)(k + 3 * ST) * lda];
      const float* __restrict__ b0 = B + (int64_t)k * ldb;
      const float* __restrict__ b1 = b0 + (int64_t)ST * ldb;
      const float* __restrict__ b2 = b1 + (int64_t)ST * ldb;
      const float* __restrict__ b3 = b2 + (int64_t)ST * ldb;
#pragma unroll
      for (int n = 0; n < SK_MAXN; ++n)
        if (n < N) {
          acc[n] = fmaf(a0, __ldg(b0 + n), acc[n]); acc[n] = fmaf(a1, __ldg(b1 + n), acc[n]);
          acc[n] = fmaf(a2, __ldg(b2 + n), acc[n]); acc[n] = fmaf(a3, __ldg(b3 + n), acc[n]);
        }
    }
    for (; k < K; k += ST) {
      const float av = a[(int64_t)k * lda];
      const float* __restrict__ b = B + (int64_t)k * ldb;
#pragma unroll
      for (int n = 0; n < SK_MAXN; ++n) if (n < N) acc[n] = fmaf(av, __ldg(b + n), acc[n]);
    }
  }
#pragma unroll
  for (int n = 0; n < SK_MAXN; ++n) {
    acc[n] += __shfl_xor_sync(0xffffffffu, acc[n], 8);
    acc[n] += __shfl_xor_sync(0xffffffffu, acc[n], 16);
    if (ks == 0) red[warp][n][mi] = acc[n];
  }
  __syncthreads();
  // thread t finishes output (m = t % 8, n = t / 8)
  const int on = threadIdx.x >> 3, om = blockIdx.x * 8 + (threadIdx.x & 7);
  if (on < N && om < M) {
    float v = 0.f;
#pragma unroll
    for (int q = 0; q < SK_TN_WARPS; ++q) v += red[q][on][threadIdx.x & 7];
    float* cptr = C + (int64_t)om * ldc + on;
    *cptr = sk_epilogue(v, bias, on, cptr, relu, accumulate);
  }
}

// C(M,N) = A(M,K) . B^T with B stored (N,K), K <= 16: one thread per output, the output row is the
// coalesced stream; A row and B rows are a few cached floats.
__global__ void __launch_bounds__(256) skinny_nt_kernel(int M, int N, int K, const float* __restrict__ A, int64_t lda,
                                                        const float* __restrict__ B, int64_t ldb, float* __restrict__ C,
                                                        int64_t ldc, const float* __restrict__ bias, int relu, int accumulate) {
  mc::pdl_sync();
  const int64_t total = (int64_t)M * N;
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < total; i += (int64_t)gridDim.x * 256) {
    const int m = (int)(i / N), n = (int)(i - (int64_t)m * N);
    const float* __restrict__ a = A + (int64_t)m * lda;
    const float* __restrict__ b = B + (int64_t)n * ldb;
    float v = 0.f;
    for (int k = 0; k < K; ++k) v = fmaf(__ldg(a + k), __ldg(b + k), v);
    float* cptr = C + (int64_t)m * ldc + n;
    *cptr = sk_epilogue(v, bias, n, cptr, relu, accumulate);
  }
}

// returns true when a skinny kernel took the call
static bool gemm_skinny(int ta, int tb, int64_t M, int64_t N, int64_t K, const float* A, int64_t lda, const float* B,
                        int64_t ldb, float* C, int64_t ldc, const float* bias, int relu, int accumulate) {
  if (N <= SK_MAXN && !ta && !tb && K >= 64) {
    if (M >= 2048)
      mc::launch(skinny_nn_kernel<1>, (unsigned)((M + 7) / 8), 256, 0, (int)M, (int)N, (int)K, A, lda, B, ldb, C, ldc, bias,
                                                                       relu, accumulate);
    else if (M >= 512 && K < 1024)
      mc::launch(skinny_nn_kernel<2>, (unsigned)((M + 3) / 4), 256, 0, (int)M, (int)N, (int)K, A, lda, B, ldb, C, ldc, bias,
                                                                       relu, accumulate);
    else
      mc::launch(skinny_nn_kernel<8>, (unsigned)M, 256, 0, (int)M, (int)N, (int)K, A, lda, B, ldb, C, ldc, bias, relu,
                                                             accumulate);
  } else if (N <= SK_MAXN && ta && !tb && K >= 64) {
    mc::launch(skinny_tn_kernel, (unsigned)((M + 7) / 8), 32 * SK_TN_WARPS, 0, (int)M, (int)N, (int)K, A, lda, B, ldb, C, ldc, bias, relu,
                                                                    accumulate);
  } else if (K <= SK_MAXN && !ta && tb && M * N >= 4096) {
    mc::launch(skinny_nt_kernel, grid_for(M * N, 256), 256, 0, (int)M, (int)N, (int)K, A, lda, B, ldb, C, ldc, bias, relu,
                                                               accumulate);
  } else {
    return false;
  }
  return true;
}

// =====================================================================================
//                         TF32 tcgen05 / TMEM / TMA GEMM (sm_100a)
// =====================================================================================
namespace tc {

constexpr int BM = 128;         // UMMA M (cta_group::1): accumulator row i <-> TMEM lane i
constexpr int BK = 32;          // fp32 per stage along K = 128 B = one SWIZZLE_128B row
constexpr int THREADS = 256;    // warp 0: TMA producer, warp 1: MMA issuer; all 8: epilogue (two per TMEM lane quadrant)

// The LSTM time step as the epilogue of its recurrent GEMM (cluster split-K path only; see lstm.cu).
//   mode 1 (forward):  the GEMM is h_{t-1} . Wh' with Wh' = Wh's columns reordered unit-major (column 4u+g =
//          gate g of hidden unit u), so every float4 of the reduced tile holds the four recurrent
//          pre-activations of ONE unit; the epilogue adds the input projection ax[row, g*H + u], applies the
//          gate non-linearities and writes h_t, c_t and the gates (all gate-major, as the un-fused kernels).
//   mode 2 (backward): the GEMM is dA_t . Wh^T = the recurrent part of dh_{t-1}; every float4 holds it for four
//          consecutive units; the epilogue adds dH[t], and runs the gate backward of step t-1: dA_{t-1}, dc_{t-1}.
struct LstmEpilogue {
  const float* ax;        // fwd: (N, 4H) input projection of this step
  const float* c_prev;    // fwd: c_{t-1};            bwd: c_{t-1} of the step being differentiated
  const float* c_next;    //                           bwd: c_t of that step
  float* h_out;           // fwd: h_t
  float* c_out;           // fwd: c_t
  float* gates;           // fwd: out (N, 4H);         bwd: in, the gates of that step
  const float* dh_in;     // bwd: upstream dH of that step's h
  const float* dc_next;   // bwd: dc flowing in from the step after it (may be null)
  float* da;              // bwd: out (N, 4H)
  float* dc_out;          // bwd: out (N, H)
  int H;
};

struct GemmParams {
  float* C;
  int64_t ldc;
  const float* bias;
  float* ws;           // split-K workspace [splits][M][N] or nullptr
  int* counters;       // split-K: arrival count per output tile (self-resetting) or nullptr = separate reduce kernel
  int M, N, K;
  int relu, accumulate;
  int a_mn, b_mn;      // operand tile is MN-major (1) or K-major (0)
  int kb_per_split;    // k-blocks (of BK) per grid.z slice
  int cluster;         // > 1: the grid.z slices of a tile are one thread-block cluster; partials meet in DSMEM
  int group_m;         // tile raster: CTAs walk `group_m` M-tiles for each N-tile before moving on (0 = x-major)
  int half_from;       // persistent kernel: work items >= half_from are HALF-width tiles (two per tile), -1 = none
  int lstm_mode;       // cluster epilogue: 0 plain, 1 LSTM gate forward, 2 LSTM gate backward (lstm.cu)
  LstmEpilogue lstm;
};

// CTA q of an S-CTA cluster sums rows [q*BM/S, (q+1)*BM/S) of the parked partial tiles of all S peers
// (distributed shared memory), slice order 0..S-1, and applies the epilogue.  All S loads of an output
// chunk are issued before the first add: DSMEM latency is paid once per chunk, not S times.
template <int BN, int S>
__device__ __forceinline__ void cluster_reduce(const GemmParams& p, uint32_t smem_base, int m0, int n0, bool vec_ok) {
  constexpr int rows_per = BM / S;
  constexpr int F4 = BN / 4;
  constexpr int ITEMS = rows_per * F4;
  const int q = blockIdx.z;                           // cluster dims (1,1,S), gridDim.z == S: rank in cluster
  uint32_t peer[S];
#pragma unroll
  for (int z = 0; z < S; ++z) asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(peer[z]) : "r"(smem_base), "r"(z));
#pragma unroll (S >= 8 ? 1 : 2)
  for (int i = threadIdx.x; i < ITEMS; i += THREADS) {
    const int r = q * rows_per + i / F4, j = i % F4;
    const uint32_t off = (uint32_t)r * (BN * 4u) + (uint32_t)(j ^ (r & 7)) * 16u;
    float4 t[S];
#pragma unroll
    for (int z = 0; z < S; ++z)
      asm volatile("ld.shared::cluster.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(t[z].x), "=f"(t[z].y), "=f"(t[z].z), "=f"(t[z].w) : "r"(peer[z] + off) : "memory");
    float4 o = t[0];
#pragma unroll
    for (int z = 1; z < S; ++z) { o.x += t[z].x; o.y += t[z].y; o.z += t[z].z; o.w += t[z].w; }
    const int grow = m0 + r, col = n0 + 4 * j;
    if (p.lstm_mode == 1) {
      if (grow < p.M && col < p.N) {
        const LstmEpilogue& L = p.lstm;
        const int H = L.H, u = col >> 2;
        const int64_t gbase = (int64_t)grow * 4 * H + u, cell = (int64_t)grow * H + u;
        const float p0 = o.x + L.ax[gbase], p1 = o.y + L.ax[gbase + H], p2 = o.z + L.ax[gbase + 2 * H],
                    p3 = o.w + L.ax[gbase + 3 * H];
        const float gi = 1.0f / (1.0f + expf(-p0)), gf = 1.0f / (1.0f + expf(-p1)), go = 1.0f / (1.0f + expf(-p2)),
                    gg = tanhf(p3);
        const float c = gf * L.c_prev[cell] + gi * gg;
        L.c_out[cell] = c;
        L.h_out[cell] = go * tanhf(c);
        L.gates[gbase] = gi; L.gates[gbase + H] = gf; L.gates[gbase + 2 * H] = go; L.gates[gbase + 3 * H] = gg;
      }
      continue;
    }
    if (p.lstm_mode == 2) {
      if (grow < p.M && col < p.N) {
        const LstmEpilogue& L = p.lstm;
        const int H = L.H;
        const int64_t gbase = (int64_t)grow * 4 * H + col, cell = (int64_t)grow * H + col;
        const float4 gi = *reinterpret_cast<const float4*>(L.gates + gbase), gf = *reinterpret_cast<const float4*>(L.gates + gbase + H),
                     go = *reinterpret_cast<const float4*>(L.gates + gbase + 2 * H),
                     gg = *reinterpret_cast<const float4*>(L.gates + gbase + 3 * H);
        const float4 cn = *reinterpret_cast<const float4*>(L.c_next + cell), cp = *reinterpret_cast<const float4*>(L.c_prev + cell);
        float4 dh = *reinterpret_cast<const float4*>(L.dh_in + cell);
        dh.x += o.x; dh.y += o.y; dh.z += o.z; dh.w += o.w;
        float4 dcn = make_float4(0.f, 0.f, 0.f, 0.f);
        if (L.dc_next) dcn = *reinterpret_cast<const float4*>(L.dc_next + cell);
        float4 d0, d1, d2, d3, dpc;
#define MC_LSTM_BWD(m)                                                \
        {                                                             \
          const float tc = tanhf(cn.m);                               \
          float dc = dh.m * go.m * (1.f - tc * tc);                   \
          if (L.dc_next) dc += dcn.m;                                 \
          d0.m = dc * gg.m * gi.m * (1.f - gi.m);                     \
          d1.m = dc * cp.m * gf.m * (1.f - gf.m);                     \
          d2.m = dh.m * tc * go.m * (1.f - go.m);                     \
          d3.m = dc * gi.m * (1.f - gg.m * gg.m);                     \
          dpc.m = dc * gf.m;                                          \
        }
        MC_LSTM_BWD(x) MC_LSTM_BWD(y) MC_LSTM_BWD(z) MC_LSTM_BWD(w)
#undef MC_LSTM_BWD
        *reinterpret_cast<float4*>(L.da + gbase) = d0;
        *reinterpret_cast<float4*>(L.da + gbase + H) = d1;
        *reinterpret_cast<float4*>(L.da + gbase + 2 * H) = d2;
        *reinterpret_cast<float4*>(L.da + gbase + 3 * H) = d3;
        *reinterpret_cast<float4*>(L.dc_out + cell) = dpc;
      }
      continue;
    }
    if (grow < p.M && col < p.N) {
      float* dst = p.C + (int64_t)grow * p.ldc + col;
      if (vec_ok && col + 4 <= p.N) {
        if (p.bias) {
          const float4 b = __ldg(reinterpret_cast<const float4*>(p.bias + col));
          o.x += b.x; o.y += b.y; o.z += b.z; o.w += b.w;
        }
        if (p.accumulate) {
          const float4 c = *reinterpret_cast<const float4*>(dst);
          o.x += c.x; o.y += c.y; o.z += c.z; o.w += c.w;
        }
        if (p.relu) {
          o.x = relu_apply(o.x, p.relu); o.y = relu_apply(o.y, p.relu);
          o.z = relu_apply(o.z, p.relu); o.w = relu_apply(o.w, p.relu);
        }
        *reinterpret_cast<float4*>(dst) = o;
      } else {
        const float e[4] = {o.x, o.y, o.z, o.w};
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          if (col + u < p.N) {
            float w = e[u];
            if (p.bias) w += __ldg(p.bias + col + u);
            if (p.accumulate) w += dst[u];
            if (p.relu) w = relu_apply(w, p.relu);
            dst[u] = w;
          }
        }
      }
    }
  }
}

// Stage layout (each operand tile is a stack of SWIZZLE_128B atoms: 8 rows x 128 B):
//   K-major  tile (R rows of MN, BK of K): R rows x 128 B; 8-row groups 1024 B apart (SBO).
//            MMA k (K = 8k..8k+7) starts 32k bytes into the swizzled row.
//   MN-major tile: R/32 blocks of [BK k-rows x 128 B] (4096 B each, LBO), written by TMA with
//            SWIZZLE_128B_ATOM_32B (pattern repeats every 4 k-rows = 512 B, SBO); one MMA
//            (K = 8) covers two such groups: start + 1024k.
MC_PROBE_DEFINE(gemm)
template <int BN, int STAGES>
__global__ void __launch_bounds__(THREADS, 1) gemm_tf32_kernel(const __grid_constant__ CUtensorMap tmA,
                                                               const __grid_constant__ CUtensorMap tmB,
                                                               const GemmParams p) {
  constexpr uint32_t A_BYTES = BM * BK * 4;
  constexpr uint32_t B_BYTES = BN * BK * 4;
  constexpr uint32_t STAGE_BYTES = A_BYTES + B_BYTES;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bar_base = smem_base + STAGES * STAGE_BYTES;  // full[STAGES], empty[STAGES], tmem_full
  const uint32_t tmem_slot = bar_base + (2 * STAGES + 1) * 8;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (STAGES + s); };
  const uint32_t tmem_full_bar = bar_base + 8u * (2 * STAGES);

  MC_STAMP(gemm, 0);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // Tile raster.  CTAs are scheduled in linear block order; with the plain x-major order one wave of 148
  // CTAs of a 4096^3 problem touches ~5 A panels and ALL 32 B panels, so B (64 MB) is streamed from HBM
  // once per wave: 520 MB of DRAM reads for 134 MB of operands (profiles/r1z/prof_gemm_4096.txt).  Walking
  // `group_m` M-tiles per N-tile keeps the group's A panels L2-resident while each B panel is fetched once
  // per group.
  int tile_m = blockIdx.y, tile_n = blockIdx.x;
  if (p.group_m > 0) {
    const int tiles_m = gridDim.y, tiles_n = gridDim.x;
    const int lin = blockIdx.y * tiles_n + blockIdx.x;
    const int per_group = p.group_m * tiles_n;
    const int first_m = (lin / per_group) * p.group_m;
    const int gsize = min(tiles_m - first_m, p.group_m);
    const int r = lin % per_group;
    tile_m = first_m + r % gsize;
    tile_n = r / gsize;
  }
  const int m0 = tile_m * BM, n0 = tile_n * BN;
  const int num_kb_total = (p.K + BK - 1) / BK;
  const int kb_begin = blockIdx.z * p.kb_per_split;
  const int kb_end = min(num_kb_total, kb_begin + p.kb_per_split);
  const int num_kb = kb_end - kb_begin;

  if (warp == 0 && lane == 0) {
    // descriptor fetch overlaps the previous kernel's tail (the maps are kernel parameters: legal before pdl_sync)
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmA)) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmB)) : "memory");
    for (int s = 0; s < STAGES; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
    mbar_init(tmem_full_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"((uint32_t)BN)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  MC_STAMP(gemm, 1);
  mc::pdl_sync();      // prologue done (on-chip state only); global memory from here on
  MC_STAMP(gemm, 2);
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot) : "memory");

  if (warp == 0 && elect_one()) {
    // ------------------------------ TMA producer ------------------------------------
    // (one thread: stage / phase / coordinates are carried, not recomputed -- see conv_tc.cu)
    const bool a_mn = p.a_mn != 0, b_mn = p.b_mn != 0;
    int s = 0, k0 = kb_begin * BK;
    uint32_t ph = 0;
    for (int i = 0; i < num_kb; ++i, k0 += BK) {
      mbar_wait(empty_bar(s), ph ^ 1u);
      const uint32_t fb = full_bar(s);
      mbar_expect_tx(fb, STAGE_BYTES);
      const uint32_t sa = smem_base + s * STAGE_BYTES, sb = sa + A_BYTES;
      if (!a_mn) {
        tma_load_2d(sa, &tmA, fb, k0, m0);
      } else {
#pragma unroll
        for (int j = 0; j < BM / 32; ++j) tma_load_2d(sa + j * 4096, &tmA, fb, m0 + 32 * j, k0);
      }
      if (!b_mn) {
        tma_load_2d(sb, &tmB, fb, k0, n0);
      } else {
#pragma unroll
        for (int j = 0; j < BN / 32; ++j) tma_load_2d(sb + j * 4096, &tmB, fb, n0 + 32 * j, k0);
      }
      if (++s == STAGES) { s = 0; ph ^= 1u; }
    }
    MC_STAMP(gemm, 3);
  } else if (warp == 1 && elect_one()) {
    // ------------------------------ MMA issuer --------------------------------------
    // Instruction descriptor (cute UMMA::InstrDescriptor): c_format F32 (1<<4), a/b TF32
    // (2<<7, 2<<10), a_major bit 15, b_major bit 16, N>>3 at [17,23), M>>4 at [24,29).
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)p.a_mn << 15) | ((uint32_t)p.b_mn << 16) |
                           ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
    // shared-memory descriptor = layout bits | (address >> 4): one K step (8 tf32) is +32 B inside a
    // K-major swizzled row (+2) or +1024 B = 8 k-rows of an MN-major block (+64)
    const uint64_t adesc0 = p.a_mn ? make_desc(0, 4096, 512, 1) : make_desc(0, 16, 1024, 2);
    const uint64_t bdesc0 = p.b_mn ? make_desc(0, 4096, 512, 1) : make_desc(0, 16, 1024, 2);
    const uint32_t astep = p.a_mn ? 64u : 2u, bstep = p.b_mn ? 64u : 2u;
    int s = 0;
    uint32_t ph = 0;
    for (int i = 0; i < num_kb; ++i) {
      mbar_wait(full_bar(s), ph);
      tcgen05_fence_after();
      const uint32_t sa = smem_base + s * STAGE_BYTES, sb = sa + A_BYTES;
      const uint64_t adesc = adesc0 + (uint64_t)((sa & 0x3FFFFu) >> 4);
      const uint64_t bdesc = bdesc0 + (uint64_t)((sb & 0x3FFFFu) >> 4);
#pragma unroll
      for (int k = 0; k < BK / UMMA_K; ++k)
        umma_tf32(tmem_base, adesc + (uint64_t)(astep * k), bdesc + (uint64_t)(bstep * k), idesc, (i > 0 || k > 0) ? 1u : 0u);
      umma_commit(empty_bar(s));  // frees the smem stage once these MMAs have read it
      if (++s == STAGES) { s = 0; ph ^= 1u; }
    }
    umma_commit(tmem_full_bar);   // accumulator complete
  }
  __syncwarp();

  // ---------------------------------- epilogue ----------------------------------------
  // Eight warps: warp w drains TMEM lanes 32 (w & 3) .. +31 (the quadrant a warp may read), columns
  // [(w >> 2) BN / 2, +BN / 2).  With four warps -- one per scheduler -- every dependent-issue stall of the
  // epilogue's instruction stream was exposed: 1.0 us per 32-column pass, 2.5 us of a 7.2 us (1024,512,512) GEMM
  // (tools/probe.py gemm, profiles/r2l), whatever the store pattern.
  const int quad = warp & 3, chalf = warp >> 2;
  const bool vec_ok = ((p.ldc & 3) == 0) && ((reinterpret_cast<uintptr_t>(p.C) & 15u) == 0);
  const int j4 = (lane & 7) * 4, rsub = lane >> 3;
  // interior tile, aligned output: no bounds checks, no per-element branches
  const bool ws_mode = p.ws != nullptr;
  const bool fast = (m0 + BM <= p.M) && (n0 + BN <= p.N) && p.cluster <= 1 &&
                    (ws_mode ? ((p.N & 3) == 0) : (vec_ok && (!p.bias || (reinterpret_cast<uintptr_t>(p.bias) & 15u) == 0)));
  float4 bq[BN / 64];                                   // this lane's bias columns, fetched under the main loop
#pragma unroll
  for (int c = 0; c < BN / 64; ++c) {
    bq[c] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (fast && !ws_mode && p.bias) bq[c] = __ldg(reinterpret_cast<const float4*>(p.bias + n0 + chalf * (BN / 2) + c * 32 + j4));
  }
  mbar_wait_warp(tmem_full_bar, 0);
  tcgen05_fence_after();
  __syncwarp();
  MC_STAMP(gemm, 4);
  if (p.cluster > 1) {
    // Split-K inside a thread-block cluster: the S = gridDim.z CTAs of one output tile each hold a
    // partial accumulator.  Every CTA parks its partial in its own shared memory (the operand ring is
    // idle now), the cluster synchronises, and CTA q sums rows [q*BM/S, (q+1)*BM/S) over the S peers
    // through distributed shared memory, in slice order 0..S-1 (deterministic), and applies the
    // epilogue.  No workspace in HBM, no second kernel (splitk_reduce_kernel + its launch gap were
    // 12% of the C3 step).  Parked layout: row r at r*BN*4, its 16-byte chunk j at slot j ^ (r & 7).
    const int r_local = quad * 32 + lane;
#pragma unroll 1
    for (int c0 = chalf * (BN / 2); c0 < (chalf + 1) * (BN / 2); c0 += 32) {
      float v[32];
      tmem_ld32(tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)c0, v);
      const uint32_t rbase = smem_base + (uint32_t)r_local * (BN * 4u);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const uint32_t slot = (uint32_t)((c0 >> 2) + j) ^ (uint32_t)(r_local & 7);
        asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(rbase + slot * 16u), "f"(v[4 * j]), "f"(v[4 * j + 1]),
                     "f"(v[4 * j + 2]), "f"(v[4 * j + 3]) : "memory");
      }
    }
    tcgen05_fence_before();
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
    const bool vec_c = vec_ok;
    switch (p.cluster) {
      case 2: cluster_reduce<BN, 2>(p, smem_base, m0, n0, vec_c); break;
      case 4: cluster_reduce<BN, 4>(p, smem_base, m0, n0, vec_c); break;
      case 8: cluster_reduce<BN, 8>(p, smem_base, m0, n0, vec_c); break;
      default: cluster_reduce<BN, 16>(p, smem_base, m0, n0, vec_c); break;
    }
    // peers may still be reading this CTA's shared memory
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
    if (warp == 1) {
      asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)BN) : "memory");
    }
    return;
  }
  // A thread reads one ROW of the accumulator (32 columns a pass).  The warp's 32 x 32 block is turned through
  // shared memory (the operand ring is idle once tmem_full has fired; 4 KB a warp, 16-byte chunks XOR-swizzled by
  // row) so that eight lanes write one 128-byte row segment and a warp store covers four whole lines.
  const uint32_t stage = smem_base + (uint32_t)warp * 4096u;
  const uint32_t st_addr = stage + (uint32_t)lane * 128u, st_x = (uint32_t)(lane & 7);
  const int row0 = m0 + quad * 32;
  if (fast) {
    float* base;
    int64_t ld;
    if (ws_mode) { base = p.ws + ((int64_t)blockIdx.z * p.M + row0 + rsub) * p.N + n0 + j4; ld = p.N; }
    else { base = p.C + (int64_t)(row0 + rsub) * p.ldc + n0 + j4; ld = p.ldc; }
    const int mode = ws_mode ? 0 : (p.accumulate ? 2 : (p.relu ? 1 : 0));
    const int relu = ws_mode ? 0 : p.relu;
    const float negv = (relu == 2) ? -0.0f : 0.f;
#pragma unroll
    for (int c = 0; c < BN / 64; ++c) {
      const int c0 = chalf * (BN / 2) + c * 32;
      float v[32];
      tmem_ld32(tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)c0, v);
      if (num_kb <= 0) continue;
#pragma unroll
      for (int j = 0; j < 8; ++j)
        asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(st_addr + (((uint32_t)j ^ st_x) * 16u)), "f"(v[4 * j]),
                     "f"(v[4 * j + 1]), "f"(v[4 * j + 2]), "f"(v[4 * j + 3]) : "memory");
      __syncwarp();
      const float4 b4 = bq[c];
      float* d = base + c0;
      float4 o[8];
#pragma unroll
      for (int it = 0; it < 8; ++it) {
        const int rr = it * 4 + rsub;
        asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(o[it].x), "=f"(o[it].y), "=f"(o[it].z), "=f"(o[it].w)
                     : "r"(stage + (uint32_t)rr * 128u + (uint32_t)(((lane & 7) ^ (rr & 7)) * 16)) : "memory");
      }
      float4 cv[8];
      if (mode == 2) {
#pragma unroll
        for (int it = 0; it < 8; ++it) cv[it] = *reinterpret_cast<const float4*>(d + (int64_t)(it * 4) * ld);
      }
#pragma unroll
      for (int it = 0; it < 8; ++it) {
        float4 t = make_float4(o[it].x + b4.x, o[it].y + b4.y, o[it].z + b4.z, o[it].w + b4.w);
        if (mode == 2) { t.x += cv[it].x; t.y += cv[it].y; t.z += cv[it].z; t.w += cv[it].w; }
        if (relu) {                                     // relu_apply without branches: NaN passes, -x -> negv, +-0 -> +0
          t.x = (t.x < 0.f) ? negv : ((t.x == 0.f) ? 0.f : t.x); t.y = (t.y < 0.f) ? negv : ((t.y == 0.f) ? 0.f : t.y);
          t.z = (t.z < 0.f) ? negv : ((t.z == 0.f) ? 0.f : t.z); t.w = (t.w < 0.f) ? negv : ((t.w == 0.f) ? 0.f : t.w);
        }
        *reinterpret_cast<float4*>(d + (int64_t)(it * 4) * ld) = t;
      }
      __syncwarp();                                     // the block is reused by the next pass
    }
  } else {
    const bool ws_vec = (p.N & 3) == 0;
#pragma unroll 1
    for (int c0 = chalf * (BN / 2); c0 < (chalf + 1) * (BN / 2); c0 += 32) {
      float v[32];
      tmem_ld32(tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)c0, v);
      if (num_kb <= 0) continue;
#pragma unroll
      for (int j = 0; j < 8; ++j)
        asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(st_addr + (((uint32_t)j ^ st_x) * 16u)), "f"(v[4 * j]),
                     "f"(v[4 * j + 1]), "f"(v[4 * j + 2]), "f"(v[4 * j + 3]) : "memory");
      __syncwarp();
      const int col = n0 + c0 + j4;                     // this lane's four columns, the same in every row
      const bool full4 = col + 4 <= p.N;
      float4 b4 = make_float4(0.f, 0.f, 0.f, 0.f);
      if (!p.ws && p.bias) {
        if (col < p.N) b4.x = __ldg(p.bias + col);
        if (col + 1 < p.N) b4.y = __ldg(p.bias + col + 1);
        if (col + 2 < p.N) b4.z = __ldg(p.bias + col + 2);
        if (col + 3 < p.N) b4.w = __ldg(p.bias + col + 3);
      }
#pragma unroll 1
      for (int it = 0; it < 8; ++it) {
        const int rr = it * 4 + rsub;
        float4 o;
        asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(o.x), "=f"(o.y), "=f"(o.z), "=f"(o.w)
                     : "r"(stage + (uint32_t)rr * 128u + (uint32_t)(((lane & 7) ^ (rr & 7)) * 16)) : "memory");
        const int orow = row0 + rr;
        if (orow >= p.M || col >= p.N) continue;
        if (p.ws) {
          float* dst = p.ws + ((int64_t)blockIdx.z * p.M + orow) * p.N + col;
          if (ws_vec && full4) *reinterpret_cast<float4*>(dst) = o;
          else {
            dst[0] = o.x;
            if (col + 1 < p.N) dst[1] = o.y;
            if (col + 2 < p.N) dst[2] = o.z;
            if (col + 3 < p.N) dst[3] = o.w;
          }
        } else {
          float* dst = p.C + (int64_t)orow * p.ldc + col;
          o.x += b4.x; o.y += b4.y; o.z += b4.z; o.w += b4.w;
          if (vec_ok && full4) {
            if (p.accumulate) {
              const float4 cv = *reinterpret_cast<const float4*>(dst);
              o.x += cv.x; o.y += cv.y; o.z += cv.z; o.w += cv.w;
            }
            if (p.relu) {
              o.x = relu_apply(o.x, p.relu); o.y = relu_apply(o.y, p.relu);
              o.z = relu_apply(o.z, p.relu); o.w = relu_apply(o.w, p.relu);
            }
            *reinterpret_cast<float4*>(dst) = o;
          } else {
            const float ov[4] = {o.x, o.y, o.z, o.w};
#pragma unroll
            for (int e2 = 0; e2 < 4; ++e2) {
              if (col + e2 < p.N) {
                float t = ov[e2];
                if (p.accumulate) t += dst[e2];
                if (p.relu) t = relu_apply(t, p.relu);
                dst[e2] = t;
              }
            }
          }
        }
      }
      __syncwarp();                                     // the block is reused by the next pass
    }
  }
  MC_STAMP(gemm, 5);
  tcgen05_fence_before();
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)BN) : "memory");
  }
  MC_STAMP(gemm, 6);
  if (p.ws && p.counters) {
    // Split-K without a second kernel: the LAST slice CTA of a tile to arrive adds the partials in
    // slice order 0..S-1 (the same order whichever CTA happens to be last: deterministic) and
    // applies the epilogue.  The counter resets itself for the next call.
    __shared__ int s_last;
    __threadfence();
    __syncthreads();
    const int tile_id = tile_m * gridDim.x + tile_n;
    if (threadIdx.x == 0) {
      const int ticket = atomicAdd(p.counters + tile_id, 1);
      s_last = (ticket == (int)gridDim.z - 1) ? 1 : 0;
      if (s_last) p.counters[tile_id] = 0;
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    const int splits = gridDim.z;
    const int64_t slice = (int64_t)p.M * p.N;
    const int rows = min(BM, p.M - m0), cols = min(BN, p.N - n0);
    const bool vec = ((p.N & 3) == 0) && ((p.ldc & 3) == 0) && ((reinterpret_cast<uintptr_t>(p.C) & 15u) == 0) &&
                     ((reinterpret_cast<uintptr_t>(p.ws) & 15u) == 0);
    if (vec) {
      const int c4n = cols >> 2;                      // N % 4 == 0 and n0 % 4 == 0 -> cols % 4 == 0
      for (int i = threadIdx.x; i < rows * c4n; i += THREADS) {
        const int r = i / c4n, c = (i - r * c4n) << 2;
        const float* src = p.ws + (int64_t)(m0 + r) * p.N + n0 + c;
        float4 o = __ldcg(reinterpret_cast<const float4*>(src));
        for (int z = 1; z < splits; ++z) {
          const float4 q = __ldcg(reinterpret_cast<const float4*>(src + (int64_t)z * slice));
          o.x += q.x; o.y += q.y; o.z += q.z; o.w += q.w;
        }
        float* dst = p.C + (int64_t)(m0 + r) * p.ldc + n0 + c;
        if (p.bias) {
          const float4 b = __ldg(reinterpret_cast<const float4*>(p.bias + n0 + c));
          o.x += b.x; o.y += b.y; o.z += b.z; o.w += b.w;
        }
        if (p.accumulate) {
          const float4 q = *reinterpret_cast<const float4*>(dst);
          o.x += q.x; o.y += q.y; o.z += q.z; o.w += q.w;
        }
        if (p.relu) {
          o.x = relu_apply(o.x, p.relu); o.y = relu_apply(o.y, p.relu);
          o.z = relu_apply(o.z, p.relu); o.w = relu_apply(o.w, p.relu);
        }
        *reinterpret_cast<float4*>(dst) = o;
      }
    } else {
      for (int i = threadIdx.x; i < rows * cols; i += THREADS) {
        const int r = i / cols, c = i - r * cols;
        const float* src = p.ws + (int64_t)(m0 + r) * p.N + n0 + c;
        float o = __ldcg(src);
        for (int z = 1; z < splits; ++z) o += __ldcg(src + (int64_t)z * slice);
        float* dst = p.C + (int64_t)(m0 + r) * p.ldc + n0 + c;
        if (p.bias) o += __ldg(p.bias + n0 + c);
        if (p.accumulate) o += *dst;
        if (p.relu) o = relu_apply(o, p.relu);
        *dst = o;
      }
    }
  }
}

// ------------------------------- host side --------------------------------------------
EncodeTiledFn tensor_map_encoder() {
  static EncodeTiledFn g_encode = nullptr;
  if (g_encode) return g_encode;
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult qres;
  cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
  if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !fn) {
    set_error("cuTensorMapEncodeTiled unavailable (%s)", cudaGetErrorString(e));
    return nullptr;
  }
  g_encode = (EncodeTiledFn)fn;
  return g_encode;
}

int make_tmap(CUtensorMap* map, const float* ptr, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
              const uint32_t* box, bool atom32) {
  EncodeTiledFn enc = tensor_map_encoder();
  if (!enc) return 1;
  cuuint64_t d[5], st[4];
  cuuint32_t bx[5], es[5];
  for (int i = 0; i < rank; ++i) { d[i] = dims[i]; bx[i] = box[i]; es[i] = 1; }
  for (int i = 0; i + 1 < rank; ++i) st[i] = strides_bytes[i];
  static int promo = -1;
  if (promo < 0) { const char* e = getenv("MC_TMAP_PROMO"); promo = e ? atoi(e) : 3; }
  const CUtensorMapL2promotion l2p = promo == 0 ? CU_TENSOR_MAP_L2_PROMOTION_NONE
                                     : promo == 1 ? CU_TENSOR_MAP_L2_PROMOTION_L2_64B
                                     : promo == 2 ? CU_TENSOR_MAP_L2_PROMOTION_L2_128B : CU_TENSOR_MAP_L2_PROMOTION_L2_256B;
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, (cuuint32_t)rank, const_cast<float*>(ptr), d, st, bx, es,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, atom32 ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
                   l2p, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed with CUresult %d (rank %d, dims %llu x %llu, box %u x %u)", (int)r, rank,
              (unsigned long long)dims[0], (unsigned long long)(rank > 1 ? dims[1] : 1), box[0], rank > 1 ? box[1] : 1);
    return 1;
  }
  return 0;
}

// 2-D fp32 matrix, `inner` contiguous elements per row, `outer` rows of pitch ld elements.
int make_tmap_2d(CUtensorMap* map, const float* ptr, int64_t inner, int64_t outer, int64_t ld, int box_inner,
                 int box_outer, bool mn_major) {
  const uint64_t dims[2] = {(uint64_t)inner, (uint64_t)outer};
  const uint64_t strides[1] = {(uint64_t)ld * 4};
  const uint32_t box[2] = {(uint32_t)box_inner, (uint32_t)box_outer};
  return make_tmap(map, ptr, 2, dims, strides, box, mn_major);
}

// ---- persistent variant (grids beyond one wave) ----------------------------------------------
// One CTA per SM walks the rastered tile list.  Six warps: 0-3 drain accumulators (warp q owns TMEM lanes
// 32q..32q+31), warp 4 is the TMA producer, warp 5 issues the MMAs.  TWO accumulators in TMEM (2 x BN
// columns): the MMA warp fills accumulator i+1 while the epilogue warps drain accumulator i, and the operand
// ring never stops at a tile boundary -- the non-persistent kernel above pays TMEM allocation, barrier
// set-up, a pipeline fill and an un-overlapped epilogue per 128 x BN tile (profiles/r1z/prof_gemm_4096.txt:
// 54.7 % tensor-pipe active at 4096^3).
constexpr int PTHREADS = 320;   // warps 0-7: epilogue (two per TMEM lane quadrant, half the columns each), 8: TMA producer, 9: MMA issuer

__device__ __forceinline__ void tile_of(const GemmParams& p, int lin, int tiles_m, int tiles_n, int& tile_m, int& tile_n) {
  if (p.group_m > 0) {
    const int per_group = p.group_m * tiles_n;
    const int first_m = (lin / per_group) * p.group_m;
    const int gsize = min(tiles_m - first_m, p.group_m);
    const int r = lin % per_group;
    tile_m = first_m + r % gsize;
    tile_n = r / gsize;
  } else {
    tile_m = lin / tiles_n;
    tile_n = lin % tiles_n;
  }
}

template <int BN, int STAGES>
__global__ void __launch_bounds__(PTHREADS, 1) gemm_tf32_persistent_kernel(const __grid_constant__ CUtensorMap tmA,
                                                                           const __grid_constant__ CUtensorMap tmB,
                                                                           const __grid_constant__ CUtensorMap tmBh,
                                                                           const GemmParams p) {
  constexpr uint32_t A_BYTES = BM * BK * 4;
  constexpr uint32_t B_BYTES = BN * BK * 4;
  constexpr uint32_t STAGE_BYTES = A_BYTES + B_BYTES;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bar_base = smem_base + STAGES * STAGE_BYTES;   // full[S], empty[S], tfull[2], tempty[2]
  // Work list: items below `half_from` are whole tiles; from there on item i is half (i - half_from) & 1 of tile
  // half_from + (i - half_from) / 2, BN / 2 columns wide (tmBh: the K-major B box of that width).  The host splits
  // the tiles of the last, less-than-half-full wave so that it costs half a tile time (4096^3: 3 waves + 68 tiles).
  const int half_from = p.half_from;
  auto item_of = [&](int lin, int& tile_lin, int& half) {
    if (half_from < 0 || lin < half_from) { tile_lin = lin; half = -1; }
    else { tile_lin = half_from + ((lin - half_from) >> 1); half = (lin - half_from) & 1; }
  };
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (STAGES + s); };
  auto tfull_bar = [&](int a) { return bar_base + 8u * (2 * STAGES + a); };
  auto tempty_bar = [&](int a) { return bar_base + 8u * (2 * STAGES + 2 + a); };
  const uint32_t tmem_slot = bar_base + 8u * (2 * STAGES + 4);
  const uint32_t turn_base = (tmem_slot + 16u + 127u) & ~127u;  // 8 x 4 KB: one 32 x 32 block per epilogue warp

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tiles_m = (p.M + BM - 1) / BM, tiles_n = (p.N + BN - 1) / BN;
  const int total = half_from < 0 ? tiles_m * tiles_n : half_from + 2 * (tiles_m * tiles_n - half_from);
  const int num_kb = (p.K + BK - 1) / BK;

  if (warp == 8 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmA)) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmB)) : "memory");
    for (int s = 0; s < STAGES; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
    for (int a = 0; a < 2; ++a) { mbar_init(tfull_bar(a), 1); mbar_init(tempty_bar(a), 8); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  if (warp == 9) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"((uint32_t)(2 * BN))
                 : "memory");
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
      const bool a_mn = p.a_mn != 0, b_mn = p.b_mn != 0;
      int s = 0;
      uint32_t ph = 0;
      for (int lin = blockIdx.x; lin < total; lin += gridDim.x) {
        int tile_m, tile_n, tile_lin, half;
        item_of(lin, tile_lin, half);
        tile_of(p, tile_lin, tiles_m, tiles_n, tile_m, tile_n);
        const int m0 = tile_m * BM, n0 = tile_n * BN + (half > 0 ? BN / 2 : 0);
        const int nbox = half < 0 ? BN / 32 : BN / 64;
        const uint32_t stage_tx = half < 0 ? STAGE_BYTES : A_BYTES + B_BYTES / 2;
        int k0 = 0;
        for (int i = 0; i < num_kb; ++i, k0 += BK) {
          mbar_wait(empty_bar(s), ph ^ 1u);
          const uint32_t fb = full_bar(s);
          mbar_expect_tx(fb, stage_tx);
          const uint32_t sa = smem_base + s * STAGE_BYTES, sb = sa + A_BYTES;
          if (!a_mn) {
            tma_load_2d(sa, &tmA, fb, k0, m0);
          } else {
#pragma unroll
            for (int j = 0; j < BM / 32; ++j) tma_load_2d(sa + j * 4096, &tmA, fb, m0 + 32 * j, k0);
          }
          if (!b_mn) {
            tma_load_2d(sb, half < 0 ? &tmB : &tmBh, fb, k0, n0);
          } else {
            for (int j = 0; j < nbox; ++j) tma_load_2d(sb + j * 4096, &tmB, fb, n0 + 32 * j, k0);
          }
          if (++s == STAGES) { s = 0; ph ^= 1u; }
        }
      }
    }
  } else if (warp == 9) {
    if (elect_one()) {
      // ------------------------------ MMA issuer --------------------------------------
      const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)p.a_mn << 15) | ((uint32_t)p.b_mn << 16) |
                             ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
      const uint64_t adesc0 = p.a_mn ? make_desc(0, 4096, 512, 1) : make_desc(0, 16, 1024, 2);
      const uint64_t bdesc0 = p.b_mn ? make_desc(0, 4096, 512, 1) : make_desc(0, 16, 1024, 2);
      const uint32_t astep = p.a_mn ? 64u : 2u, bstep = p.b_mn ? 64u : 2u;
      int s = 0, acc = 0;
      uint32_t ph = 0, aph = 0;
      const uint32_t idesc_half = (idesc & ~(0x3Fu << 17)) | ((uint32_t)(BN >> 4) << 17);
      for (int lin = blockIdx.x; lin < total; lin += gridDim.x) {
        const uint32_t idesc_t = (half_from >= 0 && lin >= half_from) ? idesc_half : idesc;
        mbar_wait(tempty_bar(acc), aph ^ 1u);          // the epilogue has drained this accumulator
        tcgen05_fence_after();
        const uint32_t dacc = tmem_base + (uint32_t)(acc * BN);
        for (int i = 0; i < num_kb; ++i) {
          mbar_wait(full_bar(s), ph);
          tcgen05_fence_after();
          const uint32_t sa = smem_base + s * STAGE_BYTES, sb = sa + A_BYTES;
          const uint64_t adesc = adesc0 + (uint64_t)((sa & 0x3FFFFu) >> 4);
          const uint64_t bdesc = bdesc0 + (uint64_t)((sb & 0x3FFFFu) >> 4);
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k)
            umma_tf32(dacc, adesc + (uint64_t)(astep * k), bdesc + (uint64_t)(bstep * k), idesc_t, (i > 0 || k > 0) ? 1u : 0u);
          umma_commit(empty_bar(s));
          if (++s == STAGES) { s = 0; ph ^= 1u; }
        }
        umma_commit(tfull_bar(acc));
        if (++acc == 2) { acc = 0; aph ^= 1u; }
      }
    }
  } else {
    // ---------------------------------- epilogue (warps 0-7) ----------------------------
    // Warp w drains TMEM lanes 32 (w & 3) .. +31, columns [(w >> 2) width / 2, + width / 2).  Interior tiles: every
    // 32 x 32 block is turned through 4 KB of shared memory (16-byte chunks XOR-swizzled by row) so that eight lanes
    // write one 128-byte row segment and a warp store covers four whole lines -- the pattern of gemm_tf32_kernel.
    // (Four warps storing one ROW per thread took ~1 us per 32-column pass: (12800, 4096, 128), whose tiles have four
    //  k-blocks of mainloop, ran at 110 us against 33 us of output traffic.)
    const bool vec_ok = ((p.ldc & 3) == 0) && ((reinterpret_cast<uintptr_t>(p.C) & 15u) == 0) &&
                        (!p.bias || (reinterpret_cast<uintptr_t>(p.bias) & 15u) == 0);
    const int quad = warp & 3, chalf = warp >> 2;
    const uint32_t turn = turn_base + (uint32_t)warp * 4096u;
    const uint32_t st_addr = turn + (uint32_t)lane * 128u, st_x = (uint32_t)(lane & 7);
    const int j4 = (lane & 7) * 4, rsub = lane >> 3;
    const float negv = (p.relu == 2) ? -0.0f : 0.f;
    int acc = 0;
    uint32_t aph = 0;
    for (int lin = blockIdx.x; lin < total; lin += gridDim.x) {
      int tile_m, tile_n, tile_lin, half;
      item_of(lin, tile_lin, half);
      tile_of(p, tile_lin, tiles_m, tiles_n, tile_m, tile_n);
      const int m0 = tile_m * BM, n0 = tile_n * BN + (half > 0 ? BN / 2 : 0);
      const int width = half < 0 ? BN : BN / 2;
      const int cbeg = chalf * (width / 2), cend = cbeg + width / 2;
      mbar_wait_warp(tfull_bar(acc), aph);
      tcgen05_fence_after();
      const uint32_t tacc = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(acc * BN);
      if (vec_ok && m0 + BM <= p.M && n0 + width <= p.N) {
        float* base = p.C + (int64_t)(m0 + quad * 32 + rsub) * p.ldc + n0 + j4;
#pragma unroll 1
        for (int c0 = cbeg; c0 < cend; c0 += 32) {
          float v[32];
          tmem_ld32(tacc + (uint32_t)c0, v);
#pragma unroll
          for (int j = 0; j < 8; ++j)
            asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(st_addr + (((uint32_t)j ^ st_x) * 16u)), "f"(v[4 * j]),
                         "f"(v[4 * j + 1]), "f"(v[4 * j + 2]), "f"(v[4 * j + 3]) : "memory");
          __syncwarp();
          float4 b4 = make_float4(0.f, 0.f, 0.f, 0.f);
          if (p.bias) b4 = __ldg(reinterpret_cast<const float4*>(p.bias + n0 + c0 + j4));
          float* d = base + c0;
          float4 o[8];
#pragma unroll
          for (int it = 0; it < 8; ++it) {
            const int rr = it * 4 + rsub;
            asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(o[it].x), "=f"(o[it].y), "=f"(o[it].z), "=f"(o[it].w)
                         : "r"(turn + (uint32_t)rr * 128u + (uint32_t)(((lane & 7) ^ (rr & 7)) * 16)) : "memory");
          }
          float4 cv[8];
          if (p.accumulate) {
#pragma unroll
            for (int it = 0; it < 8; ++it) cv[it] = *reinterpret_cast<const float4*>(d + (int64_t)(it * 4) * p.ldc);
          }
#pragma unroll
          for (int it = 0; it < 8; ++it) {
            float4 t = make_float4(o[it].x + b4.x, o[it].y + b4.y, o[it].z + b4.z, o[it].w + b4.w);
            if (p.accumulate) { t.x += cv[it].x; t.y += cv[it].y; t.z += cv[it].z; t.w += cv[it].w; }
            if (p.relu) {                                   // relu_apply without branches: NaN passes, -x -> negv, +-0 -> +0
              t.x = (t.x < 0.f) ? negv : ((t.x == 0.f) ? 0.f : t.x); t.y = (t.y < 0.f) ? negv : ((t.y == 0.f) ? 0.f : t.y);
              t.z = (t.z < 0.f) ? negv : ((t.z == 0.f) ? 0.f : t.z); t.w = (t.w < 0.f) ? negv : ((t.w == 0.f) ? 0.f : t.w);
            }
            *reinterpret_cast<float4*>(d + (int64_t)(it * 4) * p.ldc) = t;
          }
          __syncwarp();                                     // the block is reused by the next pass
        }
      } else {
        // edge tiles / unaligned output: one row per thread, bounds checked per element
        const int row = m0 + quad * 32 + lane;
        const bool vec_c = ((p.ldc & 3) == 0) && ((reinterpret_cast<uintptr_t>(p.C) & 15u) == 0);
#pragma unroll 1
        for (int c0 = cbeg; c0 < cend; c0 += 32) {
          float v[32];
          tmem_ld32(tacc + (uint32_t)c0, v);
          const int col0 = n0 + c0;
          if (row < p.M && col0 < p.N) {
            float* dst = p.C + (int64_t)row * p.ldc + col0;
            if (vec_c && col0 + 32 <= p.N && (!p.bias || (reinterpret_cast<uintptr_t>(p.bias) & 15u) == 0)) {
#pragma unroll
              for (int j = 0; j < 32; j += 4) {
                float4 o = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
                if (p.bias) {
                  const float4 b = __ldg(reinterpret_cast<const float4*>(p.bias + col0 + j));
                  o.x += b.x; o.y += b.y; o.z += b.z; o.w += b.w;
                }
                if (p.accumulate) {
                  const float4 c = *reinterpret_cast<const float4*>(dst + j);
                  o.x += c.x; o.y += c.y; o.z += c.z; o.w += c.w;
                }
                if (p.relu) {
                  o.x = relu_apply(o.x, p.relu); o.y = relu_apply(o.y, p.relu);
                  o.z = relu_apply(o.z, p.relu); o.w = relu_apply(o.w, p.relu);
                }
                *reinterpret_cast<float4*>(dst + j) = o;
              }
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j) {
                if (col0 + j < p.N) {
                  float o = v[j];
                  if (p.bias) o += __ldg(p.bias + col0 + j);
                  if (p.accumulate) o += dst[j];
                  if (p.relu) o = relu_apply(o, p.relu);
                  dst[j] = o;
                }
              }
            }
          }
        }
      }
      tcgen05_fence_before();
      __syncwarp();
      if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(tempty_bar(acc)) : "memory");
      if (++acc == 2) { acc = 0; aph ^= 1u; }
    }
  }
  tcgen05_fence_before();
  __syncthreads();
  if (warp == 9) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)(2 * BN)) : "memory");
  }
}

template <int BN, int STAGES>
static int launch_tf32(const CUtensorMap& tmA, const CUtensorMap& tmB, const GemmParams& p, int splits) {
  constexpr size_t SMEM = (size_t)STAGES * (BM * BK * 4 + BN * BK * 4) + (2 * STAGES + 1) * 8 + 16 + 1024;
  static uint32_t configured = 0;      // function attributes are per device
  if (first_use_on_device(&configured))
    MC_CHECK(cudaFuncSetAttribute(gemm_tf32_kernel<BN, STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM));
  dim3 grid((unsigned)((p.N + BN - 1) / BN), (unsigned)((p.M + BM - 1) / BM), (unsigned)splits);
  if (p.cluster > 1) {
    static uint32_t big_ok = 0;
    if (p.cluster > 8 && first_use_on_device(&big_ok))
      MC_CHECK(cudaFuncSetAttribute(gemm_tf32_kernel<BN, STAGES>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = dim3(THREADS); cfg.dynamicSmemBytes = SMEM; cfg.stream = stream();
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 1; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = (unsigned)p.cluster;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = pdl_mode() != 0 ? 2 : 1;
    MC_CHECK(cudaLaunchKernelEx(&cfg, gemm_tf32_kernel<BN, STAGES>, tmA, tmB, p));
    count_launch();
    return 0;
  }
  mc::launch_pdl(gemm_tf32_kernel<BN, STAGES>, grid, THREADS, SMEM, tmA, tmB, p);
  MC_LAUNCHED();
  return 0;
}

}  // namespace tc

static thread_local int g_last_path = -1;

static int gemm_fp32(int ta, int tb, int64_t M, int64_t N, int64_t K, const float* A, int64_t lda, const float* B,
                     int64_t ldb, float* C, int64_t ldc, const float* bias, int relu, int accumulate) {
  const int64_t tiles = ((M + SG_BM - 1) / SG_BM) * ((N + SG_BN - 1) / SG_BN);
  int splits = 1;
  if (tiles < sm_count() && K >= 512) {
    splits = (int)((2 * sm_count()) / tiles);
    const int maxs = (int)(K / 128);
    if (splits > maxs) splits = maxs;
    if (splits > 64) splits = 64;
    if (splits < 1) splits = 1;
  }
  int k_per = (int)((K + splits - 1) / splits);
  k_per = (k_per + SG_BK - 1) / SG_BK * SG_BK;
  splits = (int)((K + k_per - 1) / k_per);
  float* ws = nullptr;
  if (splits > 1) MC_CHECK(cudaMallocAsync((void**)&ws, sizeof(float) * splits * M * N, stream()));
  dim3 grid((unsigned)((N + SG_BN - 1) / SG_BN), (unsigned)((M + SG_BM - 1) / SG_BM), (unsigned)splits);
#define SG_LAUNCH(TA, TB)                                                                                          \
  mc::launch(sgemm_kernel<TA, TB>, grid, SG_THREADS, 0, (int)M, (int)N, (int)K, A, lda, B, ldb, C, ldc, bias, relu, \
                                                          accumulate, k_per, ws)
  if (ta && tb) SG_LAUNCH(true, true);
  else if (ta) SG_LAUNCH(true, false);
  else if (tb) SG_LAUNCH(false, true);
  else SG_LAUNCH(false, false);
#undef SG_LAUNCH
  MC_LAUNCHED();
  if (splits > 1) {
    mc::launch(splitk_reduce_kernel, grid_for(M * N / 4 + 1, 256), 256, 0, ws, splits, M, N, C, ldc, bias, relu, accumulate);
    MC_LAUNCHED();
    MC_CHECK(cudaFreeAsync(ws, stream()));
  }
  return 0;
}

template <int BN, int STAGES>
static int launch_tf32_persistent(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmBh, tc::GemmParams p,
                                  int64_t tiles, bool split_tail) {
  using namespace tc;
  constexpr size_t SMEM = (size_t)STAGES * (BM * BK * 4 + BN * BK * 4) + (2 * STAGES + 4) * 8 + 16 + 128 + 8 * 4096 + 1024;
  static uint32_t configured = 0;
  if (first_use_on_device(&configured))
    MC_CHECK(cudaFuncSetAttribute(gemm_tf32_persistent_kernel<BN, STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM));
  const unsigned grid = (unsigned)(tiles < sm_count() ? tiles : sm_count());
  p.half_from = -1;
  if (split_tail && tiles > grid) {
    const int64_t tail = tiles % grid;
    if (tail > 0 && 2 * tail <= grid) p.half_from = (int)(tiles - tail);
  }
  mc::launch_pdl(gemm_tf32_persistent_kernel<BN, STAGES>, grid, PTHREADS, SMEM, tmA, tmB, tmBh, p);
  MC_LAUNCHED();
  return 0;
}

constexpr int kMaxFixupTiles = 16384;
static int* splitk_counters() {
  static int* g_counters[16] = {};      // device memory: one block per device
  int*& c = g_counters[device() & 15];
  if (!c) {
    if (cudaMalloc((void**)&c, sizeof(int) * kMaxFixupTiles) != cudaSuccess) { c = nullptr; return nullptr; }
    cudaMemset(c, 0, sizeof(int) * kMaxFixupTiles);
  }
  return c;
}

// Split-K partials live in ONE persistent block per device (grown on demand, old blocks are kept alive
// because captured graphs may still reference them).  Every kernel of the library runs on one stream, so
// reuse in program order is safe; a cudaMallocAsync / cudaFreeAsync pair per call put two memory nodes
// per GEMM into the captured step graph (suspected part of the per-node cost).
static float* splitk_workspace(size_t bytes) {
  static float* ws[32] = {};            // [device][main / side stream]
  static size_t cap[32] = {};
  const int d = ((device() & 15) << 1) | side_index();
  if (cap[d] < bytes) {
    size_t want = bytes < ((size_t)32 << 20) ? ((size_t)32 << 20) : bytes * 2;
    float* fresh = nullptr;
    if (cudaMalloc((void**)&fresh, want) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    ws[d] = fresh;
    cap[d] = want;
  }
  return ws[d];
}

static bool tma_ok(const float* p, int64_t ld) { return aligned16(p) && (ld % 4 == 0); }

// `ext_ws` non-null = raw-partials mode (gemm_tf32_partials below): the K slices land in
// ext_ws[z][M][N], z < *splits_out <= max_splits, and nobody adds them up here.
static int gemm_tf32(int ta, int tb, int64_t M, int64_t N, int64_t K, const float* A, int64_t lda, const float* B,
                     int64_t ldb, float* C, int64_t ldc, const float* bias, int relu, int accumulate,
                     float* ext_ws = nullptr, int max_splits = 0, int* splits_out = nullptr) {
  using namespace tc;
  static int force_bn = -1, min_kb = -1;
  if (force_bn < 0) { const char* e = getenv("MC_GEMM_BN"); force_bn = e ? atoi(e) : 0; }
  if (min_kb < 0) { const char* e = getenv("MC_GEMM_MIN_KB"); min_kb = e ? atoi(e) : 8; }
  static int split_from_kb = -1;     // smallest k-block count that is split at all (experiments: MC_GEMM_SPLIT_FROM_KB)
  if (split_from_kb < 0) { const char* e = getenv("MC_GEMM_SPLIT_FROM_KB"); split_from_kb = e ? atoi(e) : 32; }
  // Tile width: 128 columns, except (measured, profiles/r1za gemm sweeps) for problems that are both few
  // tiles and short in K -- there 64-column tiles double the CTAs without deepening split-K:
  // (1024,512,512) 9.4 -> 8.2 us, (512,512,1024) 9.5 -> 8.3, (256,512,3072) 10.7 -> 9.5, (128,1024,4096)
  // 11.0 -> 10.2; the large-K / many-tile shapes ((1024,512,3072) 13.9 vs 16.7 us) keep 128.
  int BN = (N > 64) ? 128 : 64;
  if (BN == 128) {
    const int64_t tiles128 = ((M + BM - 1) / BM) * ((N + 127) / 128);
    const int64_t kb = (K + BK - 1) / BK;
    if (tiles128 <= 32 && kb >= 16 && tiles128 * kb <= 1024) BN = 64;
    // raw-partials mode with a long K (the per-step backward GEMM of the LSTM, (128, 1024, 4096)): alone the 64-column
    // tiles win (9.2 vs 9.7 us), inside the sequence -- GEMM, gate kernel, GEMM ... -- 128-column tiles with 16 slices
    // do: 10.03 -> 9.63 us per step (less operand traffic per CTA: 256 KB instead of 360 KB)
    if (ext_ws && kb >= 64 && N % 128 == 0) BN = 128;
  }
  if (force_bn) BN = force_bn;
  CUtensorMap tmA, tmB;
  // A operand: op(A) is (M, K).  Stored (M,K) row-major -> K-major tile {BK x BM};
  // stored (K,M) (trans_a) -> MN-major: boxes of {32 (M) x BK (K)}.
  if (!ta) { if (make_tmap_2d(&tmA, A, K, M, lda, BK, BM, false)) return 1; }
  else     { if (make_tmap_2d(&tmA, A, M, K, lda, 32, BK, true)) return 1; }
  // B operand: op(B) is (K, N).  Stored (K,N) row-major -> MN-major; stored (N,K) (trans_b) -> K-major.
  if (!tb) { if (make_tmap_2d(&tmB, B, N, K, ldb, 32, BK, true)) return 1; }
  else     { if (make_tmap_2d(&tmB, B, K, N, ldb, BK, BN, false)) return 1; }

  const int64_t tiles = ((M + BM - 1) / BM) * ((N + BN - 1) / BN);
  const int num_kb = (int)((K + BK - 1) / BK);
  int splits = 1;
  // K >= 1024 only: with 16..31 k-blocks the combine launch costs more than the second slice saves
  // (profiles/r1za: (1024,512,512) 8.2 us split in two, 7.6 us whole)
  // On the side stream (weight gradients beside the backward chain) a GEMM has slack and its CTAs compete with the
  // critical chain for SMs: K is split there only from 64 k-blocks on (mlp_c3: 28 -> 25 launches, 0.1465 -> 0.1458 ms).
  const int split_from = (side_index() && split_from_kb < 64) ? 64 : split_from_kb;
  if (tiles < sm_count() && num_kb >= split_from) {
    splits = (int)(sm_count() / tiles);
    const int maxs = num_kb / min_kb;  // at least 8 k-blocks (K = 256) per slice
    if (splits > maxs) splits = maxs;
    if (splits < 1) splits = 1;
  }
  if (ext_ws && splits > max_splits) splits = max_splits;
  int kb_per = (num_kb + splits - 1) / splits;
  splits = (num_kb + kb_per - 1) / kb_per;

  GemmParams p;
  p.C = C; p.ldc = ldc; p.bias = bias; p.ws = nullptr;
  p.M = (int)M; p.N = (int)N; p.K = (int)K;
  p.relu = relu; p.accumulate = accumulate;
  p.a_mn = ta ? 1 : 0;
  p.b_mn = tb ? 0 : 1;
  p.kb_per_split = kb_per;
  p.counters = nullptr;
  p.cluster = 0;
  p.half_from = -1;
  p.lstm_mode = 0;
  p.lstm = LstmEpilogue();
  // raster only when the grid is larger than one wave: keep <= ~48 MB of A panels resident in the 126 MB L2
  p.group_m = 0;
  {
    static int raster = -1;
    if (raster < 0) { const char* e = getenv("MC_GEMM_RASTER"); raster = (e && e[0] == '0') ? 0 : 1; }
    const int64_t tiles_m = (M + BM - 1) / BM;
    if (raster && tiles > sm_count() && tiles_m > 1) {
      const int64_t panel = (int64_t)BM * K * 4;
      int64_t g = ((int64_t)48 << 20) / (panel > 0 ? panel : 1);
      if (g < 2) g = 2;
      if (g > tiles_m) g = tiles_m;
      p.group_m = (int)g;
    }
  }
  // Cluster split-K (partials reduced through distributed shared memory, S in {2,4,8,16} slices, every
  // slice non-empty) is OFF by default: it removes splitk_reduce_kernel and the HBM workspace, and is
  // -1..+1 us per call in isolation, but whole steps got slower (C3 0.341 -> 0.348 ms, C5 3.83 -> 4.06 ms,
  // profiles/r1z: S co-scheduled SMs per tile delay the next kernel's start).  MC_GEMM_CLUSTER=<max S>
  // enables it for experiments.
  static int use_cluster = -1;
  if (use_cluster < 0) { const char* e = getenv("MC_GEMM_CLUSTER"); use_cluster = e ? atoi(e) : 0; }
  if (ext_ws) {
    if (splits > 1) p.ws = ext_ws;
    else { p.C = ext_ws; p.ldc = N; p.bias = nullptr; p.relu = 0; p.accumulate = 0; }
    *splits_out = splits;
    return (BN == 128) ? launch_tf32<128, 6>(tmA, tmB, p, splits) : launch_tf32<64, 8>(tmA, tmB, p, splits);
  }
  if (splits > 1 && use_cluster > 1) {
    int S = 2;
    while (S * 2 <= splits && S * 2 <= use_cluster && S * 2 <= 16) S *= 2;
    for (; S > 1; S /= 2) {
      const int per = (num_kb + S - 1) / S;
      if ((S - 1) * per < num_kb) { kb_per = per; break; }
    }
    if (S > 1) { splits = S; p.kb_per_split = kb_per; p.cluster = S; }
  }
  if (splits > 1 && !p.cluster) {
    p.ws = splitk_workspace(sizeof(float) * splits * M * N);
    MC_REQUIRE(p.ws != nullptr, "split-K workspace allocation failed");
    // In-kernel fixup (last-arriving slice CTA reduces) is OFF by default: one SM streaming S x 64 KB
    // of partials is latency-bound (c3_fwd2: 9.9 -> 27 us measured, r1t); the separate reduce kernel
    // spreads the same bytes over every SM.  MC_GEMM_FIXUP=1 re-enables it for experiments.
    static int fixup = -1;
    if (fixup < 0) { const char* e = getenv("MC_GEMM_FIXUP"); fixup = (e && atoi(e) == 1) ? 1 : 0; }
    if (fixup && tiles <= kMaxFixupTiles) p.counters = splitk_counters();
  }
  // grids beyond one wave, no split: the persistent, TMEM-double-buffered kernel (profiles/r2d: 4096^3 535 -> 570
  // TF/s, 8192^3 560 -> 585; MC_GEMM_PERSISTENT=0 selects the one-tile-per-CTA kernel)
  static int persistent = -1;
  if (persistent < 0) { const char* e = getenv("MC_GEMM_PERSISTENT"); persistent = (e && e[0] == '0') ? 0 : 1; }
  if (persistent && splits == 1 && BN == 128 && tiles > sm_count()) {
    // 256-column tiles: one tcgen05.mma of M128 x N256 x K8 reads 4 KB of A + 8 KB of B from shared memory for 138
    // clocks of tensor work (87 B/clk) where the 128 x 128 tile reads 8 KB per 69 clocks (118 B/clk of the 128 B/clk an
    // SM has), and A panels are re-read half as often.  Both accumulators fill TMEM (2 x 256 columns).
    // MEASURED (profiles/r2t): 4096^3 562 -> 629 TF/s, 8192^3 573 -> 707 TF/s.  Taken when its waves are not worse:
    // a 256-wide tile costs ~1.6 of a 128-wide one (MC_GEMM_BN256=0 keeps 128 everywhere).
    static int bn256 = -1;
    if (bn256 < 0) { const char* e = getenv("MC_GEMM_BN256"); bn256 = e ? atoi(e) : 1; }
    const int64_t tiles256 = ((M + BM - 1) / BM) * ((N + 255) / 256);
    const int64_t sms = sm_count();
    const int64_t tail256 = tiles256 > sms ? tiles256 % sms : 0;      // a last wave at most half full runs as half tiles
    const double t128 = (double)((tiles + sms - 1) / sms);
    const double t256 = 1.6 * ((double)(tiles256 / sms) + (tiles256 % sms == 0 ? 0.0 : ((tail256 > 0 && 2 * tail256 <= sms) ? 0.5 : 1.0)));
    if (bn256 && N % 256 == 0 && t256 <= t128) {
      CUtensorMap tmB2;
      if (!tb) { if (make_tmap_2d(&tmB2, B, N, K, ldb, 32, BK, true)) return 1; }
      else     { if (make_tmap_2d(&tmB2, B, K, N, ldb, BK, 256, false)) return 1; }
      // (tmB = the 128-wide box of the same matrix: the half tiles of a split last wave)
      static int tail = -1;
      if (tail < 0) { const char* e = getenv("MC_GEMM_TAIL_SPLIT"); tail = (e && e[0] == '0') ? 0 : 1; }
      return launch_tf32_persistent<256, 4>(tmA, tmB2, tmB, p, tiles256, tail != 0);
    }
    return launch_tf32_persistent<128, 6>(tmA, tmB, tmB, p, tiles, false);
  }
  int rc = (BN == 128) ? launch_tf32<128, 6>(tmA, tmB, p, splits) : launch_tf32<64, 8>(tmA, tmB, p, splits);
  if (rc == 0 && splits > 1 && !p.counters && !p.cluster) {
    mc::launch(splitk_reduce_kernel, grid_for(M * N / 4 + 1, 256), 256, 0, p.ws, splits, M, N, C, ldc, bias, relu, accumulate);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { set_error("splitk reduce launch failed: %s", cudaGetErrorString(e)); rc = 1; }
    else count_launch();
  }
  return rc;
}

// One LSTM time step as ONE launch: the recurrent GEMM (cluster split-K, partials reduced through distributed
// shared memory) with the gate arithmetic as its epilogue (LstmEpilogue above).  *taken = 0 when the shape does
// not fit (the caller then runs the GEMM and the gate kernel separately).
//   mode 1: op(A) = h_{t-1} (M = batch, K = H), B = Wh' (K, N = 4H) unit-major columns
//   mode 2: op(A) = dA_t (M = batch, K = 4H), B = Wh (N = H rows of K = 4H): trans_b
int gemm_tf32_lstm(int mode, int64_t M, int64_t N, int64_t K, const float* A, int64_t lda, const float* B, int64_t ldb,
                   const tc::LstmEpilogue& epi, int* taken) {
  using namespace tc;
  *taken = 0;
  const int tb = (mode == 2) ? 1 : 0;
  if (M <= 0 || N < 64 || K < 64 || (N & 3) || (epi.H & 3) || !tma_ok(A, lda) || !tma_ok(B, ldb)) return 0;
  static int enabled = -1;
  // MEASURED (gpurun r2g, C5): correct (golden / full-size / step-graph LSTM tests pass) but SLOWER than GEMM +
  // gate kernel -- 18 vs 10.5 us per time step: the 128 epilogue threads of a CTA each walk 2-8 items one after
  // the other, every item a DSMEM round trip, then dependent global loads, then exp/tanh, while the separate gate
  // kernel spreads the same items over 32 k threads.  Kept behind MC_LSTM_FUSED=1 (tests run both paths).
  if (enabled < 0) { const char* e = getenv("MC_LSTM_FUSED"); enabled = (e && e[0] == '1') ? 1 : 0; }
  if (!enabled) return 0;
  const int BN = 64;                                      // 4H / 64 (or H / 64) tiles x S slices ~ one CTA per SM
  const int64_t tiles = ((M + BM - 1) / BM) * ((N + BN - 1) / BN);
  const int num_kb = (int)((K + BK - 1) / BK);
  int S = 16;
  while (S > 1 && (tiles * S > sm_count() || num_kb / S < 4)) S >>= 1;
  int kb_per = 0;
  for (; S > 1; S >>= 1) {
    kb_per = (num_kb + S - 1) / S;
    if ((S - 1) * kb_per < num_kb) break;                 // every slice non-empty
  }
  if (S < 2) return 0;
  CUtensorMap tmA, tmB;
  if (make_tmap_2d(&tmA, A, K, M, lda, BK, BM, false)) return 1;
  if (!tb) { if (make_tmap_2d(&tmB, B, N, K, ldb, 32, BK, true)) return 1; }
  else     { if (make_tmap_2d(&tmB, B, K, N, ldb, BK, BN, false)) return 1; }
  GemmParams p;
  p.C = nullptr; p.ldc = N; p.bias = nullptr; p.ws = nullptr; p.counters = nullptr;
  p.M = (int)M; p.N = (int)N; p.K = (int)K;
  p.relu = 0; p.accumulate = 0;
  p.a_mn = 0;
  p.b_mn = tb ? 0 : 1;
  p.kb_per_split = kb_per;
  p.cluster = S;
  p.group_m = 0;
  p.lstm_mode = mode;
  p.lstm = epi;
  *taken = 1;
  return launch_tf32<64, 8>(tmA, tmB, p, S);
}

// The split-K GEMM without its combine pass, for consumers that add the slices themselves (the LSTM gate
// kernels of lstm.cu read  a + sum_z ws[z]  and so save a launch and a round trip of the (N,4H) tile
// per time step).  Returns 0 and *splits = 0 when the problem does not take the tensor-core path
// (the caller then uses mc_gemm).
int gemm_tf32_partials(int ta, int tb, int64_t M, int64_t N, int64_t K, const float* A, int64_t lda, const float* B,
                       int64_t ldb, float* ws, int max_splits, int* splits) {
  *splits = 0;
  if (M <= 0 || N <= 0 || K <= 0 || max_splits < 1) return 0;
  if (N <= 16 || K <= 16 || !(tma_ok(A, lda) && tma_ok(B, ldb) && (M >= 32 || N >= 32)) || !aligned16(ws) || (N & 3)) return 0;
  return gemm_tf32(ta, tb, M, N, K, A, lda, B, ldb, nullptr, N, nullptr, 0, 0, ws, max_splits, splits);
}

}  // namespace mc

using namespace mc;

extern "C" {

int mc_gemm(int trans_a, int trans_b, int64_t M, int64_t N, int64_t K, const float* A, int64_t lda, const float* B,
            int64_t ldb, float* C, int64_t ldc, const float* bias, int relu, int accumulate, int precision) {
  MC_REQUIRE(M >= 0 && N >= 0 && K >= 0, "negative extent");
  MC_REQUIRE(M < 2147483647LL && N < 2147483647LL && K < 2147483647LL, "extent exceeds int32");
  if (M == 0 || N == 0) return 0;
  if (K == 0) {
    // empty contraction: C = bias (+C), NumPy gives zeros
    if (!accumulate) {
      for (int64_t m = 0; m < M; ++m) if (mc_fill(C + m * ldc, 0.f, N)) return 1;
    }
    if (bias) return mc_binary_vec(MC_ADD, C, bias, 0, 0, C, M, N);
    return 0;
  }
  if (gemm_skinny(trans_a, trans_b, M, N, K, A, lda, B, ldb, C, ldc, bias, relu, accumulate)) {
    g_last_path = 0;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { set_error("skinny gemm launch failed: %s", cudaGetErrorString(e)); return 1; }
    count_launch();
    return 0;
  }
  const bool tma = tma_ok(A, lda) && tma_ok(B, ldb) && (M >= 32 || N >= 32);
  if (precision == MC_GEMM_TF32 && tma) {
    g_last_path = 1;
    return gemm_tf32(trans_a, trans_b, M, N, K, A, lda, B, ldb, C, ldc, bias, relu, accumulate);
  }
  g_last_path = 0;
  return gemm_fp32(trans_a, trans_b, M, N, K, A, lda, B, ldb, C, ldc, bias, relu, accumulate);
}

int mc_gemm_last_path(int* path) {
  *path = g_last_path;
  return 0;
}

}  // extern "C"
