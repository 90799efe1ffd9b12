// Row softmax, log-sum-exp and the fused softmax cross-entropy forward / backward.
//
// HBM-bound: softmax 8*rows*cols bytes (read once, write once).  One warp owns one row
// and keeps it in registers (cols <= 1024: 8 float4 per lane), so max / exp / sum / scale
// need a single pass over HBM; wider rows use one CTA per row with the row staged in
// shared memory (<= 48 KB) or, beyond that, three streaming passes.
#include "common.cuh"

namespace mc {

int reduce_sum_internal(const float* x, float* out, int64_t outer, int64_t red, int64_t inner);

constexpr int SM_THREADS = 256;
constexpr int ROW_REGS = 32;  // floats cached per lane -> cols <= 1024

enum { MODE_SOFTMAX = 0, MODE_LSE = 1, MODE_CE = 2 };

// exp(x - m) as ONE fma + ONE ex2.approx (2^-22 relative): with accurate expf and a NaN-checking max
// the row kernels were instruction-bound (43 % of the HBM roofline at 1024 columns, profiles/r1a).
// NaN needs no special max: fmaxf skips it, exp(NaN - m) is NaN, so the row sum and with it every
// output of that row is NaN -- what NumPy's max-subtracted softmax yields as well.
__device__ __forceinline__ float exp_sub(float x, float m_log2e) {
  float r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(fmaf(x, 1.4426950408889634f, -m_log2e)));
  return r;
}

// One warp per row, whole row in registers.
// MODE_SOFTMAX: out = softmax(x).  MODE_LSE: out[row] = logsumexp.  MODE_CE: out = probs and
// rowloss[row] = -log p[label] (labels) or -sum_j onehot[j] * log p[j].
template <int MODE>
__global__ void __launch_bounds__(SM_THREADS) softmax_warp_kernel(const float* __restrict__ x, float* __restrict__ out,
                                                                  const int32_t* __restrict__ labels,
                                                                  const float* __restrict__ onehot,
                                                                  float* __restrict__ rowloss, int64_t rows,
                                                                  int64_t cols) {
  mc::pdl_sync();
  const int lane = threadIdx.x & 31;
  const int64_t warps_per_grid = (int64_t)gridDim.x * (SM_THREADS / 32);
  const bool vec = ((cols & 3) == 0) && ((reinterpret_cast<uintptr_t>(x) & 15u) == 0) &&
                   ((reinterpret_cast<uintptr_t>(out) & 15u) == 0);
  for (int64_t row = (int64_t)blockIdx.x * (SM_THREADS / 32) + (threadIdx.x >> 5); row < rows; row += warps_per_grid) {
    const float* __restrict__ p = x + row * cols;
    float v[ROW_REGS];
    float m = -INFINITY;
    if (vec) {
      const float4* __restrict__ p4 = reinterpret_cast<const float4*>(p);
      const int64_t c4 = cols >> 2;
#pragma unroll
      for (int j = 0; j < ROW_REGS / 4; ++j) {
        int64_t i = lane + 32 * j;
        if (i < c4) {
          float4 t = p4[i];
          v[4 * j] = t.x; v[4 * j + 1] = t.y; v[4 * j + 2] = t.z; v[4 * j + 3] = t.w;
          m = fmaxf(m, fmaxf(fmaxf(t.x, t.y), fmaxf(t.z, t.w)));
        }
      }
    } else {
#pragma unroll
      for (int j = 0; j < ROW_REGS; ++j) {
        int64_t i = lane + 32 * j;
        if (i < cols) { v[j] = p[i]; m = fmaxf(m, v[j]); }
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    float s = 0.f;
    const float ml2 = m * 1.4426950408889634f;
    if (vec) {
      const int64_t c4 = cols >> 2;
#pragma unroll
      for (int j = 0; j < ROW_REGS / 4; ++j) {
        if (lane + 32 * j < c4) {
#pragma unroll
          for (int k = 0; k < 4; ++k) { v[4 * j + k] = exp_sub(v[4 * j + k], ml2); s += v[4 * j + k]; }
        }
      }
    } else {
#pragma unroll
      for (int j = 0; j < ROW_REGS; ++j)
        if (lane + 32 * j < cols) { v[j] = exp_sub(v[j], ml2); s += v[j]; }
    }
    s = warp_sum(s);
    if (MODE == MODE_LSE) {
      if (lane == 0) out[row] = m + logf(s);
      continue;
    }
    const float inv = 1.f / s;
    float* __restrict__ q = out + row * cols;
    float loss = 0.f;
    const float logs = logf(s);
    if (vec) {
      float4* __restrict__ q4 = reinterpret_cast<float4*>(q);
      const int64_t c4 = cols >> 2;
#pragma unroll
      for (int j = 0; j < ROW_REGS / 4; ++j) {
        int64_t i = lane + 32 * j;
        if (i < c4) {
          float4 t = make_float4(v[4 * j] * inv, v[4 * j + 1] * inv, v[4 * j + 2] * inv, v[4 * j + 3] * inv);
          q4[i] = t;
          if (MODE == MODE_CE && onehot) {
            float4 y = __ldg(reinterpret_cast<const float4*>(onehot + row * cols) + i);
            // log p = log(e) - log(s), with e = exp(x - m) kept in v
            if (y.x != 0.f) loss -= y.x * (logf(v[4 * j]) - logs);
            if (y.y != 0.f) loss -= y.y * (logf(v[4 * j + 1]) - logs);
            if (y.z != 0.f) loss -= y.z * (logf(v[4 * j + 2]) - logs);
            if (y.w != 0.f) loss -= y.w * (logf(v[4 * j + 3]) - logs);
          }
        }
      }
    } else {
#pragma unroll
      for (int j = 0; j < ROW_REGS; ++j) {
        int64_t i = lane + 32 * j;
        if (i < cols) {
          q[i] = v[j] * inv;
          if (MODE == MODE_CE && onehot) {
            float y = __ldg(onehot + row * cols + i);
            if (y != 0.f) loss -= y * (logf(v[j]) - logs);
          }
        }
      }
    }
    if (MODE == MODE_CE) {
      if (labels) {
        // -log p[label] = log(s) - (x[label] - m)
        if (lane == 0) {
          int32_t y = labels[row];
          rowloss[row] = (y >= 0 && y < cols) ? (logs - (p[y] - m)) : 0.f;
        }
      } else {
        loss = warp_sum(loss);
        if (lane == 0) rowloss[row] = loss;
      }
    }
  }
}

// One CTA per row for wide rows: three passes over the row (max, sum-exp, write); the
// row is re-read from L2, not HBM, for any width that matters here.
template <int MODE>
__global__ void __launch_bounds__(SM_THREADS) softmax_block_kernel(const float* __restrict__ x, float* __restrict__ out,
                                                                   const int32_t* __restrict__ labels,
                                                                   const float* __restrict__ onehot,
                                                                   float* __restrict__ rowloss, int64_t cols) {
  mc::pdl_sync();
  __shared__ float smem[32];
  const int64_t row = blockIdx.x;
  const float* __restrict__ p = x + row * cols;
  float m = -INFINITY;
  for (int64_t i = threadIdx.x; i < cols; i += SM_THREADS) m = fmaxf(m, p[i]);
  m = block_reduce(m, MaxOp(), smem);
  float s = 0.f;
  const float ml2 = m * 1.4426950408889634f;
  for (int64_t i = threadIdx.x; i < cols; i += SM_THREADS) s += exp_sub(p[i], ml2);
  s = block_reduce(s, SumOp(), smem);
  if (MODE == MODE_LSE) {
    if (threadIdx.x == 0) out[row] = m + logf(s);
    return;
  }
  const float inv = 1.f / s, logs = logf(s);
  float* __restrict__ q = out + row * cols;
  float loss = 0.f;
  for (int64_t i = threadIdx.x; i < cols; i += SM_THREADS) {
    float d = p[i] - m;
    q[i] = exp_sub(p[i], ml2) * inv;
    if (MODE == MODE_CE && onehot) {
      float y = __ldg(onehot + row * cols + i);
      if (y != 0.f) loss -= y * (d - logs);
    }
  }
  if (MODE == MODE_CE) {
    if (labels) {
      if (threadIdx.x == 0) {
        int32_t y = labels[row];
        rowloss[row] = (y >= 0 && y < cols) ? (logs - (p[y] - m)) : 0.f;
      }
    } else {
      loss = block_reduce(loss, SumOp(), smem);
      if (threadIdx.x == 0) rowloss[row] = loss;
    }
  }
}

template <int MODE>
static int launch_softmax(const float* x, float* out, const int32_t* labels, const float* onehot, float* rowloss,
                          int64_t rows, int64_t cols) {
  if (cols <= 32 * ROW_REGS) {
    int grid = grid_for(rows, SM_THREADS / 32);
    mc::launch(softmax_warp_kernel<MODE>, grid, SM_THREADS, 0, x, out, labels, onehot, rowloss, rows, cols);
  } else {
    MC_REQUIRE(rows < (int64_t)2147483647, "too many rows");
    mc::launch(softmax_block_kernel<MODE>, (unsigned)rows, SM_THREADS, 0, x, out, labels, onehot, rowloss, cols);
  }
  MC_LAUNCHED();
  return 0;
}

__global__ void __launch_bounds__(SM_THREADS) scale_scalar_kernel(float* __restrict__ v, float s) {
  mc::pdl_sync();
  if (threadIdx.x == 0 && blockIdx.x == 0) v[0] *= s;
}

__global__ void __launch_bounds__(SM_THREADS) onehot_kernel(const int32_t* __restrict__ idx, float* __restrict__ out,
                                                            int64_t rows, int64_t cols) {
  mc::pdl_sync();
  const int64_t stride = (int64_t)gridDim.x * SM_THREADS;
  for (int64_t r = (int64_t)blockIdx.x * SM_THREADS + threadIdx.x; r < rows; r += stride) {
    int32_t c = idx[r];
    if (c < 0) c += (int32_t)cols;  // NumPy negative indexing
    if (c >= 0 && c < cols) out[r * cols + c] = 1.f;
  }
}

__global__ void __launch_bounds__(SM_THREADS) softmax_ce_bwd_kernel(const float* __restrict__ probs,
                                                                    const int32_t* __restrict__ labels,
                                                                    const float* __restrict__ onehot,
                                                                    float* __restrict__ dx, int64_t rows, int64_t cols,
                                                                    float inv_rows) {
  mc::pdl_sync();
  const int64_t n = rows * cols;
  const int64_t stride = (int64_t)gridDim.x * SM_THREADS;
  for (int64_t t = (int64_t)blockIdx.x * SM_THREADS + threadIdx.x; t < n; t += stride) {
    float y;
    if (labels) {
      const int64_t r = t / cols, c = t % cols;
      y = (labels[r] == (int32_t)c) ? 1.f : 0.f;
    } else {
      y = onehot[t];
    }
    dx[t] = (probs[t] - y) * inv_rows;
  }
}

// softmax_loss for a classifier head (cols <= 32, integer labels): ONE launch does everything the three
// launches of mc_softmax_ce_fwd plus mc_softmax_ce_bwd do -- probabilities, the mean cross-entropy and, because
// the VJP of the reference's softmax_output ignores its upstream gradient (numpy_core.py:170-175), the
// gradient (p - onehot) / rows itself.  One warp per row; the block that finishes last (ticket counter, reset
// for the next call) adds the row losses in a fixed order, so the mean is deterministic.
// (First version: a single 1024-thread block looping over 32 rows per warp -- 45 us for 1024 x 10, every
// iteration a dependent global-load round trip; profiles/r2d.)
constexpr int CE_THREADS = 256;
__global__ void __launch_bounds__(CE_THREADS) softmax_ce_small_kernel(const float* __restrict__ x, const int32_t* __restrict__ labels,
                                                                      float* __restrict__ probs, float* __restrict__ dx,
                                                                      float* __restrict__ loss, float* __restrict__ rowloss,
                                                                      int* __restrict__ counter, int rows, int cols,
                                                                      float inv_rows) {
  mc::pdl_sync();
  __shared__ float red[32];
  __shared__ int s_last;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int row = blockIdx.x * (CE_THREADS / 32) + warp;
  if (row < rows) {
    const int64_t base = (int64_t)row * cols;
    const float v = lane < cols ? x[base + lane] : -INFINITY;
    const int lab = labels[row];
    const float m = warp_max(v);
    const float e = lane < cols ? expf(v - m) : 0.f;
    const float s = warp_sum(e);
    const float pr = e / s;
    if (lane < cols) {
      if (probs) probs[base + lane] = pr;
      if (dx) dx[base + lane] = (pr - (lane == lab ? 1.f : 0.f)) * inv_rows;
    }
    const float pl = __shfl_sync(0xffffffffu, pr, lab & 31);
    if (lane == 0) rowloss[row] = -logf(pl);
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const int ticket = atomicAdd(counter, 1);
    s_last = (ticket == (int)gridDim.x - 1) ? 1 : 0;
    if (s_last) *counter = 0;
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  float acc = 0.f;
#pragma unroll 8
  for (int r = threadIdx.x; r < rows; r += CE_THREADS) acc += __ldcg(rowloss + r);
  acc = block_reduce(acc, SumOp(), red);
  if (threadIdx.x == 0) *loss = acc * inv_rows;
}

// The classifier head in ONE launch: logits = h . W + b (nn/layers.py:27), softmax_loss and its gradient
// dlogits = (p - onehot) / rows (layers.py:201-221; numpy_core.py:170-175), and the gradient with respect to the
// head's input, dh = dlogits . W^T -- what affine + softmax_loss + the two VJPs compute in three launches on the
// chain (skinny GEMM, loss kernel, skinny GEMM).  One warp per row; W (K x C, C <= 16) lives in shared memory;
// FP32 throughout, like the skinny kernels it replaces.  The block that finishes last sums the row losses.
constexpr int HEAD_ROWS = 8;      // rows per block
constexpr int HEAD_RPW = 4;       // rows a warp carries through its chunks of K
constexpr int HEAD_RG = HEAD_ROWS / HEAD_RPW;
constexpr int HEAD_SLOTS = 8;     // warps that split K
constexpr int HEAD_THREADS = 32 * HEAD_RG * HEAD_SLOTS;
constexpr int HEAD_MAXC = 16;
MC_PROBE_DEFINE(head)   // tools/probe.sh: %globaltimer stamps at the phase boundaries of head_softmax_kernel
#define HEAD_STAMP(i) MC_STAMP(head, i)
// One row of W (C floats at w_s + f * C) into registers with the widest shared-memory loads its alignment allows.
template <int C>
__device__ __forceinline__ void head_load_w(const float* __restrict__ wr, float (&w)[C]) {
  if constexpr (C % 4 == 0) {
#pragma unroll
    for (int c = 0; c < C; c += 4) {
      const float4 v = *reinterpret_cast<const float4*>(wr + c);
      w[c] = v.x, w[c + 1] = v.y, w[c + 2] = v.z, w[c + 3] = v.w;
    }
  } else if constexpr (C % 2 == 0) {
#pragma unroll
    for (int c = 0; c < C; c += 2) {
      const float2 v = *reinterpret_cast<const float2*>(wr + c);
      w[c] = v.x, w[c + 1] = v.y;
    }
  } else {
#pragma unroll
    for (int c = 0; c < C; ++c) w[c] = wr[c];
  }
}

// Sum v[0..N) over the 32 lanes by halving: at each step a lane keeps one half of its values and trades the other
// half with its partner, so N values cost about N shuffles instead of 5 N.  Afterwards the lane holds, in v[0..R),
// the totals of the original indices base + i, for those below `limit` (R = N halved five times, rounding up).
template <int N, int MASK>
__device__ __forceinline__ void head_fold(float* v, int lane, int& base, int& limit) {
  if constexpr (MASK >= 1) {
    constexpr int H = (N + 1) / 2;
    const bool upper = (lane & MASK) != 0;
#pragma unroll
    for (int i = 0; i < H; ++i) {
      const float lo = v[i];
      const float hi = (i + H < N) ? v[i + H] : 0.f;
      const float send = upper ? lo : hi;
      const float keep = upper ? hi : lo;
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, MASK);
    }
    if (upper) base += H;
    else limit = min(limit, base + H);
    head_fold<H, MASK / 2>(v, lane, base, limit);
  }
}
constexpr int head_fold_left(int n) {
  for (int s = 0; s < 5; ++s) n = (n + 1) / 2;
  return n;
}

// Layout of the work (profiles/r2l has the in-kernel phase timings that led here).  The first version gave each
// row one warp and a run-time class count: every W element was a predicated scalar LDS feeding one FMA, and each
// row re-read all of W from shared memory, twice (logits, then dh) -- 12.5 us at 1024 x 512 x 10, 27 us at K = 2048,
// bound by LDS latency and then by shared-memory bandwidth.  Now: C is a template parameter (LDS.64 / LDS.128 rows
// of W, no predicates); a warp owns 32-element chunks of K for FOUR rows at a time, so a W row read from shared
// memory feeds 4 x C FMAs; the per-lane partial logits are folded across lanes by halving (head_fold), the eight
// K-slots are summed in a fixed order by the warp that owns the row, which also does the softmax with one class
// per lane; thread 0 takes the ticket while the other warps write dh.
template <int C>
__global__ void __launch_bounds__(HEAD_THREADS) head_softmax_kernel(const float* __restrict__ h, const float* __restrict__ W,
                                                                    const float* __restrict__ bias, const int32_t* __restrict__ labels,
                                                                    float* __restrict__ logits, float* __restrict__ dlogits,
                                                                    float* __restrict__ dh, float* __restrict__ loss,
                                                                    float* __restrict__ blockloss, int* __restrict__ counter, int rows,
                                                                    int K, float inv_rows) {
  extern __shared__ __align__(16) float w_s[];            // K * C
  __shared__ __align__(16) float part[HEAD_SLOTS][HEAD_ROWS * HEAD_MAXC];
  __shared__ __align__(16) float dls[HEAD_ROWS][HEAD_MAXC];
  __shared__ float rl[HEAD_ROWS];
  __shared__ int s_last;
  HEAD_STAMP(0);
  mc::pdl_sync();
  HEAD_STAMP(1);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int rg = warp % HEAD_RG, slot = warp / HEAD_RG;   // group of four rows, K-slot
  const int row0 = blockIdx.x * HEAD_ROWS + rg * HEAD_RPW;
  constexpr int PASS_CHUNKS = 4;                          // chunks of 32 a warp handles per pass
  // what the softmax warps need later, fetched now
  const int my_row = blockIdx.x * HEAD_ROWS + warp;
  const bool owner = warp < HEAD_ROWS && my_row < rows;
  const int lab = owner ? labels[my_row] : 0;
  const float my_bias = (owner && bias && lane < C) ? __ldg(bias + lane) : 0.f;

  auto load_h = [&](int pass, float (&hv)[HEAD_RPW * PASS_CHUNKS]) {
#pragma unroll
    for (int j = 0; j < HEAD_RPW; ++j) {
#pragma unroll
      for (int i = 0; i < PASS_CHUNKS; ++i) {
        const int f = ((pass * PASS_CHUNKS + i) * HEAD_SLOTS + slot) * 32 + lane;
        hv[j * PASS_CHUNKS + i] = (row0 + j < rows && f < K) ? __ldg(h + (int64_t)(row0 + j) * K + f) : 0.f;
      }
    }
  };
  float hv[HEAD_RPW * PASS_CHUNKS];
  load_h(0, hv);
  {
    const int n = K * C;
    if ((n & 3) == 0 && ((reinterpret_cast<uintptr_t>(W) & 15u) == 0)) {
      const float4* __restrict__ W4 = reinterpret_cast<const float4*>(W);
      float4* w4 = reinterpret_cast<float4*>(w_s);
#pragma unroll 4
      for (int i = threadIdx.x; i < (n >> 2); i += HEAD_THREADS) w4[i] = __ldg(W4 + i);
    } else {
#pragma unroll 4
      for (int i = threadIdx.x; i < n; i += HEAD_THREADS) w_s[i] = __ldg(W + i);
    }
  }
  __syncthreads();
  HEAD_STAMP(2);
  float acc[HEAD_RPW * C];
#pragma unroll
  for (int i = 0; i < HEAD_RPW * C; ++i) acc[i] = 0.f;
  const int npass = (K + 32 * HEAD_SLOTS * PASS_CHUNKS - 1) / (32 * HEAD_SLOTS * PASS_CHUNKS);
  for (int pass = 0; pass < npass; ++pass) {
    float hn[HEAD_RPW * PASS_CHUNKS];
    if (pass + 1 < npass) load_h(pass + 1, hn);          // the next pass's slice of h is in flight during this one
#pragma unroll
    for (int i = 0; i < PASS_CHUNKS; ++i) {
      const int f = ((pass * PASS_CHUNKS + i) * HEAD_SLOTS + slot) * 32 + lane;
      if (f - lane >= K) break;
      float w[C];
      if (f - lane + 32 <= K) {
        head_load_w<C>(w_s + f * C, w);
      } else {
#pragma unroll
        for (int c = 0; c < C; ++c) w[c] = (f < K) ? w_s[f * C + c] : 0.f;
      }
#pragma unroll
      for (int j = 0; j < HEAD_RPW; ++j) {
#pragma unroll
        for (int c = 0; c < C; ++c) acc[j * C + c] = fmaf(hv[j * PASS_CHUNKS + i], w[c], acc[j * C + c]);
      }
    }
    if (pass + 1 < npass) {
#pragma unroll
      for (int q = 0; q < HEAD_RPW * PASS_CHUNKS; ++q) hv[q] = hn[q];
    }
  }
  {
    int base = 0, limit = HEAD_RPW * C;
    head_fold<HEAD_RPW * C, 16>(acc, lane, base, limit);
#pragma unroll
    for (int i = 0; i < head_fold_left(HEAD_RPW * C); ++i) {
      const int idx = base + i;                           // (row in group) * C + class
      if (idx < limit) part[slot][(rg * HEAD_RPW + idx / C) * HEAD_MAXC + idx % C] = acc[i];
    }
  }
  __syncthreads();
  if (warp < HEAD_ROWS) {                                 // warp r finishes row r of the block: one class per lane
    if (owner) {
      float t = -INFINITY;
      if (lane < C) {
        t = part[0][warp * HEAD_MAXC + lane];
#pragma unroll
        for (int sl = 1; sl < HEAD_SLOTS; ++sl) t += part[sl][warp * HEAD_MAXC + lane];
        t += my_bias;
      }
      const float m = warp_max(t);
      const float ex = (lane < C) ? expf(t - m) : 0.f;
      const float ssum = warp_sum(ex);
      const float pr = ex / ssum;
      const float d = (pr - (lane == lab ? 1.f : 0.f)) * inv_rows;
      const float tl = __shfl_sync(0xffffffffu, t, lab & 31);
      const float lab_logit = (lab >= 0 && lab < C) ? tl : 0.f;
      if (lane < C) {
        if (logits) logits[(int64_t)my_row * C + lane] = t;
        dlogits[(int64_t)my_row * C + lane] = d;
        dls[warp][lane] = d;
      }
      if (lane == 0) rl[warp] = -(lab_logit - m - logf(ssum));
    } else if (lane == 0) {
      rl[warp] = 0.f;
    }
  }
  __syncthreads();
  HEAD_STAMP(3);
  if (threadIdx.x == 0) {                                // the block's loss and the ticket, while the others write dh
    float t = 0.f;
#pragma unroll
    for (int r = 0; r < HEAD_ROWS; ++r) t += rl[r];
    blockloss[blockIdx.x] = t;
    __threadfence();
    const int ticket = atomicAdd(counter, 1);
    s_last = (ticket == (int)gridDim.x - 1) ? 1 : 0;
    if (s_last) *counter = 0;
  }
  if (dh && row0 < rows) {
    float dl[HEAD_RPW][C];
#pragma unroll
    for (int j = 0; j < HEAD_RPW; ++j) {
      if (row0 + j < rows) {
        head_load_w<C>(&dls[rg * HEAD_RPW + j][0], dl[j]);
      } else {
#pragma unroll
        for (int c = 0; c < C; ++c) dl[j][c] = 0.f;
      }
    }
#pragma unroll 2
    for (int f = slot * 32 + lane; f < K; f += 32 * HEAD_SLOTS) {
      float w[C];
      head_load_w<C>(w_s + f * C, w);
#pragma unroll
      for (int j = 0; j < HEAD_RPW; ++j) {
        float v0 = 0.f, v1 = 0.f;
#pragma unroll
        for (int c = 0; c < C; ++c) {
          if (c & 1) v1 = fmaf(dl[j][c], w[c], v1);
          else v0 = fmaf(dl[j][c], w[c], v0);
        }
        if (row0 + j < rows) dh[(int64_t)(row0 + j) * K + f] = v0 + v1;
      }
    }
  }
  HEAD_STAMP(4);
  __syncthreads();
  HEAD_STAMP(5);
  HEAD_STAMP(6);
  if (!s_last || warp != 0) return;
  __threadfence();
  float a = 0.f;
  for (int b = lane; b < (int)gridDim.x; b += 32) a += __ldcg(blockloss + b);
  a = warp_sum(a);
  if (lane == 0) *loss = a * inv_rows;
  HEAD_STAMP(7);
}

// temporal_softmax_loss (nn/layers.py:444-480) over rows = N*T score vectors: one warp per row computes the
// softmax, the masked cross-entropy term  -mask * log p[y] / N  and the gradient  mask * (p - onehot) / N.
// The reference composes it from ~12 primitives (30 launches forward + backward on the device).
__global__ void __launch_bounds__(SM_THREADS) temporal_ce_kernel(const float* __restrict__ x, const int32_t* __restrict__ y,
                                                                 const float* __restrict__ mask, float* __restrict__ dx,
                                                                 float* __restrict__ rowloss, int64_t rows, int64_t cols,
                                                                 float inv_n) {
  mc::pdl_sync();
  const int lane = threadIdx.x & 31;
  const int64_t warps_per_grid = (int64_t)gridDim.x * (SM_THREADS / 32);
  for (int64_t row = (int64_t)blockIdx.x * (SM_THREADS / 32) + (threadIdx.x >> 5); row < rows; row += warps_per_grid) {
    const float* __restrict__ xr = x + row * cols;
    float m = -INFINITY;
    for (int64_t c = lane; c < cols; c += 32) m = fmaxf(m, xr[c]);
    m = warp_max(m);
    float s = 0.f;
    for (int64_t c = lane; c < cols; c += 32) s += expf(xr[c] - m);
    s = warp_sum(s);
    const float w = mask ? mask[row] : 1.f;
    int32_t lab = y[row];
    if (lab < 0) lab += (int32_t)cols;
    float* __restrict__ dr = dx + row * cols;
    const float scale = w * inv_n;
    for (int64_t c = lane; c < cols; c += 32) {
      const float pr = expf(xr[c] - m) / s;
      dr[c] = (pr - (c == lab ? 1.f : 0.f)) * scale;
    }
    if (lane == 0) rowloss[row] = (w != 0.f) ? -(xr[lab] - m - logf(s)) * scale : 0.f;
  }
}

}  // namespace mc

using namespace mc;

extern "C" {

int mc_softmax_rows(const float* x, float* out, int64_t rows, int64_t cols) {
  if (rows * cols <= 0) return 0;
  return launch_softmax<MODE_SOFTMAX>(x, out, nullptr, nullptr, nullptr, rows, cols);
}

int mc_logsumexp_rows(const float* x, float* out, int64_t rows, int64_t cols) {
  if (rows <= 0) return 0;
  MC_REQUIRE(cols > 0, "logsumexp of empty rows");
  return launch_softmax<MODE_LSE>(x, out, nullptr, nullptr, nullptr, rows, cols);
}

int mc_onehot(const int32_t* idx, float* out, int64_t rows, int64_t cols) {
  if (rows <= 0) return 0;
  mc::launch(onehot_kernel, grid_for(rows, SM_THREADS), SM_THREADS, 0, idx, out, rows, cols);
  MC_LAUNCHED();
  return 0;
}

int mc_softmax_ce_fwd(const float* x, const int32_t* labels, const float* onehot, float* probs, float* loss,
                      int64_t rows, int64_t cols) {
  MC_REQUIRE((labels != nullptr) != (onehot != nullptr), "exactly one of labels / onehot");
  MC_REQUIRE(rows > 0 && cols > 0, "empty input");
  ScratchFrame scratch;
  float* rowloss = scratch.floats((size_t)rows);
  if (!rowloss) return 1;
  int rc = launch_softmax<MODE_CE>(x, probs, labels, onehot, rowloss, rows, cols);
  if (rc == 0) rc = reduce_sum_internal(rowloss, loss, 1, rows, 1);
  if (rc == 0) {
    mc::launch(scale_scalar_kernel, 1, 32, 0, loss, 1.f / (float)rows);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { mc::set_error("scale launch failed: %s", cudaGetErrorString(e)); rc = 1; }
    else mc::count_launch();
  }
  return rc;
}

static int* ticket_counter() {
  static int* counters[16] = {};             // one ticket counter per device, zero between calls
  int*& counter = counters[device() & 15];
  if (!counter) {
    if (cudaMalloc((void**)&counter, sizeof(int)) != cudaSuccess) { counter = nullptr; return nullptr; }
    cudaMemsetAsync(counter, 0, sizeof(int), stream());
  }
  return counter;
}

int mc_softmax_ce_fused(const float* x, const int32_t* labels, const float* onehot, float* probs, float* dx, float* loss,
                        int64_t rows, int64_t cols) {
  MC_REQUIRE((labels != nullptr) != (onehot != nullptr), "exactly one of labels / onehot");
  MC_REQUIRE(rows > 0 && cols > 0, "empty input");
  MC_REQUIRE(probs != nullptr || dx != nullptr, "nothing to write");
  if (labels && cols <= 32 && rows <= (1 << 20)) {
    int* counter = ticket_counter();
    MC_REQUIRE(counter != nullptr, "ticket counter allocation failed");
    ScratchFrame scratch;
    float* rowloss = scratch.floats((size_t)rows);
    if (!rowloss) return 1;
    const unsigned blocks = (unsigned)((rows + CE_THREADS / 32 - 1) / (CE_THREADS / 32));
    mc::launch(softmax_ce_small_kernel, blocks, CE_THREADS, 0, x, labels, probs, dx, loss, rowloss, counter, (int)rows, (int)cols,
               1.f / (float)rows);
    MC_LAUNCHED();
    return 0;
  }
  // general shapes: the multi-launch forward, then the gradient from the probabilities
  float* p = probs ? probs : dx;
  if (mc_softmax_ce_fwd(x, labels, onehot, p, loss, rows, cols)) return 1;
  if (dx) return mc_softmax_ce_bwd(p, labels, onehot, dx, rows, cols);
  return 0;
}

// ---- softmax_loss with no allocation at all -------------------------------------------------------------
// The row losses and the ticket counter of softmax_ce_small_kernel live in a file-scope device array instead of
// the scratch arena / a cudaMalloc'd counter, for callers that cannot allocate (graph capture, serving loops).
// Every GPU that loads the library holds these 257 KB until exit.  Layout: [0, MAX_ROWS) row losses, then one
// 256-byte line whose first word is the ticket counter.
static __device__ __align__(256) float g_ce_static_scratch[MC_SOFTMAX_CE_STATIC_MAX_ROWS + 64];

// The array's name is a host-side stand-in in host code; the device address comes from cudaGetSymbolAddress,
// asked once per device with that device current.
static float* ce_static_scratch_ptr() {
  static float* ptrs[16] = {};
  const int dev = device();
  if (dev < 0 || dev >= 16) { mc::set_error("mc_init has not selected a device"); return nullptr; }
  if (!ptrs[dev]) {
    int current = -1;
    cudaError_t e = cudaGetDevice(&current);
    if (e == cudaSuccess && current != dev) e = cudaSetDevice(dev);
    void* p = nullptr;
    if (e == cudaSuccess) e = cudaGetSymbolAddress(&p, g_ce_static_scratch);
    if (e != cudaSuccess || !p) {
      mc::set_error("cudaGetSymbolAddress(g_ce_static_scratch) on device %d: %s", dev, cudaGetErrorString(e));
      cudaGetLastError();
      return nullptr;
    }
    ptrs[dev] = static_cast<float*>(p);
  }
  return ptrs[dev];
}

int mc_softmax_ce_fused_static(const float* x, const int32_t* labels, float* probs, float* dx, float* loss, int64_t rows,
                               int64_t cols) {
  MC_REQUIRE(x != nullptr && labels != nullptr && loss != nullptr, "x, labels and loss are required");
  MC_REQUIRE(probs != nullptr || dx != nullptr, "nothing to write");
  MC_REQUIRE(rows > 0 && cols > 0, "empty input");
  MC_REQUIRE(cols <= 32, "more than 32 classes: beyond the static-scratch variant (use mc_softmax_ce_fused)");
  MC_REQUIRE(rows <= MC_SOFTMAX_CE_STATIC_MAX_ROWS,
             "rows beyond MC_SOFTMAX_CE_STATIC_MAX_ROWS (65536): the static scratch cannot hold them (use mc_softmax_ce_fused)");
  float* scratch = ce_static_scratch_ptr();
  if (!scratch) return 1;
  float* rowloss = scratch;
  int* counter = reinterpret_cast<int*>(scratch + MC_SOFTMAX_CE_STATIC_MAX_ROWS);
  // the array keeps the last call's contents: the ticket counter must start at zero (a memset node, no allocation);
  // every row loss the last block reads is rewritten by this call
  MC_CHECK(cudaMemsetAsync(counter, 0, sizeof(int), stream()));
  const unsigned blocks = (unsigned)((rows + CE_THREADS / 32 - 1) / (CE_THREADS / 32));
  mc::launch(softmax_ce_small_kernel, blocks, CE_THREADS, 0, x, labels, probs, dx, loss, rowloss, counter, (int)rows, (int)cols,
             1.f / (float)rows);
  MC_LAUNCHED();
  return 0;
}

int mc_temporal_softmax_loss(const float* x, const int32_t* y, const float* mask, float* dx, float* loss, int64_t rows,
                             int64_t cols, int64_t batch) {
  MC_REQUIRE(rows > 0 && cols > 0 && batch > 0, "empty input");
  ScratchFrame scratch;
  float* rowloss = scratch.floats((size_t)rows);
  if (!rowloss) return 1;
  mc::launch(temporal_ce_kernel, grid_for(rows * 32, SM_THREADS), SM_THREADS, 0, x, y, mask, dx, rowloss, rows, cols,
             1.f / (float)batch);
  MC_LAUNCHED();
  return reduce_sum_internal(rowloss, loss, 1, rows, 1);
}

int mc_head_softmax_loss(const float* h, const float* W, const float* bias, const int32_t* labels, float* logits,
                         float* dlogits, float* dh, float* loss, int64_t rows, int64_t K, int64_t C) {
  MC_REQUIRE(rows > 0 && K > 0 && C > 0 && C <= HEAD_MAXC, "classifier head: 1 <= classes <= 16");
  const size_t smem = sizeof(float) * (size_t)K * C;
  MC_REQUIRE(smem <= 200 * 1024, "classifier head: weight matrix beyond 200 KB of shared memory");
  static uint32_t configured[HEAD_MAXC + 1] = {};
  int* counter = ticket_counter();
  MC_REQUIRE(counter != nullptr, "ticket counter allocation failed");
  ScratchFrame scratch;
  const unsigned blocks = (unsigned)((rows + HEAD_ROWS - 1) / HEAD_ROWS);
  float* rowloss = scratch.floats((size_t)blocks);      // one partial loss per block
  if (!rowloss) return 1;
  const float inv_rows = 1.f / (float)rows;
  switch ((int)C) {
#define HEAD_CASE(CC)                                                                                                  \
  case CC:                                                                                                             \
    if (first_use_on_device(&configured[CC]))                                                                          \
      MC_CHECK(cudaFuncSetAttribute(head_softmax_kernel<CC>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024)); \
    mc::launch(head_softmax_kernel<CC>, blocks, HEAD_THREADS, smem, h, W, bias, labels, logits, dlogits, dh, loss,     \
               rowloss, counter, (int)rows, (int)K, inv_rows);                                                         \
    break;
    HEAD_CASE(1) HEAD_CASE(2) HEAD_CASE(3) HEAD_CASE(4) HEAD_CASE(5) HEAD_CASE(6) HEAD_CASE(7) HEAD_CASE(8)
    HEAD_CASE(9) HEAD_CASE(10) HEAD_CASE(11) HEAD_CASE(12) HEAD_CASE(13) HEAD_CASE(14) HEAD_CASE(15) HEAD_CASE(16)
#undef HEAD_CASE
    default: MC_REQUIRE(false, "classifier head: 1 <= classes <= 16");
  }
  MC_LAUNCHED();
  return 0;
}

int mc_softmax_ce_bwd(const float* probs, const int32_t* labels, const float* onehot, float* dx, int64_t rows,
                      int64_t cols) {
  MC_REQUIRE((labels != nullptr) != (onehot != nullptr), "exactly one of labels / onehot");
  if (rows * cols <= 0) return 0;
  mc::launch(softmax_ce_bwd_kernel, grid_for(rows * cols, SM_THREADS), SM_THREADS, 0, probs, labels, onehot, dx, rows,
                                                                                       cols, 1.f / (float)rows);
  MC_LAUNCHED();
  return 0;
}

}  // extern "C"
