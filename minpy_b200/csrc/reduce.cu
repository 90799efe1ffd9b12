// Axis reductions (sum / max / min / argmax) and their VJP helpers.
//
// x is viewed as (outer, red, inner) and the middle axis is reduced.  HBM-bound:
// 4(outer*red*inner + outer*inner) bytes.  Two kernels cover every case:
//   * inner == 1  -> "row" reduction: one warp or one CTA per (row, split) segment,
//     float4 loads, warp-shuffle then shared-memory tree;
//   * inner  > 1  -> "column" reduction: threads map to the contiguous inner axis
//     (coalesced, float4 = 4 columns per thread), 8 row-lanes per CTA combined in
//     shared memory.
// When the natural grid would leave SMs idle the reduced axis is split across CTAs
// into a workspace and combined by a second (deterministic) pass -- no atomics, so
// results are reproducible run to run.
#include "common.cuh"

namespace mc {

constexpr int RED_THREADS = 256;

template <class Op>
__global__ void __launch_bounds__(RED_THREADS) reduce_rows_block_kernel(const float* __restrict__ x,
                                                                        float* __restrict__ out, int64_t red,
                                                                        int splits, int64_t seg) {
  mc::pdl_sync();
  // block b handles row = b / splits, segment s = b % splits: [s*seg, min(red,(s+1)*seg))
  __shared__ float smem[32];
  Op op;
  const int64_t row = blockIdx.x / splits;
  const int s = blockIdx.x % splits;
  const int64_t begin = (int64_t)s * seg;
  int64_t end = begin + seg;
  if (end > red) end = red;
  const float* __restrict__ p = x + row * red;
  float acc = Op::identity();
  const bool vec = ((reinterpret_cast<uintptr_t>(p + begin) & 15u) == 0);
  if (vec) {
    const int64_t n4 = (end - begin) >> 2;
    const float4* __restrict__ p4 = reinterpret_cast<const float4*>(p + begin);
    int64_t i = threadIdx.x;
    // 4 independent loads in flight per thread
    for (; i + 3 * RED_THREADS < n4; i += 4 * RED_THREADS) {
      float4 a = p4[i], b = p4[i + RED_THREADS], c = p4[i + 2 * RED_THREADS], d = p4[i + 3 * RED_THREADS];
      acc = op(acc, op(op(op(a.x, a.y), op(a.z, a.w)), op(op(b.x, b.y), op(b.z, b.w))));
      acc = op(acc, op(op(op(c.x, c.y), op(c.z, c.w)), op(op(d.x, d.y), op(d.z, d.w))));
    }
    for (; i < n4; i += RED_THREADS) {
      float4 a = p4[i];
      acc = op(acc, op(op(a.x, a.y), op(a.z, a.w)));
    }
    for (int64_t t = begin + (n4 << 2) + threadIdx.x; t < end; t += RED_THREADS) acc = op(acc, p[t]);
  } else {
    for (int64_t t = begin + threadIdx.x; t < end; t += RED_THREADS) acc = op(acc, p[t]);
  }
  acc = block_reduce(acc, op, smem);
  if (threadIdx.x == 0) out[blockIdx.x] = acc;
}

// One warp per row, for many short rows.
template <class Op>
__global__ void __launch_bounds__(RED_THREADS) reduce_rows_warp_kernel(const float* __restrict__ x,
                                                                       float* __restrict__ out, int64_t rows,
                                                                       int64_t red) {
  mc::pdl_sync();
  Op op;
  const int lane = threadIdx.x & 31;
  const int64_t warps_per_grid = (int64_t)gridDim.x * (RED_THREADS / 32);
  for (int64_t row = (int64_t)blockIdx.x * (RED_THREADS / 32) + (threadIdx.x >> 5); row < rows; row += warps_per_grid) {
    const float* __restrict__ p = x + row * red;
    float acc = Op::identity();
    if (((reinterpret_cast<uintptr_t>(p) & 15u) == 0) && (red & 3) == 0) {
      const float4* __restrict__ p4 = reinterpret_cast<const float4*>(p);
      for (int64_t i = lane; i < (red >> 2); i += 32) {
        float4 a = p4[i];
        acc = op(acc, op(op(a.x, a.y), op(a.z, a.w)));
      }
    } else {
      for (int64_t i = lane; i < red; i += 32) acc = op(acc, p[i]);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc = op(acc, __shfl_xor_sync(0xffffffffu, acc, o));
    if (lane == 0) out[row] = acc;
  }
}

// Small column reductions (the bias gradient of a (1024, 512) layer is 2 MB) in ONE launch: a block
// owns 8 columns (one 32-byte sector per row), thread = (row lane 0..127, column half); four rows
// in flight per thread; 16 row lanes combine by shuffle, the 8 warps through shared memory, in a
// fixed order.  The split + second-pass path below costs two launches for ~1 us of memory time.
template <class Op>
__global__ void __launch_bounds__(RED_THREADS) reduce_cols_slab_kernel(const float* __restrict__ x, float* __restrict__ out,
                                                                       int64_t red, int64_t inner) {
  mc::pdl_sync();
  __shared__ float sm[8][2][4];
  Op op;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, half = threadIdx.x & 1, rl = threadIdx.x >> 1;
  const int64_t col = (int64_t)blockIdx.x * 8 + 4 * half;
  const float* __restrict__ p = x + (int64_t)blockIdx.y * red * inner + col;
  float acc[4] = {Op::identity(), Op::identity(), Op::identity(), Op::identity()};
  int64_t r = rl;
  for (; r + 384 < red; r += 512) {
    const float4 a = __ldg(reinterpret_cast<const float4*>(p + r * inner));
    const float4 b = __ldg(reinterpret_cast<const float4*>(p + (r + 128) * inner));
    const float4 c = __ldg(reinterpret_cast<const float4*>(p + (r + 256) * inner));
    const float4 d = __ldg(reinterpret_cast<const float4*>(p + (r + 384) * inner));
    acc[0] = op(acc[0], op(op(a.x, b.x), op(c.x, d.x))); acc[1] = op(acc[1], op(op(a.y, b.y), op(c.y, d.y)));
    acc[2] = op(acc[2], op(op(a.z, b.z), op(c.z, d.z))); acc[3] = op(acc[3], op(op(a.w, b.w), op(c.w, d.w)));
  }
  for (; r < red; r += 128) {
    const float4 a = __ldg(reinterpret_cast<const float4*>(p + r * inner));
    acc[0] = op(acc[0], a.x); acc[1] = op(acc[1], a.y); acc[2] = op(acc[2], a.z); acc[3] = op(acc[3], a.w);
  }
#pragma unroll
  for (int k = 0; k < 4; ++k) {
#pragma unroll
    for (int o = 2; o < 32; o <<= 1) acc[k] = op(acc[k], __shfl_xor_sync(0xffffffffu, acc[k], o));
    if (lane < 2) sm[warp][half][k] = acc[k];
  }
  __syncthreads();
  if (threadIdx.x < 8) {
    const int h = threadIdx.x >> 2, k = threadIdx.x & 3;
    float v = sm[0][h][k];
#pragma unroll
    for (int w = 1; w < 8; ++w) v = op(v, sm[w][h][k]);
    out[(int64_t)blockIdx.y * inner + (int64_t)blockIdx.x * 8 + threadIdx.x] = v;
  }
}

// Column reduction: block = 32 (x) * 8 (y) threads; each x-thread owns VEC columns.
// grid = (col_tiles, splits, outer).  out index: ((o*splits + s) * inner + col).
template <class Op, int VEC>
__global__ void __launch_bounds__(RED_THREADS) reduce_cols_kernel(const float* __restrict__ x,
                                                                  float* __restrict__ out, int64_t red,
                                                                  int64_t inner, int64_t seg) {
  mc::pdl_sync();
  __shared__ float smem[8][32 * VEC + 1];
  Op op;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int64_t col = ((int64_t)blockIdx.x * 32 + tx) * VEC;
  const int64_t o = blockIdx.z;
  const int splits = gridDim.y;
  const int64_t begin = (int64_t)blockIdx.y * seg;
  int64_t end = begin + seg;
  if (end > red) end = red;
  float acc[VEC];
#pragma unroll
  for (int v = 0; v < VEC; ++v) acc[v] = Op::identity();
  if (col < inner) {
    const float* __restrict__ p = x + o * red * inner + col;
    if (VEC == 4) {
      int64_t r = begin + ty;
      for (; r + 24 < end; r += 32) {  // four independent 16-byte loads in flight per thread
        float4 a = __ldg(reinterpret_cast<const float4*>(p + r * inner));
        float4 b = __ldg(reinterpret_cast<const float4*>(p + (r + 8) * inner));
        float4 c = __ldg(reinterpret_cast<const float4*>(p + (r + 16) * inner));
        float4 d = __ldg(reinterpret_cast<const float4*>(p + (r + 24) * inner));
        acc[0] = op(acc[0], op(op(a.x, b.x), op(c.x, d.x))); acc[1] = op(acc[1], op(op(a.y, b.y), op(c.y, d.y)));
        acc[2] = op(acc[2], op(op(a.z, b.z), op(c.z, d.z))); acc[3] = op(acc[3], op(op(a.w, b.w), op(c.w, d.w)));
      }
      for (; r < end; r += 8) {
        float4 a = *reinterpret_cast<const float4*>(p + r * inner);
        acc[0] = op(acc[0], a.x); acc[1] = op(acc[1], a.y); acc[2] = op(acc[2], a.z); acc[3] = op(acc[3], a.w);
      }
    } else {
      for (int64_t r = begin + ty; r < end; r += 8) acc[0] = op(acc[0], p[r * inner]);
    }
  }
#pragma unroll
  for (int v = 0; v < VEC; ++v) smem[ty][tx * VEC + v] = acc[v];
  __syncthreads();
  if (ty == 0 && col < inner) {
#pragma unroll
    for (int v = 0; v < VEC; ++v) {
      float r = smem[0][tx * VEC + v];
#pragma unroll
      for (int k = 1; k < 8; ++k) r = op(r, smem[k][tx * VEC + v]);
      out[(o * splits + blockIdx.y) * inner + col + v] = r;
    }
  }
}

template <class Op>
static int reduce_impl(const float* x, float* out, int64_t outer, int64_t red, int64_t inner) {
  const int sms = sm_count();
  if (inner == 1) {
    // ---- row reduction ----
    if (red <= 1024 && outer >= 64) {
      int grid = grid_for(outer, RED_THREADS / 32);
      mc::launch(reduce_rows_warp_kernel<Op>, grid, RED_THREADS, 0, x, out, outer, red);
      MC_LAUNCHED();
      return 0;
    }
    int splits = 1;
    if (outer < 8 * sms && red > 8192) {     // 8 CTAs of 256 threads per SM keep enough loads in flight
      int64_t want = (8 * sms + outer - 1) / outer;
      int64_t maxs = red / 4096;
      splits = (int)(want < maxs ? want : maxs);
      if (splits < 1) splits = 1;
    }
    int64_t seg = (red + splits - 1) / splits;
    seg = (seg + 3) & ~(int64_t)3;  // keep segment starts float4-aligned
    splits = (int)((red + seg - 1) / seg);
    MC_REQUIRE(outer * splits < (int64_t)2147483647, "too many rows");
    if (splits == 1) {
      mc::launch(reduce_rows_block_kernel<Op>, (unsigned)outer, RED_THREADS, 0, x, out, red, 1, seg);
      MC_LAUNCHED();
      return 0;
    }
    ScratchFrame scratch;
    float* ws = scratch.floats((size_t)(outer * splits));
    if (!ws) return 1;
    mc::launch(reduce_rows_block_kernel<Op>, (unsigned)(outer * splits), RED_THREADS, 0, x, ws, red, splits, seg);
    MC_LAUNCHED();
    return reduce_impl<Op>(ws, out, outer, splits, 1);
  }
  // ---- column reduction ----
  if ((inner % 8) == 0 && aligned16(x) && red >= 128 && red * inner <= (1 << 20) && outer <= 65535) {
    dim3 grid((unsigned)(inner / 8), (unsigned)outer);
    mc::launch(reduce_cols_slab_kernel<Op>, grid, RED_THREADS, 0, x, out, red, inner);
    MC_LAUNCHED();
    return 0;
  }
  const bool vec = (inner % 4 == 0) && aligned16(x);
  const int vecw = vec ? 4 : 1;
  const int64_t col_tiles = (inner + 32 * vecw - 1) / (32 * vecw);
  MC_REQUIRE(outer <= 65535, "outer too large for column reduction");
  int splits = 1;
  if (col_tiles * outer < 8 * sms && red > 64) {
    int64_t want = (8 * sms + col_tiles * outer - 1) / (col_tiles * outer);
    int64_t maxs = red / 64;
    splits = (int)(want < maxs ? want : maxs);
    if (splits < 1) splits = 1;
    if (splits > 1024) splits = 1024;
  }
  int64_t seg = (red + splits - 1) / splits;
  splits = (int)((red + seg - 1) / seg);
  dim3 grid((unsigned)col_tiles, (unsigned)splits, (unsigned)outer);
  float* dst = out;
  float* ws = nullptr;
  ScratchFrame scratch;
  if (splits > 1) {
    ws = scratch.floats((size_t)(outer * splits * inner));
    if (!ws) return 1;
    dst = ws;
  }
  if (vec) mc::launch(reduce_cols_kernel<Op, 4>, grid, RED_THREADS, 0, x, dst, red, inner, seg);
  else mc::launch(reduce_cols_kernel<Op, 1>, grid, RED_THREADS, 0, x, dst, red, inner, seg);
  MC_LAUNCHED();
  if (splits > 1) {
    return reduce_impl<Op>(ws, out, outer, splits, inner);
  }
  return 0;
}

// Internal entry used by other translation units (softmax.cu, layers.cu).
int reduce_sum_internal(const float* x, float* out, int64_t outer, int64_t red, int64_t inner) {
  return reduce_impl<SumOp>(x, out, outer, red, inner);
}

// ---- argmax: first index of the maximum; a NaN counts as the maximum (NumPy) --------
__device__ __forceinline__ bool arg_better(float v, int64_t i, float bv, int64_t bi) {
  // true if (v,i) should replace (bv,bi)
  const bool vn = (v != v), bn = (bv != bv);
  if (vn || bn) return vn && (!bn || i < bi);
  return (v > bv) || (v == bv && i < bi);
}

__global__ void __launch_bounds__(RED_THREADS) argmax_thread_kernel(const float* __restrict__ x,
                                                                    int64_t* __restrict__ out, int64_t outer,
                                                                    int64_t red, int64_t inner) {
  mc::pdl_sync();
  const int64_t n = outer * inner;
  const int64_t stride = (int64_t)gridDim.x * RED_THREADS;
  for (int64_t t = (int64_t)blockIdx.x * RED_THREADS + threadIdx.x; t < n; t += stride) {
    const int64_t o = t / inner, i = t % inner;
    const float* __restrict__ p = x + o * red * inner + i;
    float bv = p[0];
    int64_t bi = 0;
    for (int64_t r = 1; r < red; ++r) {
      float v = p[r * inner];
      if (arg_better(v, r, bv, bi)) { bv = v; bi = r; }
    }
    out[t] = bi;
  }
}

__global__ void __launch_bounds__(RED_THREADS) argmax_rows_warp_kernel(const float* __restrict__ x,
                                                                       int64_t* __restrict__ out, int64_t rows,
                                                                       int64_t red) {
  mc::pdl_sync();
  const int lane = threadIdx.x & 31;
  const int64_t warps_per_grid = (int64_t)gridDim.x * (RED_THREADS / 32);
  for (int64_t row = (int64_t)blockIdx.x * (RED_THREADS / 32) + (threadIdx.x >> 5); row < rows; row += warps_per_grid) {
    const float* __restrict__ p = x + row * red;
    float bv = -INFINITY;
    int64_t bi = INT64_MAX;
    for (int64_t i = lane; i < red; i += 32) {
      float v = p[i];
      if (bi == INT64_MAX || arg_better(v, i, bv, bi)) { bv = v; bi = i; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      float ov = __shfl_xor_sync(0xffffffffu, bv, o);
      int64_t oi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (oi != INT64_MAX && (bi == INT64_MAX || arg_better(ov, oi, bv, bi))) { bv = ov; bi = oi; }
    }
    if (lane == 0) out[row] = bi;
  }
}

__global__ void __launch_bounds__(RED_THREADS) bcast_mid_kernel(const float* __restrict__ g, float* __restrict__ out,
                                                                float scale, int64_t red, int64_t inner, int64_t n) {
  mc::pdl_sync();
  const int64_t stride = (int64_t)gridDim.x * RED_THREADS;
  const int64_t ri = red * inner;
  for (int64_t t = (int64_t)blockIdx.x * RED_THREADS + threadIdx.x; t < n; t += stride) {
    const int64_t o = t / ri, i = t % inner;
    out[t] = __ldg(g + o * inner + i) * scale;
  }
}

__global__ void __launch_bounds__(RED_THREADS) select_grad_kernel(const float* __restrict__ x,
                                                                  const float* __restrict__ ans,
                                                                  const float* __restrict__ g, float* __restrict__ out,
                                                                  int64_t red, int64_t inner, int64_t n) {
  mc::pdl_sync();
  const int64_t stride = (int64_t)gridDim.x * RED_THREADS;
  const int64_t ri = red * inner;
  for (int64_t t = (int64_t)blockIdx.x * RED_THREADS + threadIdx.x; t < n; t += stride) {
    const int64_t o = t / ri, i = t % inner;
    const float a = __ldg(ans + o * inner + i);
    out[t] = (x[t] == a) ? __ldg(g + o * inner + i) : 0.f;
  }
}

}  // namespace mc

using namespace mc;

extern "C" {

int mc_reduce(int op, const float* x, float* out, int64_t outer, int64_t red, int64_t inner) {
  MC_REQUIRE(outer >= 0 && red >= 0 && inner >= 0, "negative extent");
  if (outer * inner == 0) return 0;
  if (red == 0) {
    MC_REQUIRE(op == MC_RED_SUM, "max/min of an empty axis has no identity");
    return mc_fill(out, 0.f, outer * inner);
  }
  switch (op) {
    case MC_RED_SUM: return reduce_impl<SumOp>(x, out, outer, red, inner);
    case MC_RED_MAX: return reduce_impl<MaxOp>(x, out, outer, red, inner);
    case MC_RED_MIN: return reduce_impl<MinOp>(x, out, outer, red, inner);
    default: mc::set_error("unknown reduce op %d", op); return 2;
  }
}

int mc_argmax(const float* x, int64_t* out, int64_t outer, int64_t red, int64_t inner) {
  MC_REQUIRE(red > 0, "argmax of an empty axis");
  if (outer * inner == 0) return 0;
  if (inner == 1 && red >= 64) {
    mc::launch(argmax_rows_warp_kernel, grid_for(outer, RED_THREADS / 32), RED_THREADS, 0, x, out, outer, red);
  } else {
    mc::launch(argmax_thread_kernel, grid_for(outer * inner, RED_THREADS), RED_THREADS, 0, x, out, outer, red, inner);
  }
  MC_LAUNCHED();
  return 0;
}

int mc_bcast_mid(const float* g, float* out, float scale, int64_t outer, int64_t red, int64_t inner) {
  const int64_t n = outer * red * inner;
  if (n <= 0) return 0;
  mc::launch(bcast_mid_kernel, grid_for(n, RED_THREADS), RED_THREADS, 0, g, out, scale, red, inner, n);
  MC_LAUNCHED();
  return 0;
}

int mc_select_grad(const float* x, const float* ans, const float* g, float* out, int64_t outer, int64_t red,
                   int64_t inner) {
  const int64_t n = outer * red * inner;
  if (n <= 0) return 0;
  mc::launch(select_grad_kernel, grid_for(n, RED_THREADS), RED_THREADS, 0, x, ans, g, out, red, inner, n);
  MC_LAUNCHED();
  return 0;
}

}  // extern "C"
