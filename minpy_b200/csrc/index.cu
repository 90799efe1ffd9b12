// Indexing kernels: strided slice / permute copies, gathers, and their scatter-add VJPs.
// Integer work here is bit-exact by construction (pure data movement); indices are int32
// as in the reference (minpy/array.py:494-495 casts index arrays to int32).
#include "common.cuh"

namespace mc {

constexpr int IX_THREADS = 256;

struct StridedDesc {
  int ndim;
  int64_t shape[MC_MAX_DIMS];
  int64_t strides[MC_MAX_DIMS];
  int64_t offset;
};

// ADD == false: dst[i] = src[off(i)]         (gather into dense)
// ADD == true : dst[off(i)] += src[i]        (scatter-add from dense; offsets are unique
//                                             for a basic slice so no atomics are needed)
template <bool ADD>
__global__ void __launch_bounds__(IX_THREADS) strided_kernel(const float* __restrict__ src, float* __restrict__ dst,
                                                             int64_t n, const __grid_constant__ StridedDesc d) {
  mc::pdl_sync();
  const int64_t stride = (int64_t)gridDim.x * IX_THREADS;
  for (int64_t i = (int64_t)blockIdx.x * IX_THREADS + threadIdx.x; i < n; i += stride) {
    int64_t rem = i, off = d.offset;
#pragma unroll
    for (int k = MC_MAX_DIMS - 1; k >= 0; --k) {
      if (k < d.ndim) {
        off += (rem % d.shape[k]) * d.strides[k];
        rem /= d.shape[k];
      }
    }
    if (ADD) dst[off] += src[i];
    else dst[i] = src[off];
  }
}

// The same, when the innermost dimension is contiguous on both sides and everything is 16-byte aligned
// (axis swaps of (N, T, H) sequences, slices along outer axes): one float4 per item, the index
// arithmetic once per four elements.  n4 = number of float4 items; shape[ndim-1] is in float4 units here.
template <bool ADD>
__global__ void __launch_bounds__(IX_THREADS) strided_rows4_kernel(const float4* __restrict__ src, float4* __restrict__ dst,
                                                                   int64_t n4, const __grid_constant__ StridedDesc d) {
  mc::pdl_sync();
  const int64_t stride = (int64_t)gridDim.x * IX_THREADS;
  for (int64_t i = (int64_t)blockIdx.x * IX_THREADS + threadIdx.x; i < n4; i += stride) {
    int64_t rem = i, off = d.offset;
#pragma unroll
    for (int k = MC_MAX_DIMS - 1; k >= 0; --k) {
      if (k < d.ndim) {
        off += (rem % d.shape[k]) * d.strides[k];
        rem /= d.shape[k];
      }
    }
    if (ADD) {
      float4 a = dst[off];
      const float4 b = src[i];
      a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
      dst[off] = a;
    } else {
      dst[i] = src[off];
    }
  }
}

// float4 view of a strided description, or false when it does not apply
static bool rows4_desc(const StridedDesc& d, const void* dense, const void* strided, StridedDesc* out) {
  if (d.ndim < 1 || d.strides[d.ndim - 1] != 1 || (d.shape[d.ndim - 1] & 3) || (d.offset & 3)) return false;
  if (!aligned16(dense) || !aligned16(strided)) return false;
  *out = d;
  for (int k = 0; k < d.ndim - 1; ++k) {
    if (d.strides[k] & 3) return false;
    out->strides[k] = d.strides[k] >> 2;
  }
  out->shape[d.ndim - 1] = d.shape[d.ndim - 1] >> 2;
  out->offset = d.offset >> 2;
  return true;
}

// 2-D tiled transpose through shared memory: both the read and the write are coalesced.
// src is (rows, cols) with row pitch `ld`, dst is (cols, rows) dense; batched over z.
__global__ void __launch_bounds__(256) transpose_tile_kernel(const float* __restrict__ src, float* __restrict__ dst,
                                                             int64_t rows, int64_t cols) {
  mc::pdl_sync();
  __shared__ float tile[32][33];
  const int64_t b = blockIdx.z;
  const float* __restrict__ s = src + b * rows * cols;
  float* __restrict__ t = dst + b * rows * cols;
  const int64_t c0 = (int64_t)blockIdx.x * 32, r0 = (int64_t)blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
#pragma unroll
  for (int j = 0; j < 32; j += 8) {
    int64_t r = r0 + ty + j, c = c0 + tx;
    if (r < rows && c < cols) tile[ty + j][tx] = s[r * cols + c];
  }
  __syncthreads();
#pragma unroll
  for (int j = 0; j < 32; j += 8) {
    int64_t c = c0 + ty + j, r = r0 + tx;
    if (r < rows && c < cols) t[c * rows + r] = tile[tx][ty + j];
  }
}

__global__ void __launch_bounds__(IX_THREADS) pick_cols_kernel(const float* __restrict__ x, const int32_t* __restrict__ idx,
                                                               float* __restrict__ out, int64_t rows, int64_t cols) {
  mc::pdl_sync();
  const int64_t stride = (int64_t)gridDim.x * IX_THREADS;
  for (int64_t r = (int64_t)blockIdx.x * IX_THREADS + threadIdx.x; r < rows; r += stride) {
    int32_t c = idx[r];
    if (c < 0) c += (int32_t)cols;
    out[r] = x[r * cols + c];
  }
}

__global__ void __launch_bounds__(IX_THREADS) pick_cols_grad_kernel(const float* __restrict__ g,
                                                                    const int32_t* __restrict__ idx, float* __restrict__ dx,
                                                                    int64_t rows, int64_t cols) {
  mc::pdl_sync();
  const int64_t stride = (int64_t)gridDim.x * IX_THREADS;
  for (int64_t r = (int64_t)blockIdx.x * IX_THREADS + threadIdx.x; r < rows; r += stride) {
    int32_t c = idx[r];
    if (c < 0) c += (int32_t)cols;
    dx[r * cols + c] += g[r];  // one writer per row
  }
}

__global__ void __launch_bounds__(IX_THREADS) take_rows_kernel(const float* __restrict__ x, const int32_t* __restrict__ idx,
                                                               float* __restrict__ out, int64_t nidx, int64_t row_elems,
                                                               int64_t src_rows) {
  mc::pdl_sync();
  const int64_t n = nidx * row_elems;
  const int64_t stride = (int64_t)gridDim.x * IX_THREADS;
  for (int64_t t = (int64_t)blockIdx.x * IX_THREADS + threadIdx.x; t < n; t += stride) {
    const int64_t i = t / row_elems, e = t % row_elems;
    int64_t r = idx[i];
    if (r < 0) r += src_rows;
    out[t] = x[r * row_elems + e];
  }
}

__global__ void __launch_bounds__(IX_THREADS) take_rows_grad_kernel(const float* __restrict__ g,
                                                                    const int32_t* __restrict__ idx, float* __restrict__ dx,
                                                                    int64_t nidx, int64_t row_elems, int64_t src_rows) {
  mc::pdl_sync();
  const int64_t n = nidx * row_elems;
  const int64_t stride = (int64_t)gridDim.x * IX_THREADS;
  for (int64_t t = (int64_t)blockIdx.x * IX_THREADS + threadIdx.x; t < n; t += stride) {
    const int64_t i = t / row_elems, e = t % row_elems;
    int64_t r = idx[i];
    if (r < 0) r += src_rows;
    atomicAdd(dx + r * row_elems + e, g[t]);  // duplicates accumulate (np.add.at)
  }
}

__global__ void __launch_bounds__(IX_THREADS) concat_mid_kernel(const float* __restrict__ a, const float* __restrict__ b,
                                                                float* __restrict__ out, int64_t na, int64_t nb,
                                                                int64_t inner, int64_t n) {
  mc::pdl_sync();
  const int64_t stride = (int64_t)gridDim.x * IX_THREADS;
  const int64_t mid = na + nb;
  for (int64_t t = (int64_t)blockIdx.x * IX_THREADS + threadIdx.x; t < n; t += stride) {
    const int64_t i = t % inner;
    const int64_t m = (t / inner) % mid;
    const int64_t o = t / (inner * mid);
    out[t] = (m < na) ? a[(o * na + m) * inner + i] : b[(o * nb + (m - na)) * inner + i];
  }
}

static int fill_desc(StridedDesc& d, int ndim, const int64_t* shape, const int64_t* strides, int64_t offset, int64_t* n) {
  if (ndim < 0 || ndim > MC_MAX_DIMS) return 2;
  d.ndim = ndim;
  d.offset = offset;
  *n = 1;
  for (int k = 0; k < MC_MAX_DIMS; ++k) {
    d.shape[k] = k < ndim ? shape[k] : 1;
    d.strides[k] = k < ndim ? strides[k] : 0;
    *n *= d.shape[k];
  }
  return 0;
}

}  // namespace mc

using namespace mc;

extern "C" {

int mc_strided_copy(const float* src, float* dst, int ndim, const int64_t* shape, const int64_t* src_strides,
                    int64_t src_offset) {
  StridedDesc d;
  int64_t n;
  MC_REQUIRE(fill_desc(d, ndim, shape, src_strides, src_offset, &n) == 0, "ndim must be in 0..6");
  if (n <= 0) return 0;
  // (batch, rows, cols) -> (batch, cols, rows): the np.transpose / .T hot case
  if (ndim >= 2 && src_strides[ndim - 1] == shape[ndim - 2] && src_strides[ndim - 2] == 1 && src_offset == 0) {
    int64_t rows = shape[ndim - 1], cols = shape[ndim - 2];  // src is (rows, cols) dense
    int64_t batch = 1;
    bool ok = true;
    int64_t expect = rows * cols;
    for (int k = ndim - 3; k >= 0; --k) {
      if (src_strides[k] != expect) ok = false;
      expect *= shape[k];
      batch *= shape[k];
    }
    if (ok && batch <= 65535 && (rows + 31) / 32 <= 65535) {
      dim3 grid((unsigned)((cols + 31) / 32), (unsigned)((rows + 31) / 32), (unsigned)batch);
      mc::launch(transpose_tile_kernel, grid, 256, 0, src, dst, rows, cols);
      MC_LAUNCHED();
      return 0;
    }
  }
  StridedDesc d4;
  if (rows4_desc(d, dst, src, &d4)) {
    mc::launch(strided_rows4_kernel<false>, grid_for(n / 4, IX_THREADS), IX_THREADS, 0, reinterpret_cast<const float4*>(src),
               reinterpret_cast<float4*>(dst), n / 4, d4);
    MC_LAUNCHED();
    return 0;
  }
  mc::launch(strided_kernel<false>, grid_for(n, IX_THREADS), IX_THREADS, 0, src, dst, n, d);
  MC_LAUNCHED();
  return 0;
}

int mc_strided_add(float* dst, const float* src, int ndim, const int64_t* shape, const int64_t* dst_strides,
                   int64_t dst_offset) {
  StridedDesc d;
  int64_t n;
  MC_REQUIRE(fill_desc(d, ndim, shape, dst_strides, dst_offset, &n) == 0, "ndim must be in 0..6");
  if (n <= 0) return 0;
  StridedDesc d4;
  if (rows4_desc(d, src, dst, &d4)) {
    mc::launch(strided_rows4_kernel<true>, grid_for(n / 4, IX_THREADS), IX_THREADS, 0, reinterpret_cast<const float4*>(src),
               reinterpret_cast<float4*>(dst), n / 4, d4);
    MC_LAUNCHED();
    return 0;
  }
  mc::launch(strided_kernel<true>, grid_for(n, IX_THREADS), IX_THREADS, 0, src, dst, n, d);
  MC_LAUNCHED();
  return 0;
}

int mc_pick_cols(const float* x, const int32_t* idx, float* out, int64_t rows, int64_t cols) {
  if (rows <= 0) return 0;
  mc::launch(pick_cols_kernel, grid_for(rows, IX_THREADS), IX_THREADS, 0, x, idx, out, rows, cols);
  MC_LAUNCHED();
  return 0;
}

int mc_pick_cols_grad(const float* g, const int32_t* idx, float* dx, int64_t rows, int64_t cols) {
  if (rows <= 0) return 0;
  mc::launch(pick_cols_grad_kernel, grid_for(rows, IX_THREADS), IX_THREADS, 0, g, idx, dx, rows, cols);
  MC_LAUNCHED();
  return 0;
}

int mc_take_rows(const float* x, const int32_t* idx, float* out, int64_t nidx, int64_t row_elems) {
  if (nidx * row_elems <= 0) return 0;
  mc::launch(take_rows_kernel, grid_for(nidx * row_elems, IX_THREADS), IX_THREADS, 0, x, idx, out, nidx, row_elems,
                                                                                       (int64_t)1 << 40);
  MC_LAUNCHED();
  return 0;
}

int mc_take_rows_grad(const float* g, const int32_t* idx, float* dx, int64_t nidx, int64_t row_elems) {
  if (nidx * row_elems <= 0) return 0;
  mc::launch(take_rows_grad_kernel, grid_for(nidx * row_elems, IX_THREADS), IX_THREADS, 0, g, idx, dx, nidx, row_elems,
                                                                                            (int64_t)1 << 40);
  MC_LAUNCHED();
  return 0;
}

int mc_concat_mid(const float* a, const float* b, float* out, int64_t outer, int64_t na, int64_t nb, int64_t inner) {
  const int64_t n = outer * (na + nb) * inner;
  if (n <= 0) return 0;
  mc::launch(concat_mid_kernel, grid_for(n, IX_THREADS), IX_THREADS, 0, a, b, out, na, nb, inner, n);
  MC_LAUNCHED();
  return 0;
}

}  // extern "C"
