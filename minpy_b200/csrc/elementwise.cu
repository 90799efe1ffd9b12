// Elementwise ufuncs, broadcasting, casts, and the fused VJP-chain kernel.
//
// All kernels here are HBM-bound (unary 8n B, binary 12n B).  Design:
// 128-bit coalesced loads/stores, 4 independent float4 requests in flight per thread
// (>= 32 KB in flight per SM at full occupancy), grid sized to a multiple of the SM
// count and grid-strided, no shared memory (no reuse to exploit).
#include "common.cuh"

namespace mc {

constexpr int EW_THREADS = 256;
constexpr int EW_UNROLL = 4;  // float4 requests in flight per thread

// ------------------------------------------------------------------ functors --------
template <int OP>
__device__ __forceinline__ float unary_apply(float x) {
  if constexpr (OP == MC_NEG) return -x;
  else if constexpr (OP == MC_ABS) return fabsf(x);
  else if constexpr (OP == MC_SIGN) return (x != x) ? x : (float)((x > 0.f) - (x < 0.f));
  else if constexpr (OP == MC_ROUND) return rintf(x);  // half-to-even, as numpy.round
  else if constexpr (OP == MC_CEIL) return ceilf(x);
  else if constexpr (OP == MC_FLOOR) return floorf(x);
  else if constexpr (OP == MC_SQRT) return sqrtf(x);
  else if constexpr (OP == MC_EXP) return expf(x);
  else if constexpr (OP == MC_LOG) return logf(x);
  else if constexpr (OP == MC_SIN) return sinf(x);
  else if constexpr (OP == MC_COS) return cosf(x);
  else if constexpr (OP == MC_TANH) return tanhf(x);
  else if constexpr (OP == MC_SIGMOID) {
    // numerically stable form of numpy_core.py:32-43
    float z = expf(-fabsf(x));
    return (x >= 0.f) ? 1.f / (1.f + z) : z / (1.f + z);
  }
  else if constexpr (OP == MC_SQUARE) return x * x;
  else if constexpr (OP == MC_RECIP) return 1.f / x;
  else if constexpr (OP == MC_RELU) return (x != x) ? x : fmaxf(x, 0.f);
  else if constexpr (OP == MC_COPY) return x;
  else if constexpr (OP == MC_COSH) return coshf(x);
  else if constexpr (OP == MC_LOGICAL_NOT) return (x == 0.f) ? 1.f : 0.f;
  else return x;
}

__device__ __forceinline__ float floored_mod(float a, float b) {
  float r = fmodf(a, b);
  if (r != 0.f && ((r < 0.f) != (b < 0.f))) r += b;
  return r;
}

template <int OP>
__device__ __forceinline__ float binary_apply(float a, float b) {
  if constexpr (OP == MC_ADD) return a + b;
  else if constexpr (OP == MC_SUB) return a - b;
  else if constexpr (OP == MC_MUL) return a * b;
  else if constexpr (OP == MC_DIV) return a / b;
  else if constexpr (OP == MC_POW) return powf(a, b);
  else if constexpr (OP == MC_MAX) return nanmax(a, b);
  else if constexpr (OP == MC_MIN) return nanmin(a, b);
  else if constexpr (OP == MC_MOD) return floored_mod(a, b);
  else if constexpr (OP == MC_EQ) return (a == b) ? 1.f : 0.f;
  else if constexpr (OP == MC_NE) return (a != b) ? 1.f : 0.f;
  else if constexpr (OP == MC_LT) return (a < b) ? 1.f : 0.f;
  else if constexpr (OP == MC_GT) return (a > b) ? 1.f : 0.f;
  else if constexpr (OP == MC_LE) return (a <= b) ? 1.f : 0.f;
  else if constexpr (OP == MC_GE) return (a >= b) ? 1.f : 0.f;
  else return a;
}

__device__ __forceinline__ float binary_apply_rt(int op, float a, float b) {
  switch (op) {
    case MC_ADD: return a + b;
    case MC_SUB: return a - b;
    case MC_MUL: return a * b;
    case MC_DIV: return a / b;
    case MC_POW: return powf(a, b);
    case MC_MAX: return nanmax(a, b);
    case MC_MIN: return nanmin(a, b);
    case MC_MOD: return floored_mod(a, b);
    case MC_EQ: return (a == b) ? 1.f : 0.f;
    case MC_NE: return (a != b) ? 1.f : 0.f;
    case MC_LT: return (a < b) ? 1.f : 0.f;
    case MC_GT: return (a > b) ? 1.f : 0.f;
    case MC_LE: return (a <= b) ? 1.f : 0.f;
    default: return (a >= b) ? 1.f : 0.f;
  }
}

// ------------------------------------------------------------------ kernels ---------
// Vector body: element i4 indexes float4s.  `n4` float4s, then up to 3 tail scalars.
template <int OP>
__global__ void __launch_bounds__(EW_THREADS) unary_vec_kernel(const float* __restrict__ x,
                                                               float* __restrict__ y, int64_t n) {
  mc::pdl_sync();
  const int64_t n4 = n >> 2;
  const float4* __restrict__ x4 = reinterpret_cast<const float4*>(x);
  float4* __restrict__ y4 = reinterpret_cast<float4*>(y);
  const int64_t stride = (int64_t)gridDim.x * EW_THREADS * EW_UNROLL;
  for (int64_t base = (int64_t)blockIdx.x * EW_THREADS * EW_UNROLL + threadIdx.x; base < n4; base += stride) {
    float4 v[EW_UNROLL];
#pragma unroll
    for (int j = 0; j < EW_UNROLL; ++j) {
      int64_t i = base + (int64_t)j * EW_THREADS;
      if (i < n4) v[j] = x4[i];
    }
#pragma unroll
    for (int j = 0; j < EW_UNROLL; ++j) {
      int64_t i = base + (int64_t)j * EW_THREADS;
      if (i < n4) {
        float4 r;
        r.x = unary_apply<OP>(v[j].x);
        r.y = unary_apply<OP>(v[j].y);
        r.z = unary_apply<OP>(v[j].z);
        r.w = unary_apply<OP>(v[j].w);
        y4[i] = r;
      }
    }
  }
  if (blockIdx.x == 0) {
    int64_t t = (n4 << 2) + threadIdx.x;
    if (t < n) y[t] = unary_apply<OP>(x[t]);
  }
}

template <int OP>
__global__ void __launch_bounds__(EW_THREADS) unary_scalar_kernel(const float* __restrict__ x,
                                                                  float* __restrict__ y, int64_t n) {
  mc::pdl_sync();
  const int64_t stride = (int64_t)gridDim.x * EW_THREADS;
  for (int64_t i = (int64_t)blockIdx.x * EW_THREADS + threadIdx.x; i < n; i += stride)
    y[i] = unary_apply<OP>(x[i]);
}

template <int OP>
__global__ void __launch_bounds__(EW_THREADS) binary_vec_kernel(const float* __restrict__ a,
                                                                const float* __restrict__ b,
                                                                float* __restrict__ out, int64_t n) {
  mc::pdl_sync();
  const int64_t n4 = n >> 2;
  const float4* __restrict__ a4 = reinterpret_cast<const float4*>(a);
  const float4* __restrict__ b4 = reinterpret_cast<const float4*>(b);
  float4* __restrict__ o4 = reinterpret_cast<float4*>(out);
  const int64_t stride = (int64_t)gridDim.x * EW_THREADS * EW_UNROLL;
  for (int64_t base = (int64_t)blockIdx.x * EW_THREADS * EW_UNROLL + threadIdx.x; base < n4; base += stride) {
    float4 va[EW_UNROLL], vb[EW_UNROLL];
#pragma unroll
    for (int j = 0; j < EW_UNROLL; ++j) {
      int64_t i = base + (int64_t)j * EW_THREADS;
      if (i < n4) { va[j] = a4[i]; vb[j] = b4[i]; }
    }
#pragma unroll
    for (int j = 0; j < EW_UNROLL; ++j) {
      int64_t i = base + (int64_t)j * EW_THREADS;
      if (i < n4) {
        float4 r;
        r.x = binary_apply<OP>(va[j].x, vb[j].x);
        r.y = binary_apply<OP>(va[j].y, vb[j].y);
        r.z = binary_apply<OP>(va[j].z, vb[j].z);
        r.w = binary_apply<OP>(va[j].w, vb[j].w);
        o4[i] = r;
      }
    }
  }
  if (blockIdx.x == 0) {
    int64_t t = (n4 << 2) + threadIdx.x;
    if (t < n) out[t] = binary_apply<OP>(a[t], b[t]);
  }
}

template <int OP>
__global__ void __launch_bounds__(EW_THREADS) binary_scalar_elems_kernel(const float* __restrict__ a,
                                                                         const float* __restrict__ b,
                                                                         float* __restrict__ out, int64_t n) {
  mc::pdl_sync();
  const int64_t stride = (int64_t)gridDim.x * EW_THREADS;
  for (int64_t i = (int64_t)blockIdx.x * EW_THREADS + threadIdx.x; i < n; i += stride)
    out[i] = binary_apply<OP>(a[i], b[i]);
}

template <int OP, bool LEFT>
__global__ void __launch_bounds__(EW_THREADS) binary_const_kernel(const float* __restrict__ a, float s,
                                                                  float* __restrict__ out, int64_t n, int vec_ok) {
  mc::pdl_sync();
  if (vec_ok) {
    const int64_t n4 = n >> 2;
    const float4* __restrict__ a4 = reinterpret_cast<const float4*>(a);
    float4* __restrict__ o4 = reinterpret_cast<float4*>(out);
    const int64_t stride = (int64_t)gridDim.x * EW_THREADS * EW_UNROLL;
    for (int64_t base = (int64_t)blockIdx.x * EW_THREADS * EW_UNROLL + threadIdx.x; base < n4; base += stride) {
      float4 v[EW_UNROLL];
#pragma unroll
      for (int j = 0; j < EW_UNROLL; ++j) {
        int64_t i = base + (int64_t)j * EW_THREADS;
        if (i < n4) v[j] = a4[i];
      }
#pragma unroll
      for (int j = 0; j < EW_UNROLL; ++j) {
        int64_t i = base + (int64_t)j * EW_THREADS;
        if (i < n4) {
          float4 r;
          if (LEFT) {
            r.x = binary_apply<OP>(s, v[j].x); r.y = binary_apply<OP>(s, v[j].y);
            r.z = binary_apply<OP>(s, v[j].z); r.w = binary_apply<OP>(s, v[j].w);
          } else {
            r.x = binary_apply<OP>(v[j].x, s); r.y = binary_apply<OP>(v[j].y, s);
            r.z = binary_apply<OP>(v[j].z, s); r.w = binary_apply<OP>(v[j].w, s);
          }
          o4[i] = r;
        }
      }
    }
    if (blockIdx.x == 0) {
      int64_t t = (n4 << 2) + threadIdx.x;
      if (t < n) out[t] = LEFT ? binary_apply<OP>(s, a[t]) : binary_apply<OP>(a[t], s);
    }
  } else {
    const int64_t stride = (int64_t)gridDim.x * EW_THREADS;
    for (int64_t i = (int64_t)blockIdx.x * EW_THREADS + threadIdx.x; i < n; i += stride)
      out[i] = LEFT ? binary_apply<OP>(s, a[i]) : binary_apply<OP>(a[i], s);
  }
}

// (rows, cols) o row-vector(cols) / col-vector(rows).  vec4 path needs cols % 4 == 0.
template <int OP, bool VEC_IS_COL, bool VEC_LEFT>
__global__ void __launch_bounds__(EW_THREADS) binary_vec_bcast_kernel(const float* __restrict__ a,
                                                                      const float* __restrict__ v,
                                                                      float* __restrict__ out, int64_t rows,
                                                                      int64_t cols, int vec_ok) {
  mc::pdl_sync();
  const int64_t n = rows * cols;
  if (vec_ok) {
    const int64_t n4 = n >> 2;
    const int64_t cols4 = cols >> 2;
    const float4* __restrict__ a4 = reinterpret_cast<const float4*>(a);
    const float4* __restrict__ v4 = reinterpret_cast<const float4*>(v);
    float4* __restrict__ o4 = reinterpret_cast<float4*>(out);
    const int64_t stride = (int64_t)gridDim.x * EW_THREADS * EW_UNROLL;
    for (int64_t base = (int64_t)blockIdx.x * EW_THREADS * EW_UNROLL + threadIdx.x; base < n4; base += stride) {
      float4 va[EW_UNROLL], vv[EW_UNROLL];
#pragma unroll
      for (int j = 0; j < EW_UNROLL; ++j) {
        int64_t i = base + (int64_t)j * EW_THREADS;
        if (i < n4) {
          va[j] = a4[i];
          if (VEC_IS_COL) {
            float s = __ldg(v + i / cols4);
            vv[j] = make_float4(s, s, s, s);
          } else {
            vv[j] = __ldg(v4 + (i % cols4));
          }
        }
      }
#pragma unroll
      for (int j = 0; j < EW_UNROLL; ++j) {
        int64_t i = base + (int64_t)j * EW_THREADS;
        if (i < n4) {
          float4 r;
          if (VEC_LEFT) {
            r.x = binary_apply<OP>(vv[j].x, va[j].x); r.y = binary_apply<OP>(vv[j].y, va[j].y);
            r.z = binary_apply<OP>(vv[j].z, va[j].z); r.w = binary_apply<OP>(vv[j].w, va[j].w);
          } else {
            r.x = binary_apply<OP>(va[j].x, vv[j].x); r.y = binary_apply<OP>(va[j].y, vv[j].y);
            r.z = binary_apply<OP>(va[j].z, vv[j].z); r.w = binary_apply<OP>(va[j].w, vv[j].w);
          }
          o4[i] = r;
        }
      }
    }
  } else {
    const int64_t stride = (int64_t)gridDim.x * EW_THREADS;
    for (int64_t i = (int64_t)blockIdx.x * EW_THREADS + threadIdx.x; i < n; i += stride) {
      float s = VEC_IS_COL ? __ldg(v + i / cols) : __ldg(v + i % cols);
      out[i] = VEC_LEFT ? binary_apply<OP>(s, a[i]) : binary_apply<OP>(a[i], s);
    }
  }
}

struct BcastDesc {
  int ndim;
  int op;
  int64_t shape[MC_MAX_DIMS];
  int64_t sa[MC_MAX_DIMS];
  int64_t sb[MC_MAX_DIMS];
};

__global__ void __launch_bounds__(EW_THREADS) binary_bcast_kernel(const float* __restrict__ a,
                                                                  const float* __restrict__ b,
                                                                  float* __restrict__ out, int64_t n,
                                                                  const __grid_constant__ BcastDesc d) {
  mc::pdl_sync();
  const int64_t stride = (int64_t)gridDim.x * EW_THREADS;
  for (int64_t i = (int64_t)blockIdx.x * EW_THREADS + threadIdx.x; i < n; i += stride) {
    int64_t rem = i, oa = 0, ob = 0;
#pragma unroll
    for (int k = MC_MAX_DIMS - 1; k >= 0; --k) {
      if (k < d.ndim) {
        int64_t idx = rem % d.shape[k];
        rem /= d.shape[k];
        oa += idx * d.sa[k];
        ob += idx * d.sb[k];
      }
    }
    out[i] = binary_apply_rt(d.op, __ldg(a + oa), __ldg(b + ob));
  }
}

__global__ void __launch_bounds__(EW_THREADS) fill_kernel(float* __restrict__ dst, float v, int64_t n, int vec_ok) {
  mc::pdl_sync();
  if (vec_ok) {
    const int64_t n4 = n >> 2;
    float4* d4 = reinterpret_cast<float4*>(dst);
    const float4 vv = make_float4(v, v, v, v);
    const int64_t stride = (int64_t)gridDim.x * EW_THREADS;
    for (int64_t i = (int64_t)blockIdx.x * EW_THREADS + threadIdx.x; i < n4; i += stride) d4[i] = vv;
    if (blockIdx.x == 0) {
      int64_t t = (n4 << 2) + threadIdx.x;
      if (t < n) dst[t] = v;
    }
  } else {
    const int64_t stride = (int64_t)gridDim.x * EW_THREADS;
    for (int64_t i = (int64_t)blockIdx.x * EW_THREADS + threadIdx.x; i < n; i += stride) dst[i] = v;
  }
}

__global__ void __launch_bounds__(EW_THREADS) arange_kernel(float* __restrict__ dst, float start, float step, int64_t n) {
  mc::pdl_sync();
  const int64_t stride = (int64_t)gridDim.x * EW_THREADS;
  for (int64_t i = (int64_t)blockIdx.x * EW_THREADS + threadIdx.x; i < n; i += stride)
    dst[i] = start + step * (float)i;
}

template <class S, class D>
__global__ void __launch_bounds__(EW_THREADS) cast_kernel(const S* __restrict__ src, D* __restrict__ dst, int64_t n) {
  mc::pdl_sync();
  const int64_t stride = (int64_t)gridDim.x * EW_THREADS;
  for (int64_t i = (int64_t)blockIdx.x * EW_THREADS + threadIdx.x; i < n; i += stride) dst[i] = (D)src[i];
}

// ---------------------------------------------------------------- VJP chain ---------
struct ChainDesc {
  int nsteps;
  int g_mode;
  int64_t g_inner;
  float g_scalar;
  mc_vjp_step steps[MC_VJP_MAX_STEPS];
};

__device__ __forceinline__ float opd_load(const float* __restrict__ p, int mode, int64_t inner, int64_t i) {
  switch (mode) {
    case MC_OPD_FULL: return __ldg(p + i);
    case MC_OPD_ROW: return __ldg(p + (i % inner));
    case MC_OPD_COL: return __ldg(p + (i / inner));
    default: return __ldg(p);
  }
}

__device__ __forceinline__ float4 opd_load4(const float* __restrict__ p, int mode, int64_t inner, int64_t i, bool vec) {
  // i is a multiple of 4; `vec` says FULL/ROW vector loads are legal for this operand
  if (mode == MC_OPD_FULL && vec) return __ldg(reinterpret_cast<const float4*>(p + i));
  if (mode == MC_OPD_ROW && vec && (inner & 3) == 0) return __ldg(reinterpret_cast<const float4*>(p + (i % inner)));
  if (mode == MC_OPD_ONE) { float s = __ldg(p); return make_float4(s, s, s, s); }
  return make_float4(opd_load(p, mode, inner, i), opd_load(p, mode, inner, i + 1),
                     opd_load(p, mode, inner, i + 2), opd_load(p, mode, inner, i + 3));
}

__device__ __forceinline__ float vjp_factor(int op, float g, float a, float b, float s) {
  switch (op) {
    case MC_VJP_MUL_A: return g * a;
    case MC_VJP_MUL_S: return g * s;
    case MC_VJP_NEG: return -g;
    case MC_VJP_DIV_A: return g / a;
    case MC_VJP_DIV_S: return g / s;
    case MC_VJP_TANH_ANS: return g * (1.f - a * a);
    case MC_VJP_SQRT_ANS: return g * 0.5f / a;
    case MC_VJP_SIN_X: return g * cosf(a);
    case MC_VJP_COS_X: return -g * sinf(a);
    case MC_VJP_ABS_X: return g * (float)((a > 0.f) - (a < 0.f));
    case MC_VJP_SIGMOID_ANS: return g * a * (1.f - a);
    case MC_VJP_EQ_AB: return (a == b) ? g : g * 0.f;
    case MC_VJP_EQ_AS: return (a == s) ? g : g * 0.f;
    case MC_VJP_DIV_Y: return -g * a / (b * b);
    case MC_VJP_S_DIV_Y: return -g * s / (a * a);
    case MC_VJP_POW_XS: return g * s * powf(a, s - 1.f);
    case MC_VJP_POW_SY: return g * logf(s) * a;
    case MC_VJP_POW_X: return g * b * powf(a, b - 1.f);
    case MC_VJP_POW_Y: return g * logf(a) * b;
    case MC_VJP_RECIP_ANS: return -g * a * a;
    case MC_VJP_GT0_A: return (a > 0.f) ? g : g * 0.f;
    case MC_VJP_GE0_A: return (a >= 0.f) ? g : g * 0.f;
    case MC_VJP_SIGNPOS_A: return (__float_as_uint(a) >> 31) ? g * 0.f : g;
    default: return g;
  }
}

__global__ void __launch_bounds__(EW_THREADS) vjp_chain_kernel(const float* __restrict__ g,
                                                               const float* __restrict__ acc,
                                                               float* __restrict__ out, int64_t n, int vec_ok,
                                                               const __grid_constant__ ChainDesc d) {
  mc::pdl_sync();
  if (vec_ok) {
    const int64_t n4 = n >> 2;
    const int64_t stride = (int64_t)gridDim.x * EW_THREADS;
    for (int64_t i4 = (int64_t)blockIdx.x * EW_THREADS + threadIdx.x; i4 < n4; i4 += stride) {
      const int64_t i = i4 << 2;
      float4 gv = g ? opd_load4(g, d.g_mode, d.g_inner, i, true)
                    : make_float4(d.g_scalar, d.g_scalar, d.g_scalar, d.g_scalar);
      for (int k = 0; k < d.nsteps; ++k) {
        const mc_vjp_step& st = d.steps[k];
        float4 a = st.a ? opd_load4(st.a, st.a_mode, st.inner, i, true) : make_float4(0, 0, 0, 0);
        float4 b = st.b ? opd_load4(st.b, st.b_mode, st.inner, i, true) : make_float4(0, 0, 0, 0);
        gv.x = vjp_factor(st.op, gv.x, a.x, b.x, st.scalar);
        gv.y = vjp_factor(st.op, gv.y, a.y, b.y, st.scalar);
        gv.z = vjp_factor(st.op, gv.z, a.z, b.z, st.scalar);
        gv.w = vjp_factor(st.op, gv.w, a.w, b.w, st.scalar);
      }
      if (acc) {
        float4 c = __ldg(reinterpret_cast<const float4*>(acc + i));
        gv.x += c.x; gv.y += c.y; gv.z += c.z; gv.w += c.w;
      }
      *reinterpret_cast<float4*>(out + i) = gv;
    }
    if (blockIdx.x == 0) {
      int64_t i = (n4 << 2) + threadIdx.x;
      if (i < n) {
        float gv = g ? opd_load(g, d.g_mode, d.g_inner, i) : d.g_scalar;
        for (int k = 0; k < d.nsteps; ++k) {
          const mc_vjp_step& st = d.steps[k];
          float a = st.a ? opd_load(st.a, st.a_mode, st.inner, i) : 0.f;
          float b = st.b ? opd_load(st.b, st.b_mode, st.inner, i) : 0.f;
          gv = vjp_factor(st.op, gv, a, b, st.scalar);
        }
        if (acc) gv += acc[i];
        out[i] = gv;
      }
    }
  } else {
    const int64_t stride = (int64_t)gridDim.x * EW_THREADS;
    for (int64_t i = (int64_t)blockIdx.x * EW_THREADS + threadIdx.x; i < n; i += stride) {
      float gv = g ? opd_load(g, d.g_mode, d.g_inner, i) : d.g_scalar;
      for (int k = 0; k < d.nsteps; ++k) {
        const mc_vjp_step& st = d.steps[k];
        float a = st.a ? opd_load(st.a, st.a_mode, st.inner, i) : 0.f;
        float b = st.b ? opd_load(st.b, st.b_mode, st.inner, i) : 0.f;
        gv = vjp_factor(st.op, gv, a, b, st.scalar);
      }
      if (acc) gv += acc[i];
      out[i] = gv;
    }
  }
}


// ------------------------------------------------------------- multi-tensor copy ----
// Packs up to MT_MAX tensors into (or out of) a flat buffer with one launch: the gradient
// bucket handed to the all-reduce.  Block b serves tile (b - first[t]) of tensor t.
constexpr int MT_MAX = 48;
constexpr int MT_TILE = EW_THREADS * 4 * 4;   // floats per block
struct MultiCopyDesc {
  int count;
  int first[MT_MAX + 1];        // first block index of each tensor
  const float* src[MT_MAX];
  float* dst[MT_MAX];
  int64_t n[MT_MAX];
};

__global__ void __launch_bounds__(EW_THREADS) multi_copy_kernel(const __grid_constant__ MultiCopyDesc d) {
  mc::pdl_sync();
  int t = 0;
  while (t + 1 < d.count && (int)blockIdx.x >= d.first[t + 1]) ++t;
  const float* __restrict__ src = d.src[t];
  float* __restrict__ dst = d.dst[t];
  const int64_t n = d.n[t];
  const int64_t lo = (int64_t)(blockIdx.x - d.first[t]) * MT_TILE;
  const int64_t hi = (lo + MT_TILE < n) ? lo + MT_TILE : n;
  if (aligned16_dev(src) && aligned16_dev(dst)) {
    const int64_t hi4 = lo + ((hi - lo) & ~(int64_t)3);
    for (int64_t i = lo + 4 * (int64_t)threadIdx.x; i < hi4; i += 4 * EW_THREADS)
      *reinterpret_cast<float4*>(dst + i) = __ldg(reinterpret_cast<const float4*>(src + i));
    for (int64_t i = hi4 + threadIdx.x; i < hi; i += EW_THREADS) dst[i] = src[i];
  } else {
    for (int64_t i = lo + threadIdx.x; i < hi; i += EW_THREADS) dst[i] = src[i];
  }
}

// ------------------------------------------------------------ host dispatch ---------
template <int OP>
static int launch_unary(const float* x, float* y, int64_t n) {
  if (aligned16(x) && aligned16(y)) {
    int grid = grid_for((n >> 2) + 1, EW_THREADS * EW_UNROLL);
    mc::launch(unary_vec_kernel<OP>, grid, EW_THREADS, 0, x, y, n);
  } else {
    int grid = grid_for(n, EW_THREADS);
    mc::launch(unary_scalar_kernel<OP>, grid, EW_THREADS, 0, x, y, n);
  }
  MC_LAUNCHED();
  return 0;
}

template <int OP>
static int launch_binary(const float* a, const float* b, float* out, int64_t n) {
  if (aligned16(a) && aligned16(b) && aligned16(out)) {
    int grid = grid_for((n >> 2) + 1, EW_THREADS * EW_UNROLL);
    mc::launch(binary_vec_kernel<OP>, grid, EW_THREADS, 0, a, b, out, n);
  } else {
    int grid = grid_for(n, EW_THREADS);
    mc::launch(binary_scalar_elems_kernel<OP>, grid, EW_THREADS, 0, a, b, out, n);
  }
  MC_LAUNCHED();
  return 0;
}

template <int OP>
static int launch_binary_const(const float* a, float s, int left, float* out, int64_t n) {
  int vec_ok = aligned16(a) && aligned16(out);
  int grid = vec_ok ? grid_for((n >> 2) + 1, EW_THREADS * EW_UNROLL) : grid_for(n, EW_THREADS);
  if (left) mc::launch(binary_const_kernel<OP, true>, grid, EW_THREADS, 0, a, s, out, n, vec_ok);
  else mc::launch(binary_const_kernel<OP, false>, grid, EW_THREADS, 0, a, s, out, n, vec_ok);
  MC_LAUNCHED();
  return 0;
}

template <int OP>
static int launch_binary_vecb(const float* a, const float* v, int is_col, int left, float* out, int64_t rows,
                              int64_t cols) {
  int vec_ok = aligned16(a) && aligned16(out) && aligned16(v) && (cols % 4 == 0);
  int64_t n = rows * cols;
  int grid = vec_ok ? grid_for((n >> 2) + 1, EW_THREADS * EW_UNROLL) : grid_for(n, EW_THREADS);
  if (is_col) {
    if (left) mc::launch(binary_vec_bcast_kernel<OP, true, true>, grid, EW_THREADS, 0, a, v, out, rows, cols, vec_ok);
    else mc::launch(binary_vec_bcast_kernel<OP, true, false>, grid, EW_THREADS, 0, a, v, out, rows, cols, vec_ok);
  } else {
    if (left) mc::launch(binary_vec_bcast_kernel<OP, false, true>, grid, EW_THREADS, 0, a, v, out, rows, cols, vec_ok);
    else mc::launch(binary_vec_bcast_kernel<OP, false, false>, grid, EW_THREADS, 0, a, v, out, rows, cols, vec_ok);
  }
  MC_LAUNCHED();
  return 0;
}

template <class S>
static int cast_from(const S* src, void* dst, int dst_dtype, int64_t n) {
  int grid = grid_for(n, EW_THREADS);
  switch (dst_dtype) {
    case MC_F32: mc::launch(cast_kernel<S, float>, grid, EW_THREADS, 0, src, (float*)dst, n); break;
    case MC_I32: mc::launch(cast_kernel<S, int32_t>, grid, EW_THREADS, 0, src, (int32_t*)dst, n); break;
    case MC_I64: mc::launch(cast_kernel<S, int64_t>, grid, EW_THREADS, 0, src, (int64_t*)dst, n); break;
    case MC_F64: mc::launch(cast_kernel<S, double>, grid, EW_THREADS, 0, src, (double*)dst, n); break;
    case MC_U8: mc::launch(cast_kernel<S, uint8_t>, grid, EW_THREADS, 0, src, (uint8_t*)dst, n); break;
    default: mc::set_error("unknown dst dtype %d", dst_dtype); return 2;
  }
  MC_LAUNCHED();
  return 0;
}

}  // namespace mc

using namespace mc;

#define UNARY_CASE(OP) case OP: return launch_unary<OP>(x, y, n);
#define BINARY_CASES(F, ...)                                                                    \
  switch (op) {                                                                                 \
    case MC_ADD: return F<MC_ADD>(__VA_ARGS__);                                                 \
    case MC_SUB: return F<MC_SUB>(__VA_ARGS__);                                                 \
    case MC_MUL: return F<MC_MUL>(__VA_ARGS__);                                                 \
    case MC_DIV: return F<MC_DIV>(__VA_ARGS__);                                                 \
    case MC_POW: return F<MC_POW>(__VA_ARGS__);                                                 \
    case MC_MAX: return F<MC_MAX>(__VA_ARGS__);                                                 \
    case MC_MIN: return F<MC_MIN>(__VA_ARGS__);                                                 \
    case MC_MOD: return F<MC_MOD>(__VA_ARGS__);                                                 \
    case MC_EQ: return F<MC_EQ>(__VA_ARGS__);                                                   \
    case MC_NE: return F<MC_NE>(__VA_ARGS__);                                                   \
    case MC_LT: return F<MC_LT>(__VA_ARGS__);                                                   \
    case MC_GT: return F<MC_GT>(__VA_ARGS__);                                                   \
    case MC_LE: return F<MC_LE>(__VA_ARGS__);                                                   \
    case MC_GE: return F<MC_GE>(__VA_ARGS__);                                                   \
    default: mc::set_error("unknown binary op %d", op); return 2;                               \
  }

extern "C" {

int mc_unary(int op, const float* x, float* y, int64_t n) {
  if (n <= 0) return 0;
  switch (op) {
    UNARY_CASE(MC_NEG) UNARY_CASE(MC_ABS) UNARY_CASE(MC_SIGN) UNARY_CASE(MC_ROUND) UNARY_CASE(MC_CEIL)
    UNARY_CASE(MC_FLOOR) UNARY_CASE(MC_SQRT) UNARY_CASE(MC_EXP) UNARY_CASE(MC_LOG) UNARY_CASE(MC_SIN)
    UNARY_CASE(MC_COS) UNARY_CASE(MC_TANH) UNARY_CASE(MC_SIGMOID) UNARY_CASE(MC_SQUARE) UNARY_CASE(MC_RECIP)
    UNARY_CASE(MC_RELU) UNARY_CASE(MC_COPY) UNARY_CASE(MC_COSH) UNARY_CASE(MC_LOGICAL_NOT)
    default: mc::set_error("unknown unary op %d", op); return 2;
  }
}

int mc_binary(int op, const float* a, const float* b, float* out, int64_t n) {
  if (n <= 0) return 0;
  BINARY_CASES(launch_binary, a, b, out, n)
}

int mc_binary_scalar(int op, const float* a, float s, int scalar_left, float* out, int64_t n) {
  if (n <= 0) return 0;
  BINARY_CASES(launch_binary_const, a, s, scalar_left, out, n)
}

int mc_binary_vec(int op, const float* a, const float* v, int vec_is_col, int vec_left, float* out,
                  int64_t rows, int64_t cols) {
  if (rows * cols <= 0) return 0;
  BINARY_CASES(launch_binary_vecb, a, v, vec_is_col, vec_left, out, rows, cols)
}

int mc_binary_bcast(int op, const float* a, const float* b, float* out, int ndim, const int64_t* out_shape,
                    const int64_t* a_strides, const int64_t* b_strides) {
  MC_REQUIRE(ndim >= 1 && ndim <= MC_MAX_DIMS, "ndim must be in 1..6");
  MC_REQUIRE(op >= 0 && op < MC_BINARY_COUNT, "unknown binary op");
  BcastDesc d;
  d.ndim = ndim;
  d.op = op;
  int64_t n = 1;
  for (int k = 0; k < MC_MAX_DIMS; ++k) {
    d.shape[k] = k < ndim ? out_shape[k] : 1;
    d.sa[k] = k < ndim ? a_strides[k] : 0;
    d.sb[k] = k < ndim ? b_strides[k] : 0;
    n *= d.shape[k];
  }
  if (n <= 0) return 0;
  int grid = grid_for(n, EW_THREADS);
  mc::launch(binary_bcast_kernel, grid, EW_THREADS, 0, a, b, out, n, d);
  MC_LAUNCHED();
  return 0;
}

int mc_fill(float* dst, float value, int64_t n) {
  if (n <= 0) return 0;
  int vec_ok = aligned16(dst);
  int grid = grid_for(vec_ok ? (n >> 2) + 1 : n, EW_THREADS);
  mc::launch(fill_kernel, grid, EW_THREADS, 0, dst, value, n, vec_ok);
  MC_LAUNCHED();
  return 0;
}

int mc_arange(float* dst, float start, float step, int64_t n) {
  if (n <= 0) return 0;
  mc::launch(arange_kernel, grid_for(n, EW_THREADS), EW_THREADS, 0, dst, start, step, n);
  MC_LAUNCHED();
  return 0;
}

int mc_cast(const void* src, int src_dtype, void* dst, int dst_dtype, int64_t n) {
  if (n <= 0) return 0;
  switch (src_dtype) {
    case MC_F32: return cast_from<float>((const float*)src, dst, dst_dtype, n);
    case MC_I32: return cast_from<int32_t>((const int32_t*)src, dst, dst_dtype, n);
    case MC_I64: return cast_from<int64_t>((const int64_t*)src, dst, dst_dtype, n);
    case MC_F64: return cast_from<double>((const double*)src, dst, dst_dtype, n);
    case MC_U8: return cast_from<uint8_t>((const uint8_t*)src, dst, dst_dtype, n);
    default: mc::set_error("unknown src dtype %d", src_dtype); return 2;
  }
}

int mc_vjp_chain(const float* g, int g_mode, int64_t g_inner, float g_scalar, const float* acc,
                 const mc_vjp_step* steps, int nsteps, float* out, int64_t n) {
  MC_REQUIRE(nsteps >= 0 && nsteps <= MC_VJP_MAX_STEPS, "too many chain steps");
  if (n <= 0) return 0;
  ChainDesc d;
  d.nsteps = nsteps;
  d.g_mode = g_mode;
  d.g_inner = g_inner > 0 ? g_inner : 1;
  d.g_scalar = g_scalar;
  int vec_ok = aligned16(out) && (!acc || aligned16(acc)) && (!g || aligned16(g));
  for (int k = 0; k < nsteps; ++k) {
    d.steps[k] = steps[k];
    MC_REQUIRE(steps[k].op >= 0 && steps[k].op < MC_VJP_COUNT, "unknown vjp op");
    if (d.steps[k].inner <= 0) d.steps[k].inner = 1;
    if (steps[k].a && !aligned16(steps[k].a)) vec_ok = 0;
    if (steps[k].b && !aligned16(steps[k].b)) vec_ok = 0;
  }
  int grid = grid_for(vec_ok ? (n >> 2) + 1 : n, EW_THREADS);
  mc::launch(vjp_chain_kernel, grid, EW_THREADS, 0, g, acc, out, n, vec_ok, d);
  MC_LAUNCHED();
  return 0;
}

int mc_multi_copy(const void* const* srcs, void* const* dsts, const int64_t* ns, int count) {
  MC_REQUIRE(count >= 0, "negative tensor count");
  for (int base = 0; base < count; base += MT_MAX) {
    MultiCopyDesc d;
    d.count = 0;
    int blocks = 0;
    for (int k = base; k < count && d.count < MT_MAX; ++k) {
      if (ns[k] <= 0) continue;
      d.first[d.count] = blocks;
      d.src[d.count] = (const float*)srcs[k];
      d.dst[d.count] = (float*)dsts[k];
      d.n[d.count] = ns[k];
      blocks += (int)((ns[k] + MT_TILE - 1) / MT_TILE);
      ++d.count;
    }
    if (d.count == 0) continue;
    d.first[d.count] = blocks;
    mc::launch(multi_copy_kernel, blocks, EW_THREADS, 0, d);
    MC_LAUNCHED();
  }
  return 0;
}

}  // extern "C"
