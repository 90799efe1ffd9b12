// lstm_temporal of nn/layers.py:383-420 as two native loops over time-major buffers.
//
// Per time step the reference runs lstm_step (layers.py:353-380): 2 dots, 4 slices, 3 sigmoids of 4
// primitives each, 2 tanh, adds / muls = 26 dispatches forward and 32 VJP records backward.  Here a
// step is TWO launches:
//   forward   K-sliced tensor-core GEMM  h_{t-1} . Wh  ->  raw partials ws[z] (no combine pass)
//             gate kernel reading  a_t + sum_z ws[z]  (a_t = x_t . Wx + b, computed for all steps at once)
//   backward  gate-backward kernel reading  dH[t+1] + sum_z ws[z]  (ws = partials of the step after it)
//             K-sliced GEMM  da_t . Wh^T  ->  ws
// i.e. the split-K combine (a launch plus a write and a read of the (N,4H) / (N,H) tile per step;
// 22% of the C5 step in profiles/r1z) is folded into the kernel that consumes the sum.
#include <initializer_list>
#include "common.cuh"

namespace mc {

int gemm_tf32_partials(int ta, int tb, int64_t M, int64_t N, int64_t K, const float* A, int64_t lda, const float* B,
                       int64_t ldb, float* ws, int max_splits, int* splits);   // gemm.cu

namespace tc {
struct LstmEpilogue {        // must match gemm.cu
  const float* ax; const float* c_prev; const float* c_next; float* h_out; float* c_out; float* gates;
  const float* dh_in; const float* dc_next; float* da; float* dc_out; int H;
};
}
int gemm_tf32_lstm(int mode, int64_t M, int64_t N, int64_t K, const float* A, int64_t lda, const float* B, int64_t ldb,
                   const tc::LstmEpilogue& epi, int* taken);                    // gemm.cu

int lstm_persist_fwd(const float* A, const float* Wh, float* hbuf, float* cbuf, float* gates, int64_t T, int64_t N, int64_t H,
                     int* taken);                                               // lstm_persist.cu
int lstm_persist_bwd(const float* gates, const float* cbuf, const float* Wh, float* dH, float* dA, int64_t T, int64_t N, int64_t H,
                     int* taken);

constexpr int LS_THREADS = 256;

// Wh (H, 4H) gate-major columns [i | f | o | g]  ->  unit-major columns: out[k, 4u + g] = in[k, g*H + u]
__global__ void __launch_bounds__(LS_THREADS) lstm_permute_wh_kernel(const float* __restrict__ in, float* __restrict__ out, int64_t H) {
  mc::pdl_sync();
  const int64_t n = H * H;                                 // one thread per (k, u): four loads, one float4 store
  const int64_t stride = (int64_t)gridDim.x * LS_THREADS;
  for (int64_t t = (int64_t)blockIdx.x * LS_THREADS + threadIdx.x; t < n; t += stride) {
    const int64_t k = t / H, u = t - k * H;
    const float* __restrict__ r = in + k * 4 * H + u;
    *reinterpret_cast<float4*>(out + k * 4 * H + 4 * u) = make_float4(r[0], r[H], r[2 * H], r[3 * H]);
  }
}

constexpr int LS_MAX_SPLITS = 16;

__device__ __forceinline__ float ls_sigmoid(float x) { return 1.0f / (1.0f + expf(-x)); }  // layers.py:243

// a (N,4H) + sum over `splits` slices of ws (each N*4H) -> gates i,f,o,g; c' = f*c + i*g; h' = o*tanh(c')
__global__ void __launch_bounds__(LS_THREADS) lstm_fwd_part_kernel(const float* __restrict__ a, const float* __restrict__ ws,
                                                                   int splits, const float* __restrict__ prev_c,
                                                                   float* __restrict__ next_h, float* __restrict__ next_c,
                                                                   float* __restrict__ gates, int64_t N, int64_t H) {
  mc::pdl_sync();
  const int64_t n = N * H, slice = n * 4;
  const int64_t stride = (int64_t)gridDim.x * LS_THREADS;
  for (int64_t t = (int64_t)blockIdx.x * LS_THREADS + threadIdx.x; t < n; t += stride) {
    const int64_t r = t / H, j = t % H;
    const int64_t base = r * 4 * H + j;
    float p0 = a[base], p1 = a[base + H], p2 = a[base + 2 * H], p3 = a[base + 3 * H];
    for (int z = 0; z < splits; ++z) {      // slice order: deterministic
      const float* __restrict__ w = ws + (int64_t)z * slice + base;
      p0 += __ldcg(w); p1 += __ldcg(w + H); p2 += __ldcg(w + 2 * H); p3 += __ldcg(w + 3 * H);
    }
    const float i = ls_sigmoid(p0), f = ls_sigmoid(p1), o = ls_sigmoid(p2), g = tanhf(p3);
    const float c = f * prev_c[t] + i * g;
    next_c[t] = c;
    next_h[t] = o * tanhf(c);
    gates[base] = i; gates[base + H] = f; gates[base + 2 * H] = o; gates[base + 3 * H] = g;
  }
}

// dh = dnext_h + sum over slices of ws (each N*H); the rest as lstm_bwd_kernel (layers.cu)
__global__ void __launch_bounds__(LS_THREADS) lstm_bwd_part_kernel(const float* __restrict__ gates, const float* __restrict__ prev_c,
                                                                   const float* __restrict__ next_c,
                                                                   const float* __restrict__ dnext_h,
                                                                   const float* __restrict__ ws, int splits,
                                                                   const float* __restrict__ dnext_c, float* __restrict__ da,
                                                                   float* __restrict__ dprev_c, int64_t N, int64_t H) {
  mc::pdl_sync();
  const int64_t n = N * H;
  const int64_t stride = (int64_t)gridDim.x * LS_THREADS;
  for (int64_t t = (int64_t)blockIdx.x * LS_THREADS + threadIdx.x; t < n; t += stride) {
    const int64_t r = t / H, j = t % H;
    const int64_t base = r * 4 * H + j;
    const float i = gates[base], f = gates[base + H], o = gates[base + 2 * H], g = gates[base + 3 * H];
    const float tc = tanhf(next_c[t]);
    float dh = dnext_h[t];
    for (int z = 0; z < splits; ++z) dh += __ldcg(ws + (int64_t)z * n + t);
    float dc = dh * o * (1.f - tc * tc);
    if (dnext_c) dc += dnext_c[t];
    da[base] = dc * g * i * (1.f - i);
    da[base + H] = dc * prev_c[t] * f * (1.f - f);
    da[base + 2 * H] = dh * tc * o * (1.f - o);
    da[base + 3 * H] = dc * i * (1.f - g * g);
    dprev_c[t] = dc * f;
  }
}

// 128-bit versions (H % 4 == 0, 16-byte aligned operands): four hidden units per thread, every operand a
// float4 load/store, the slices unrolled by the compiler -- same arithmetic per element as the scalar
// kernels above (which remain for odd H); the scalar versions ran at 0.9 TB/s on cold operands
// (profiles/r1z/prof_lstm_c5.txt), i.e. they were latency-, not bandwidth-bound.
constexpr int LS4_THREADS = 128;

__device__ __forceinline__ float4 ld4(const float* p) { return *reinterpret_cast<const float4*>(p); }
__device__ __forceinline__ float4 ld4cg(const float* p) { return __ldcg(reinterpret_cast<const float4*>(p)); }
__device__ __forceinline__ void st4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }
__device__ __forceinline__ void add4(float4& a, float4 b) { a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w; }

__global__ void __launch_bounds__(LS4_THREADS) lstm_fwd_part4_kernel(const float* __restrict__ a, const float* __restrict__ ws,
                                                                     int splits, const float* __restrict__ prev_c,
                                                                     float* __restrict__ next_h, float* __restrict__ next_c,
                                                                     float* __restrict__ gates, int64_t N, int64_t H) {
  mc::pdl_sync();
  const int64_t H4 = H >> 2, n4 = N * H4, slice = N * H * 4;
  const int64_t stride = (int64_t)gridDim.x * LS4_THREADS;
  for (int64_t t = (int64_t)blockIdx.x * LS4_THREADS + threadIdx.x; t < n4; t += stride) {
    const int64_t r = t / H4, j = (t - r * H4) << 2;
    const int64_t base = r * 4 * H + j, cell = r * H + j;
    float4 p0 = ld4(a + base), p1 = ld4(a + base + H), p2 = ld4(a + base + 2 * H), p3 = ld4(a + base + 3 * H);
    const float4 c0 = ld4(prev_c + cell);
    for (int z = 0; z < splits; ++z) {      // slice order: deterministic
      const float* __restrict__ w = ws + (int64_t)z * slice + base;
      const float4 q0 = ld4cg(w), q1 = ld4cg(w + H), q2 = ld4cg(w + 2 * H), q3 = ld4cg(w + 3 * H);
      add4(p0, q0); add4(p1, q1); add4(p2, q2); add4(p3, q3);
    }
    float4 gi, gf, go, gg, c, h;
#define LS_CELL(m)                                                                                   \
    gi.m = ls_sigmoid(p0.m); gf.m = ls_sigmoid(p1.m); go.m = ls_sigmoid(p2.m); gg.m = tanhf(p3.m);    \
    c.m = gf.m * c0.m + gi.m * gg.m; h.m = go.m * tanhf(c.m);
    LS_CELL(x) LS_CELL(y) LS_CELL(z) LS_CELL(w)
#undef LS_CELL
    st4(next_c + cell, c);
    st4(next_h + cell, h);
    st4(gates + base, gi); st4(gates + base + H, gf); st4(gates + base + 2 * H, go); st4(gates + base + 3 * H, gg);
  }
}

__global__ void __launch_bounds__(LS4_THREADS) lstm_bwd_part4_kernel(const float* __restrict__ gates, const float* __restrict__ prev_c,
                                                                     const float* __restrict__ next_c,
                                                                     const float* __restrict__ dnext_h,
                                                                     const float* __restrict__ ws, int splits,
                                                                     const float* __restrict__ dnext_c, float* __restrict__ da,
                                                                     float* __restrict__ dprev_c, int64_t N, int64_t H) {
  mc::pdl_sync();
  const int64_t H4 = H >> 2, n4 = N * H4, n = N * H;
  const int64_t stride = (int64_t)gridDim.x * LS4_THREADS;
  for (int64_t t = (int64_t)blockIdx.x * LS4_THREADS + threadIdx.x; t < n4; t += stride) {
    const int64_t r = t / H4, j = (t - r * H4) << 2;
    const int64_t base = r * 4 * H + j, cell = r * H + j;
    const float4 gi = ld4(gates + base), gf = ld4(gates + base + H), go = ld4(gates + base + 2 * H), gg = ld4(gates + base + 3 * H);
    const float4 cn = ld4(next_c + cell), cp = ld4(prev_c + cell);
    float4 dh = ld4(dnext_h + cell);
    for (int z = 0; z < splits; ++z) add4(dh, ld4cg(ws + (int64_t)z * n + cell));
    float4 dcn = make_float4(0.f, 0.f, 0.f, 0.f);
    if (dnext_c) dcn = ld4(dnext_c + cell);
    float4 d0, d1, d2, d3, dpc;
#define LS_CELL(m)                                                                       \
    {                                                                                    \
      const float tc = tanhf(cn.m);                                                      \
      float dc = dh.m * go.m * (1.f - tc * tc);                                          \
      if (dnext_c) dc += dcn.m;                                                          \
      d0.m = dc * gg.m * gi.m * (1.f - gi.m);                                            \
      d1.m = dc * cp.m * gf.m * (1.f - gf.m);                                            \
      d2.m = dh.m * tc * go.m * (1.f - go.m);                                            \
      d3.m = dc * gi.m * (1.f - gg.m * gg.m);                                            \
      dpc.m = dc * gf.m;                                                                 \
    }
    LS_CELL(x) LS_CELL(y) LS_CELL(z) LS_CELL(w)
#undef LS_CELL
    st4(da + base, d0); st4(da + base + H, d1); st4(da + base + 2 * H, d2); st4(da + base + 3 * H, d3);
    st4(dprev_c + cell, dpc);
  }
}

static inline bool ls_vec_ok(int64_t H, std::initializer_list<const void*> ptrs) {
  if (H & 3) return false;
  for (const void* p : ptrs) if (p && !aligned16(p)) return false;
  return true;
}

}  // namespace mc

using namespace mc;

extern "C" {

int mc_lstm_seq_fwd(float* A, const float* Wh, float* hbuf, float* cbuf, float* gates, int64_t T, int64_t N, int64_t H,
                    int precision) {
  MC_REQUIRE(T >= 0 && N > 0 && H > 0, "bad extents");
  const int64_t sH = N * H, sA = N * 4 * H;
  float* ws = nullptr;
  const bool try_partials = (precision == MC_GEMM_TF32);
  if (try_partials) {          // all T steps as ONE persistent launch with Wh resident in shared memory
    int taken = 0;
    if (lstm_persist_fwd(A, Wh, hbuf, cbuf, gates, T, N, H, &taken)) return 1;
    if (taken) return 0;
  }
  ScratchFrame scratch;
  if (try_partials) { ws = scratch.floats((size_t)(LS_MAX_SPLITS * sA)); if (!ws) return 1; }
  int rc = 0;
  const bool vec = ls_vec_ok(H, {A, ws, hbuf, cbuf, gates});
  // Fused step: ONE launch = recurrent GEMM against the unit-major copy of Wh + gate arithmetic in its epilogue
  float* whp = nullptr;
  bool fused = try_partials && vec && aligned16(Wh) && T > 0;
  if (fused) {
    whp = scratch.floats((size_t)(H * 4 * H));
    if (!whp) return 1;
    mc::launch(lstm_permute_wh_kernel, grid_for(H * H, LS_THREADS), LS_THREADS, 0, Wh, whp, H);
    MC_LAUNCHED();
  }
  for (int64_t t = 0; t < T && rc == 0; ++t) {
    float* a_t = A + t * sA;
    if (fused) {
      tc::LstmEpilogue e = {};
      e.ax = a_t; e.c_prev = cbuf + t * sH; e.h_out = hbuf + (t + 1) * sH; e.c_out = cbuf + (t + 1) * sH;
      e.gates = gates + t * sA; e.H = (int)H;
      int taken = 0;
      rc = gemm_tf32_lstm(1, N, 4 * H, H, hbuf + t * sH, H, whp, 4 * H, e, &taken);
      if (rc) break;
      if (taken) continue;
      fused = false;             // shape not taken: the two-launch path for the whole sequence
    }
    int splits = 0;
    if (try_partials) rc = gemm_tf32_partials(0, 0, N, 4 * H, H, hbuf + t * sH, H, Wh, 4 * H, ws, LS_MAX_SPLITS, &splits);
    if (rc) break;
    if (splits == 0) {   // exact-FP32 mode or a shape TMA cannot describe: the two public calls
      rc = mc_gemm(0, 0, N, 4 * H, H, hbuf + t * sH, H, Wh, 4 * H, a_t, 4 * H, nullptr, 0, 1, precision);
      if (rc) break;
    }
    if (vec)
      mc::launch(lstm_fwd_part4_kernel, grid_for(sH / 4, LS4_THREADS), LS4_THREADS, 0, 
          a_t, ws, splits, cbuf + t * sH, hbuf + (t + 1) * sH, cbuf + (t + 1) * sH, gates + t * sA, N, H);
    else
      mc::launch(lstm_fwd_part_kernel, grid_for(sH, LS_THREADS), LS_THREADS, 0, 
          a_t, ws, splits, cbuf + t * sH, hbuf + (t + 1) * sH, cbuf + (t + 1) * sH, gates + t * sA, N, H);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { set_error("lstm_fwd_part_kernel launch failed: %s", cudaGetErrorString(e)); rc = 1; break; }
    count_launch();
  }
  return rc;
}

int mc_lstm_seq_bwd(const float* gates, const float* cbuf, const float* Wh, float* dH, float* dA, int64_t T, int64_t N,
                    int64_t H, int precision) {
  MC_REQUIRE(T >= 0 && N > 0 && H > 0, "bad extents");
  const int64_t sH = N * H, sA = N * 4 * H;
  float *ws = nullptr, *dc = nullptr;
  const bool try_partials = (precision == MC_GEMM_TF32);
  if (try_partials) {          // all T steps as ONE persistent launch with Wh resident in shared memory
    int taken = 0;
    if (lstm_persist_bwd(gates, cbuf, Wh, dH, dA, T, N, H, &taken)) return 1;
    if (taken) return 0;
  }
  ScratchFrame scratch;
  if (try_partials) { ws = scratch.floats((size_t)(LS_MAX_SPLITS * sH)); if (!ws) return 1; }
  dc = scratch.floats((size_t)(2 * sH));
  if (!dc) return 1;
  int rc = 0, pending = 0;
  const bool vec = ls_vec_ok(H, {gates, cbuf, dH, dA, ws, dc});
  bool fused = try_partials && vec && aligned16(Wh);
  bool done_by_epilogue = false;      // the gate backward of step t was computed by the epilogue of GEMM(t + 1)
  for (int64_t t = T - 1; t >= 0 && rc == 0; --t) {
    const float* dnext_c = (t < T - 1) ? dc + ((t + 1) & 1) * sH : nullptr;
    if (fused) {
      if (!done_by_epilogue) {          // only the last time step: nothing flows in from a later one
        mc::launch(lstm_bwd_part4_kernel, grid_for(sH / 4, LS4_THREADS), LS4_THREADS, 0, gates + t * sA, cbuf + t * sH,
                   cbuf + (t + 1) * sH, dH + (t + 1) * sH, (const float*)nullptr, 0, dnext_c, dA + t * sA, dc + (t & 1) * sH,
                   N, H);
        MC_LAUNCHED();
      }
      done_by_epilogue = false;
      if (t > 0) {
        // dA_t . Wh^T (the recurrent part of dh_{t-1}) with the gate backward of step t-1 as its epilogue
        tc::LstmEpilogue e = {};
        e.gates = const_cast<float*>(gates) + (t - 1) * sA; e.c_prev = cbuf + (t - 1) * sH; e.c_next = cbuf + t * sH;
        e.dh_in = dH + t * sH; e.dc_next = dc + (t & 1) * sH; e.da = dA + (t - 1) * sA; e.dc_out = dc + ((t - 1) & 1) * sH;
        e.H = (int)H;
        int taken = 0;
        rc = gemm_tf32_lstm(2, N, H, 4 * H, dA + t * sA, 4 * H, Wh, 4 * H, e, &taken);
        if (rc) break;
        if (taken) { done_by_epilogue = true; continue; }
        fused = false;                // not taken: fall through to the two-launch path from this step on
        pending = 0;
      } else {
        rc = mc_gemm(0, 1, N, H, 4 * H, dA, 4 * H, Wh, 4 * H, dH, H, nullptr, 0, 1, precision);   // dh0 += dA_0 . Wh^T
        continue;
      }
      // (fall-through only when the fused GEMM was not taken at this t: dA[t] is computed, run its plain GEMM)
      if (try_partials) {
        rc = gemm_tf32_partials(0, 1, N, H, 4 * H, dA + t * sA, 4 * H, Wh, 4 * H, ws, LS_MAX_SPLITS, &pending);
        if (rc) break;
      }
      if (pending == 0) rc = mc_gemm(0, 1, N, H, 4 * H, dA + t * sA, 4 * H, Wh, 4 * H, dH + t * sH, H, nullptr, 0, 1, precision);
      continue;
    }
    if (vec)
      mc::launch(lstm_bwd_part4_kernel, grid_for(sH / 4, LS4_THREADS), LS4_THREADS, 0, 
          gates + t * sA, cbuf + t * sH, cbuf + (t + 1) * sH, dH + (t + 1) * sH, ws, pending, dnext_c, dA + t * sA,
          dc + (t & 1) * sH, N, H);
    else
      mc::launch(lstm_bwd_part_kernel, grid_for(sH, LS_THREADS), LS_THREADS, 0, 
          gates + t * sA, cbuf + t * sH, cbuf + (t + 1) * sH, dH + (t + 1) * sH, ws, pending, dnext_c, dA + t * sA,
          dc + (t & 1) * sH, N, H);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { set_error("lstm_bwd_part_kernel launch failed: %s", cudaGetErrorString(e)); rc = 1; break; }
    count_launch();
    pending = 0;
    // dH[t] += da_t . Wh^T: left as partials for the next gate kernel, except the last one (dh0 is read
    // by the caller, so it is combined in memory)
    if (try_partials && t > 0) {
      rc = gemm_tf32_partials(0, 1, N, H, 4 * H, dA + t * sA, 4 * H, Wh, 4 * H, ws, LS_MAX_SPLITS, &pending);
      if (rc) break;
    }
    if (pending == 0) rc = mc_gemm(0, 1, N, H, 4 * H, dA + t * sA, 4 * H, Wh, 4 * H, dH + t * sH, H, nullptr, 0, 1, precision);
  }
  return rc;
}

}  // extern "C"
