// Fused nn.layers / nn.optim kernels and the Philox RNG.
//
// Everything here is HBM-bound.  BatchNorm (nn/layers.py:93-110: 22 primitive dispatches
// forward, 23 VJP records backward in the reference) becomes two kernels per direction:
// a column "partial statistics" pass and a normalise / input-gradient pass that both
// use the same (128 columns x row-chunk) tiling, float4 loads along the contiguous
// feature axis, and a deterministic workspace combine instead of atomics.
#include "common.cuh"

namespace mc {

constexpr int L_THREADS = 256;

// ===================================== Philox4x32-10 ==================================
struct Philox {
  uint32_t k0, k1;
  __device__ Philox(uint64_t seed) : k0((uint32_t)seed), k1((uint32_t)(seed >> 32)) {}
  __device__ __forceinline__ uint4 operator()(uint64_t ctr) const {
    uint32_t c0 = (uint32_t)ctr, c1 = (uint32_t)(ctr >> 32), c2 = 0x243F6A88u, c3 = 0x85A308D3u;
    uint32_t a = k0, b = k1;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
      const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
      const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
      const uint32_t n0 = hi1 ^ c1 ^ a, n1 = lo1, n2 = hi0 ^ c3 ^ b, n3 = lo0;
      c0 = n0; c1 = n1; c2 = n2; c3 = n3;
      a += 0x9E3779B9u; b += 0xBB67AE85u;
    }
    return make_uint4(c0, c1, c2, c3);
  }
};
__device__ __forceinline__ float u01(uint32_t r) { return (float)(r >> 8) * (1.0f / 16777216.0f); }  // [0,1)

__global__ void __launch_bounds__(L_THREADS) rand_uniform_kernel(float* __restrict__ out, int64_t n, uint64_t seed,
                                                                 uint64_t offset, const unsigned long long* epoch) {
  mc::pdl_sync();
  if (epoch) offset += (uint64_t)(*epoch) << 40;     // graph replay r draws from its own counter range
  const Philox rng(seed);
  const int64_t n4 = (n + 3) >> 2;
  const int64_t stride = (int64_t)gridDim.x * L_THREADS;
  for (int64_t i4 = (int64_t)blockIdx.x * L_THREADS + threadIdx.x; i4 < n4; i4 += stride) {
    const uint4 r = rng(offset + (uint64_t)i4);
    const float v[4] = {u01(r.x), u01(r.y), u01(r.z), u01(r.w)};
    const int64_t i = i4 << 2;
#pragma unroll
    for (int k = 0; k < 4; ++k) if (i + k < n) out[i + k] = v[k];
  }
}

__global__ void __launch_bounds__(L_THREADS) rand_normal_kernel(float* __restrict__ out, int64_t n, uint64_t seed,
                                                                uint64_t offset, const unsigned long long* epoch) {
  mc::pdl_sync();
  if (epoch) offset += (uint64_t)(*epoch) << 40;
  const Philox rng(seed);
  const int64_t n4 = (n + 3) >> 2;
  const int64_t stride = (int64_t)gridDim.x * L_THREADS;
  for (int64_t i4 = (int64_t)blockIdx.x * L_THREADS + threadIdx.x; i4 < n4; i4 += stride) {
    const uint4 r = rng(offset + (uint64_t)i4);
    // Box-Muller on (0,1] x [0,1)
    const float u1 = 1.0f - u01(r.x), u2 = u01(r.y), u3 = 1.0f - u01(r.z), u4 = u01(r.w);
    const float ra = sqrtf(-2.0f * logf(u1)), rb = sqrtf(-2.0f * logf(u3));
    float s1, c1, s2, c2;
    sincospif(2.0f * u2, &s1, &c1);
    sincospif(2.0f * u4, &s2, &c2);
    const float v[4] = {ra * c1, ra * s1, rb * c2, rb * s2};
    const int64_t i = i4 << 2;
#pragma unroll
    for (int k = 0; k < 4; ++k) if (i + k < n) out[i + k] = v[k];
  }
}

// out = x * (u > p); the (seed, offset) pair regenerates the identical mask for the VJP.
__global__ void __launch_bounds__(L_THREADS) dropout_kernel(const float* __restrict__ x, float* __restrict__ out, int64_t n,
                                                            float p, uint64_t seed, uint64_t offset, int vec_ok,
                                                            const unsigned long long* epoch) {
  mc::pdl_sync();
  if (epoch) offset += (uint64_t)(*epoch) << 40;     // same shift in forward and VJP of one replay
  const Philox rng(seed);
  const int64_t n4 = (n + 3) >> 2;
  const int64_t stride = (int64_t)gridDim.x * L_THREADS;
  for (int64_t i4 = (int64_t)blockIdx.x * L_THREADS + threadIdx.x; i4 < n4; i4 += stride) {
    const uint4 r = rng(offset + (uint64_t)i4);
    const int64_t i = i4 << 2;
    if (vec_ok && i + 3 < n) {
      float4 v = *reinterpret_cast<const float4*>(x + i);
      v.x = (u01(r.x) > p) ? v.x : 0.f;
      v.y = (u01(r.y) > p) ? v.y : 0.f;
      v.z = (u01(r.z) > p) ? v.z : 0.f;
      v.w = (u01(r.w) > p) ? v.w : 0.f;
      *reinterpret_cast<float4*>(out + i) = v;
    } else {
      const float u[4] = {u01(r.x), u01(r.y), u01(r.z), u01(r.w)};
#pragma unroll
      for (int k = 0; k < 4; ++k) if (i + k < n) out[i + k] = (u[k] > p) ? x[i + k] : 0.f;
    }
  }
}

// ===================================== BatchNorm ======================================
// Tile: 32 threads x VEC columns wide, 8 row lanes; grid = (col_tiles, row_chunks).
// Partial layout: part[(chunk * 2 + which) * D + col].
template <int VEC>
struct ColVec {
  float v[VEC];
};
template <int VEC>
__device__ __forceinline__ ColVec<VEC> ld_cols(const float* __restrict__ p) {
  ColVec<VEC> r;
  if (VEC == 4) {
    const float4 t = *reinterpret_cast<const float4*>(p);
    r.v[0] = t.x; r.v[1] = t.y; r.v[2] = t.z; r.v[3] = t.w;
  } else {
#pragma unroll
    for (int k = 0; k < VEC; ++k) r.v[k] = p[k];
  }
  return r;
}
template <int VEC>
__device__ __forceinline__ void st_cols(float* __restrict__ p, const ColVec<VEC>& r) {
  if (VEC == 4) *reinterpret_cast<float4*>(p) = make_float4(r.v[0], r.v[1], r.v[2], r.v[3]);
  else {
#pragma unroll
    for (int k = 0; k < VEC; ++k) p[k] = r.v[k];
  }
}

// BWD == false: s1 = sum(x - x[0,:]), s2 = sum((x - x[0,:])^2)   (shifted sums: no cancellation blow-up)
// BWD == true : s1 = sum(dy),         s2 = sum(dy * xhat)
template <int VEC, bool BWD>
__global__ void __launch_bounds__(L_THREADS) bn_partial_kernel(const float* __restrict__ x, const float* __restrict__ dy,
                                                               const float* __restrict__ mean,
                                                               const float* __restrict__ invstd, float* __restrict__ part,
                                                               int64_t N, int64_t D, int64_t rows_per_chunk) {
  mc::pdl_sync();
  __shared__ float sm1[8][32 * VEC + 1];
  __shared__ float sm2[8][32 * VEC + 1];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int64_t col = ((int64_t)blockIdx.x * 32 + tx) * VEC;
  const int64_t r0 = (int64_t)blockIdx.y * rows_per_chunk;
  const int64_t r1 = min(N, r0 + rows_per_chunk);
  float a1[VEC], a2[VEC];
#pragma unroll
  for (int k = 0; k < VEC; ++k) a1[k] = a2[k] = 0.f;
  if (col < D) {
    ColVec<VEC> c0, c1;
    if (BWD) { c0 = ld_cols<VEC>(mean + col); c1 = ld_cols<VEC>(invstd + col); }
    else { c0 = ld_cols<VEC>(x + col); }
    for (int64_t r = r0 + ty; r < r1; r += 8) {
      const ColVec<VEC> xv = ld_cols<VEC>(x + r * D + col);
      if (BWD) {
        const ColVec<VEC> gv = ld_cols<VEC>(dy + r * D + col);
#pragma unroll
        for (int k = 0; k < VEC; ++k) { a1[k] += gv.v[k]; a2[k] += gv.v[k] * ((xv.v[k] - c0.v[k]) * c1.v[k]); }
      } else {
#pragma unroll
        for (int k = 0; k < VEC; ++k) { const float d = xv.v[k] - c0.v[k]; a1[k] += d; a2[k] += d * d; }
      }
    }
  }
#pragma unroll
  for (int k = 0; k < VEC; ++k) { sm1[ty][tx * VEC + k] = a1[k]; sm2[ty][tx * VEC + k] = a2[k]; }
  __syncthreads();
  if (ty == 0 && col < D) {
#pragma unroll
    for (int k = 0; k < VEC; ++k) {
      float s1 = 0.f, s2 = 0.f;
#pragma unroll
      for (int j = 0; j < 8; ++j) { s1 += sm1[j][tx * VEC + k]; s2 += sm2[j][tx * VEC + k]; }
      part[((int64_t)blockIdx.y * 2 + 0) * D + col + k] = s1;
      part[((int64_t)blockIdx.y * 2 + 1) * D + col + k] = s2;
    }
  }
}

// Column sums of the per-chunk partials, shared by both apply kernels: the 8 row lanes of the block
// each add every 8th chunk (vector loads, independent chains), the eight results are combined in
// a fixed order through shared memory.  The first version had every thread walk all chunks
// serially (2 * VEC dependent load chains of length `chunks`): 11 of the kernel's 11.3 us at
// (1024, 512).  Every thread of the block must call this (it synchronises).
template <int VEC>
__device__ __forceinline__ void bn_sum_partials(const float* __restrict__ part, int chunks, int64_t D, int64_t col, bool live,
                                                float (&s1)[VEC], float (&s2)[VEC]) {
  __shared__ float red1[8][32 * VEC + 1];
  __shared__ float red2[8][32 * VEC + 1];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  float p1[VEC], p2[VEC];
#pragma unroll
  for (int k = 0; k < VEC; ++k) p1[k] = p2[k] = 0.f;
  if (live) {
    for (int c = ty; c < chunks; c += 8) {
      const ColVec<VEC> a = ld_cols<VEC>(part + ((int64_t)c * 2 + 0) * D + col);
      const ColVec<VEC> b = ld_cols<VEC>(part + ((int64_t)c * 2 + 1) * D + col);
#pragma unroll
      for (int k = 0; k < VEC; ++k) { p1[k] += a.v[k]; p2[k] += b.v[k]; }
    }
  }
#pragma unroll
  for (int k = 0; k < VEC; ++k) { red1[ty][tx * VEC + k] = p1[k]; red2[ty][tx * VEC + k] = p2[k]; }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < VEC; ++k) {
    float a = 0.f, b = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) { a += red1[j][tx * VEC + k]; b += red2[j][tx * VEC + k]; }
    s1[k] = a; s2[k] = b;
  }
}

template <int VEC>
__global__ void __launch_bounds__(L_THREADS) bn_fwd_apply_kernel(const float* __restrict__ x, const float* __restrict__ gamma,
                                                                 const float* __restrict__ beta,
                                                                 const float* __restrict__ rmean,
                                                                 const float* __restrict__ rvar, const float* __restrict__ part,
                                                                 int chunks, float* __restrict__ out,
                                                                 float* __restrict__ new_rmean, float* __restrict__ new_rvar,
                                                                 float* __restrict__ save_mean, float* __restrict__ save_invstd,
                                                                 int64_t N, int64_t D, int64_t rows_per_chunk, float eps,
                                                                 float momentum) {
  mc::pdl_sync();
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int64_t col = ((int64_t)blockIdx.x * 32 + tx) * VEC;
  float t1[VEC], t2[VEC];
  bn_sum_partials<VEC>(part, chunks, D, col, col < D, t1, t2);
  if (col >= D) return;
  float mean[VEC], inv[VEC], sc[VEC], sh[VEC];
  const ColVec<VEC> x0 = ld_cols<VEC>(x + col);
  const ColVec<VEC> g = ld_cols<VEC>(gamma + col), b = ld_cols<VEC>(beta + col);
#pragma unroll
  for (int k = 0; k < VEC; ++k) {
    const float s1 = t1[k], s2 = t2[k];
    const float dm = s1 / (float)N;                 // mean - x0
    float var = s2 / (float)N - dm * dm;            // biased variance (layers.py:103)
    var = fmaxf(var, 0.f);
    mean[k] = x0.v[k] + dm;
    inv[k] = 1.0f / sqrtf(var + eps);
    sc[k] = g.v[k] * inv[k];
    sh[k] = b.v[k] - mean[k] * sc[k];
    if (blockIdx.y == 0 && ty == 0) {
      save_mean[col + k] = mean[k];
      save_invstd[col + k] = inv[k];
      new_rmean[col + k] = momentum * rmean[col + k] + (1.0f - momentum) * mean[k];
      new_rvar[col + k] = momentum * rvar[col + k] + (1.0f - momentum) * var;
    }
  }
  const int64_t r0 = (int64_t)blockIdx.y * rows_per_chunk;
  const int64_t r1 = min(N, r0 + rows_per_chunk);
  for (int64_t r = r0 + ty; r < r1; r += 8) {
    ColVec<VEC> xv = ld_cols<VEC>(x + r * D + col);
#pragma unroll
    for (int k = 0; k < VEC; ++k) xv.v[k] = fmaf(xv.v[k], sc[k], sh[k]);
    st_cols<VEC>(out + r * D + col, xv);
  }
}

template <int VEC>
__global__ void __launch_bounds__(L_THREADS) bn_bwd_apply_kernel(const float* __restrict__ x, const float* __restrict__ dy,
                                                                 const float* __restrict__ gamma,
                                                                 const float* __restrict__ mean,
                                                                 const float* __restrict__ invstd,
                                                                 const float* __restrict__ part, int chunks,
                                                                 float* __restrict__ dx, float* __restrict__ dgamma,
                                                                 float* __restrict__ dbeta, int64_t N, int64_t D,
                                                                 int64_t rows_per_chunk) {
  mc::pdl_sync();
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int64_t col = ((int64_t)blockIdx.x * 32 + tx) * VEC;
  float t1[VEC], t2[VEC];
  bn_sum_partials<VEC>(part, chunks, D, col, col < D, t1, t2);
  if (col >= D) return;
  const ColVec<VEC> g = ld_cols<VEC>(gamma + col), mu = ld_cols<VEC>(mean + col), iv = ld_cols<VEC>(invstd + col);
  float k1[VEC], k2[VEC], k3[VEC];  // dx = k1 * (dy - k2 - xhat * k3)
#pragma unroll
  for (int k = 0; k < VEC; ++k) {
    const float s1 = t1[k], s2 = t2[k];
    if (blockIdx.y == 0 && ty == 0) { dbeta[col + k] = s1; dgamma[col + k] = s2; }
    k1[k] = g.v[k] * iv.v[k];
    k2[k] = s1 / (float)N;
    k3[k] = s2 / (float)N;
  }
  const int64_t r0 = (int64_t)blockIdx.y * rows_per_chunk;
  const int64_t r1 = min(N, r0 + rows_per_chunk);
  for (int64_t r = r0 + ty; r < r1; r += 8) {
    const ColVec<VEC> xv = ld_cols<VEC>(x + r * D + col);
    ColVec<VEC> gv = ld_cols<VEC>(dy + r * D + col);
#pragma unroll
    for (int k = 0; k < VEC; ++k) {
      const float xhat = (xv.v[k] - mu.v[k]) * iv.v[k];
      gv.v[k] = k1[k] * (gv.v[k] - k2[k] - xhat * k3[k]);
    }
    st_cols<VEC>(dx + r * D + col, gv);
  }
}

// ---- small batches: statistics + apply in ONE kernel ("slab" kernels) ----------------------
// (1024, 512) is 2 MB: the two-kernel path above costs two launches (~4-6 us each inside a step
// graph) for ~1 us of memory time.  Here a block owns 4 * HALVES columns of every row and keeps its
// whole slab in registers: thread = (row lane, column half), up to N / lanes rows each.  HALVES = 2
// (8 columns = one 32-byte sector per row, 128 row lanes x 8 rows) when that still fills the machine,
// HALVES = 1 (4 columns, 256 row lanes x 4 rows) otherwise: twice the blocks, half the serial work in
// each.  Sums: per thread, then the row lanes of a warp by shuffle, then 8 warps through shared
// memory -- a fixed order, so results are deterministic.
// The kernels are latency chains, not bandwidth: the phase probe (tools/probe.py, profiles/r2l) showed
// 1.25 us of the forward kernel's 4.4 us between the barrier and the first store -- gamma / beta / running
// statistics were loaded only then -- and 1.5 us of Philox serialised behind the loads in both kernels.
// Every parameter is now fetched at entry and the keep-mask is computed while the slab loads are in flight.
constexpr int SLAB_MAXN = 1024;                // rows a slab kernel can hold
MC_PROBE_DEFINE(bnf)
MC_PROBE_DEFINE(bnb)
// Loads the optimiser may not sink towards their first use (it moved the parameter loads behind ~500 instructions
// of Philox, exposing their latency after the barrier): volatile asm keeps them where they are written.
__device__ __forceinline__ float4 ldg4_here(const float* p) {
  float4 v;
  asm volatile("ld.global.nc.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ float ldg_here(const float* p) {
  float v;
  asm volatile("ld.global.nc.f32 %0, [%1];" : "=f"(v) : "l"(p));
  return v;
}
// 1 / sqrt(v) to within an ulp without the IEEE sqrt + divide sequences (two dependent ~40-instruction chains in
// every thread of the block, on the critical path between the barrier and the first store).
__device__ __forceinline__ float rsqrt_nr(float v) {
  const float y = rsqrtf(v);
  return y * fmaf(-0.5f * v * y, y, 1.5f);
}

template <int HALVES>
__device__ __forceinline__ void slab_block_sums(float (&a)[4], float (&b)[4], float (&ta)[4], float (&tb)[4]) {
  __shared__ float red[8][2][8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, half = (HALVES == 2) ? (threadIdx.x & 1) : 0;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
#pragma unroll
    for (int o = HALVES; o < 32; o <<= 1) { a[k] += __shfl_xor_sync(0xffffffffu, a[k], o); b[k] += __shfl_xor_sync(0xffffffffu, b[k], o); }
    if (lane < HALVES) { red[warp][half][k] = a[k]; red[warp][half][4 + k] = b[k]; }
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    float x = 0.f, y = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) { x += red[w][half][k]; y += red[w][half][4 + k]; }
    ta[k] = x; tb[k] = y;
  }
}

// Keep-mask of the thread's slab: bit 4 * i + k is set when element (row i of this thread, column k) survives the
// dropout -- lane (e & 3) of Philox(seed)(offset + e / 4) for flat element index e, exactly dropout_kernel's.
template <int RPT, int LANES>
__device__ __forceinline__ uint32_t slab_keep_mask(const Philox& rng, uint64_t offset, int rl, int N, int64_t D, int64_t col,
                                                   float p) {
  uint32_t keep = 0;
#pragma unroll
  for (int i = 0; i < RPT; ++i) {
    const int r = rl + LANES * i;
    if (r < N) {
      const uint4 q = rng(offset + (uint64_t)(((int64_t)r * D + col) >> 2));
      keep |= ((u01(q.x) > p ? 1u : 0u) | (u01(q.y) > p ? 2u : 0u) | (u01(q.z) > p ? 4u : 0u) | (u01(q.w) > p ? 8u : 0u)) << (4 * i);
    }
  }
  return keep;
}

// DROP: the dropout that follows batchnorm in the reference's block (mlp_bn_dropout.py:41-54) is applied
// at the store -- out = bn(x) * (u > p) -- so the normalised activations are never written unmasked.
template <bool DROP, int HALVES>
__global__ void __launch_bounds__(L_THREADS) bn_slab_fwd_kernel(const float* __restrict__ x, const float* __restrict__ gamma,
                                                                const float* __restrict__ beta, const float* __restrict__ rmean,
                                                                const float* __restrict__ rvar, float* __restrict__ out,
                                                                float* __restrict__ new_rmean, float* __restrict__ new_rvar,
                                                                float* __restrict__ save_mean, float* __restrict__ save_invstd,
                                                                int N, int64_t D, float eps, float momentum, float p,
                                                                uint64_t seed, uint64_t offset, const unsigned long long* epoch) {
  constexpr int LANES = L_THREADS / HALVES, RPT = SLAB_MAXN / LANES;
  MC_STAMP(bnf, 0);
  mc::pdl_sync();
  MC_STAMP(bnf, 1);
  if (DROP && epoch) offset += (uint64_t)(*epoch) << 40;
  const int half = (HALVES == 2) ? (threadIdx.x & 1) : 0, rl = threadIdx.x / HALVES;
  const int64_t col = (int64_t)blockIdx.x * (4 * HALVES) + 4 * half;
  float4 v[RPT];
#pragma unroll
  for (int i = 0; i < RPT; ++i) {
    const int r = rl + LANES * i;
    if (r < N) v[i] = *reinterpret_cast<const float4*>(x + (int64_t)r * D + col);
  }
  const float4 x0 = ldg4_here(x + col);                              // shift: no cancellation blow-up
  float gm[4], bt[4], rm[4] = {0.f, 0.f, 0.f, 0.f}, rv[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    gm[k] = ldg_here(gamma + col + k);
    bt[k] = ldg_here(beta + col + k);
    if (rl == 0) { rm[k] = ldg_here(rmean + col + k); rv[k] = ldg_here(rvar + col + k); }
  }
  const float inv_n = 1.0f / (float)N;
  uint32_t keep = 0xffffffffu;
  if (DROP) keep = slab_keep_mask<RPT, LANES>(Philox(seed), offset, rl, N, D, col, p);
  float a[4] = {0.f, 0.f, 0.f, 0.f}, b[4] = {0.f, 0.f, 0.f, 0.f}, s1[4], s2[4];
#pragma unroll
  for (int i = 0; i < RPT; ++i) {
    if (rl + LANES * i < N) {
      const float d0 = v[i].x - x0.x, d1 = v[i].y - x0.y, d2 = v[i].z - x0.z, d3 = v[i].w - x0.w;
      a[0] += d0; a[1] += d1; a[2] += d2; a[3] += d3;
      b[0] += d0 * d0; b[1] += d1 * d1; b[2] += d2 * d2; b[3] += d3 * d3;
    }
  }
  MC_STAMP(bnf, 2);
  slab_block_sums<HALVES>(a, b, s1, s2);
  MC_STAMP(bnf, 3);
  const float x0v[4] = {x0.x, x0.y, x0.z, x0.w};
  float sc[4], sh[4], mean[4], inv[4], var[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const float dm = s1[k] * inv_n;
    var[k] = fmaxf(s2[k] * inv_n - dm * dm, 0.f);                     // biased variance (layers.py:103)
    mean[k] = x0v[k] + dm;
    inv[k] = rsqrt_nr(var[k] + eps);
    sc[k] = gm[k] * inv[k];
    sh[k] = bt[k] - mean[k] * sc[k];
  }
  if (rl == 0) {                                                      // one divergent region, after the arithmetic
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      save_mean[col + k] = mean[k];
      save_invstd[col + k] = inv[k];
      new_rmean[col + k] = momentum * rm[k] + (1.0f - momentum) * mean[k];
      new_rvar[col + k] = momentum * rv[k] + (1.0f - momentum) * var[k];
    }
  }
  MC_STAMP(bnf, 4);
#pragma unroll
  for (int i = 0; i < RPT; ++i) {
    const int r = rl + LANES * i;
    if (r < N) {
      float4 o = make_float4(fmaf(v[i].x, sc[0], sh[0]), fmaf(v[i].y, sc[1], sh[1]), fmaf(v[i].z, sc[2], sh[2]), fmaf(v[i].w, sc[3], sh[3]));
      if (DROP) {
        const uint32_t m = keep >> (4 * i);
        o.x = (m & 1u) ? o.x : 0.f; o.y = (m & 2u) ? o.y : 0.f;
        o.z = (m & 4u) ? o.z : 0.f; o.w = (m & 8u) ? o.w : 0.f;
      }
      *reinterpret_cast<float4*>(out + (int64_t)r * D + col) = o;
    }
  }
  MC_STAMP(bnf, 5);
}

// DROP: dy is the gradient of the dropout that followed this batchnorm; its keep-mask is regenerated from
// (seed, offset) while dy is loaded.  RELU: x is the output of a fused affine+relu (signed-zero mode, see
// relu_apply): dx is zeroed where x's sign bit is set, i.e. the result is already the gradient of the affine
// PRE-activation, and its column sums (the bias gradient of that affine layer) are written to `dbias`.
// One launch instead of dropout-bwd + batchnorm-bwd + relu-bwd + a column reduction.
template <bool DROP, bool RELU, int HALVES>
__global__ void __launch_bounds__(L_THREADS) bn_slab_bwd_kernel(const float* __restrict__ x, const float* __restrict__ dy,
                                                                const float* __restrict__ gamma, const float* __restrict__ mean,
                                                                const float* __restrict__ invstd, float* __restrict__ dx,
                                                                float* __restrict__ dgamma, float* __restrict__ dbeta,
                                                                float* __restrict__ dbias, int N, int64_t D, float p,
                                                                uint64_t seed, uint64_t offset, const unsigned long long* epoch) {
  constexpr int LANES = L_THREADS / HALVES, RPT = SLAB_MAXN / LANES;
  MC_STAMP(bnb, 0);
  mc::pdl_sync();
  MC_STAMP(bnb, 1);
  if (DROP && epoch) offset += (uint64_t)(*epoch) << 40;
  uint32_t neg = 0;            // RELU: bit 4*i+k set when x[row i of this thread][column k] has its sign bit set
  const int half = (HALVES == 2) ? (threadIdx.x & 1) : 0, rl = threadIdx.x / HALVES;
  const int64_t col = (int64_t)blockIdx.x * (4 * HALVES) + 4 * half;
  float4 xh[RPT], g[RPT];
#pragma unroll
  for (int i = 0; i < RPT; ++i) {
    const int r = rl + LANES * i;
    if (r < N) {
      xh[i] = *reinterpret_cast<const float4*>(x + (int64_t)r * D + col);
      g[i] = *reinterpret_cast<const float4*>(dy + (int64_t)r * D + col);
    }
  }
  const float4 mu = ldg4_here(mean + col), iv = ldg4_here(invstd + col);
  float gm[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) gm[k] = ldg_here(gamma + col + k);
  const float inv_n = 1.0f / (float)N;
  uint32_t keep = 0xffffffffu;
  if (DROP) keep = slab_keep_mask<RPT, LANES>(Philox(seed), offset, rl, N, D, col, p);
  float a[4] = {0.f, 0.f, 0.f, 0.f}, b[4] = {0.f, 0.f, 0.f, 0.f}, s1[4], s2[4];
#pragma unroll
  for (int i = 0; i < RPT; ++i) {
    if (rl + LANES * i < N) {
      if (DROP) {
        const uint32_t m = keep >> (4 * i);
        g[i].x = (m & 1u) ? g[i].x : 0.f; g[i].y = (m & 2u) ? g[i].y : 0.f;
        g[i].z = (m & 4u) ? g[i].z : 0.f; g[i].w = (m & 8u) ? g[i].w : 0.f;
      }
      if (RELU)
        neg |= ((__float_as_uint(xh[i].x) >> 31) | ((__float_as_uint(xh[i].y) >> 31) << 1) |
                ((__float_as_uint(xh[i].z) >> 31) << 2) | ((__float_as_uint(xh[i].w) >> 31) << 3)) << (4 * i);
      xh[i] = make_float4((xh[i].x - mu.x) * iv.x, (xh[i].y - mu.y) * iv.y, (xh[i].z - mu.z) * iv.z, (xh[i].w - mu.w) * iv.w);
      a[0] += g[i].x; a[1] += g[i].y; a[2] += g[i].z; a[3] += g[i].w;
      b[0] += g[i].x * xh[i].x; b[1] += g[i].y * xh[i].y; b[2] += g[i].z * xh[i].z; b[3] += g[i].w * xh[i].w;
    }
  }
  MC_STAMP(bnb, 2);
  slab_block_sums<HALVES>(a, b, s1, s2);
  MC_STAMP(bnb, 3);
  const float ivv[4] = {iv.x, iv.y, iv.z, iv.w};
  float k1[4], k2[4], k3[4];                       // dx = k1 * (dy - k2 - xhat * k3)
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    k1[k] = gm[k] * ivv[k];
    k2[k] = s1[k] * inv_n;
    k3[k] = s2[k] * inv_n;
  }
  if (rl == 0) {
#pragma unroll
    for (int k = 0; k < 4; ++k) { dbeta[col + k] = s1[k]; dgamma[col + k] = s2[k]; }
  }
  float c[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
  for (int i = 0; i < RPT; ++i) {
    const int r = rl + LANES * i;
    if (r < N) {
      float4 o = make_float4(k1[0] * (g[i].x - k2[0] - xh[i].x * k3[0]), k1[1] * (g[i].y - k2[1] - xh[i].y * k3[1]),
                             k1[2] * (g[i].z - k2[2] - xh[i].z * k3[2]), k1[3] * (g[i].w - k2[3] - xh[i].w * k3[3]));
      if (RELU) {
        const uint32_t m = neg >> (4 * i);
        if (m & 1u) o.x = 0.f;
        if (m & 2u) o.y = 0.f;
        if (m & 4u) o.z = 0.f;
        if (m & 8u) o.w = 0.f;
        c[0] += o.x; c[1] += o.y; c[2] += o.z; c[3] += o.w;
      }
      *reinterpret_cast<float4*>(dx + (int64_t)r * D + col) = o;
    }
  }
  MC_STAMP(bnb, 4);
  if (RELU) {
    float z[4] = {0.f, 0.f, 0.f, 0.f}, t1[4], t2[4];
    __syncthreads();                  // slab_block_sums reuses its shared buffer
    slab_block_sums<HALVES>(c, z, t1, t2);
    if (rl == 0) {
#pragma unroll
      for (int k = 0; k < 4; ++k) dbias[col + k] = t1[k];
    }
  }
  MC_STAMP(bnb, 5);
}
// 8 columns a block while that gives every SM a block, 4 columns otherwise.
static int bn_slab_halves(int64_t D) { return (D / 8 >= sm_count()) ? 2 : 1; }
static bool bn_slab_ok(int64_t N, int64_t D, bool aligned) {
  return aligned && N <= SLAB_MAXN && (D % 8) == 0 && N * D <= (1 << 20);
}

template <int VEC>
__global__ void __launch_bounds__(L_THREADS) bn_test_kernel(const float* __restrict__ x, const float* __restrict__ gamma,
                                                            const float* __restrict__ beta, const float* __restrict__ rmean,
                                                            const float* __restrict__ rvar, float* __restrict__ out, int64_t N,
                                                            int64_t D, int64_t rows_per_chunk, float eps) {
  mc::pdl_sync();
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int64_t col = ((int64_t)blockIdx.x * 32 + tx) * VEC;
  if (col >= D) return;
  const ColVec<VEC> g = ld_cols<VEC>(gamma + col), b = ld_cols<VEC>(beta + col);
  const ColVec<VEC> m = ld_cols<VEC>(rmean + col), v = ld_cols<VEC>(rvar + col);
  float sc[VEC], sh[VEC];
#pragma unroll
  for (int k = 0; k < VEC; ++k) { sc[k] = g.v[k] / sqrtf(v.v[k] + eps); sh[k] = b.v[k] - m.v[k] * sc[k]; }
  const int64_t r0 = (int64_t)blockIdx.y * rows_per_chunk;
  const int64_t r1 = min(N, r0 + rows_per_chunk);
  for (int64_t r = r0 + ty; r < r1; r += 8) {
    ColVec<VEC> xv = ld_cols<VEC>(x + r * D + col);
#pragma unroll
    for (int k = 0; k < VEC; ++k) xv.v[k] = fmaf(xv.v[k], sc[k], sh[k]);
    st_cols<VEC>(out + r * D + col, xv);
  }
}

struct BnTiling {
  int vec;
  int64_t col_tiles, chunks, rows_per_chunk;
};
static BnTiling bn_tiling(bool ptrs_aligned, int64_t N, int64_t D) {
  BnTiling t;
  t.vec = ((D % 4 == 0) && ptrs_aligned) ? 4 : 1;
  t.col_tiles = (D + 32 * t.vec - 1) / (32 * t.vec);
  int64_t want = (2 * (int64_t)sm_count() + t.col_tiles - 1) / t.col_tiles;  // ~2 CTAs per SM
  int64_t maxc = (N + 31) / 32;                                             // >= 32 rows per chunk
  t.chunks = want < maxc ? want : maxc;
  if (t.chunks < 1) t.chunks = 1;
  if (t.chunks > 4096) t.chunks = 4096;
  t.rows_per_chunk = (N + t.chunks - 1) / t.chunks;
  t.chunks = (N + t.rows_per_chunk - 1) / t.rows_per_chunk;
  return t;
}

// ===================================== LSTM cell ======================================
__device__ __forceinline__ float sigmoid_ref(float x) { return 1.0f / (1.0f + expf(-x)); }  // layers.py:243

__global__ void __launch_bounds__(L_THREADS) lstm_fwd_kernel(const float* __restrict__ a, const float* __restrict__ prev_c,
                                                             float* __restrict__ next_h, float* __restrict__ next_c,
                                                             float* __restrict__ gates, int64_t N, int64_t H) {
  mc::pdl_sync();
  const int64_t n = N * H;
  const int64_t stride = (int64_t)gridDim.x * L_THREADS;
  for (int64_t t = (int64_t)blockIdx.x * L_THREADS + threadIdx.x; t < n; t += stride) {
    const int64_t r = t / H, j = t % H;
    const float* __restrict__ ar = a + r * 4 * H;
    const float i = sigmoid_ref(ar[j]), f = sigmoid_ref(ar[H + j]), o = sigmoid_ref(ar[2 * H + j]);
    const float g = tanhf(ar[3 * H + j]);
    const float c = f * prev_c[t] + i * g;
    next_c[t] = c;
    next_h[t] = o * tanhf(c);
    float* __restrict__ gr = gates + r * 4 * H;
    gr[j] = i; gr[H + j] = f; gr[2 * H + j] = o; gr[3 * H + j] = g;
  }
}

__global__ void __launch_bounds__(L_THREADS) lstm_bwd_kernel(const float* __restrict__ gates, const float* __restrict__ prev_c,
                                                             const float* __restrict__ next_c,
                                                             const float* __restrict__ dnext_h,
                                                             const float* __restrict__ dnext_c, float* __restrict__ da,
                                                             float* __restrict__ dprev_c, int64_t N, int64_t H) {
  mc::pdl_sync();
  const int64_t n = N * H;
  const int64_t stride = (int64_t)gridDim.x * L_THREADS;
  for (int64_t t = (int64_t)blockIdx.x * L_THREADS + threadIdx.x; t < n; t += stride) {
    const int64_t r = t / H, j = t % H;
    const float* __restrict__ gr = gates + r * 4 * H;
    const float i = gr[j], f = gr[H + j], o = gr[2 * H + j], g = gr[3 * H + j];
    const float tc = tanhf(next_c[t]);
    const float dh = dnext_h ? dnext_h[t] : 0.f;
    float dc = dh * o * (1.f - tc * tc);
    if (dnext_c) dc += dnext_c[t];
    float* __restrict__ dr = da + r * 4 * H;
    dr[j] = dc * g * i * (1.f - i);
    dr[H + j] = dc * prev_c[t] * f * (1.f - f);
    dr[2 * H + j] = dh * tc * o * (1.f - o);
    dr[3 * H + j] = dc * i * (1.f - g * g);
    dprev_c[t] = dc * f;
  }
}

// ===================================== optimizers =====================================
__global__ void __launch_bounds__(L_THREADS) sgd_kernel(const float* __restrict__ w, const float* __restrict__ dw,
                                                        float* __restrict__ wo, float lr, int64_t n) {
  mc::pdl_sync();
  const int64_t stride = (int64_t)gridDim.x * L_THREADS;
  for (int64_t i = (int64_t)blockIdx.x * L_THREADS + threadIdx.x; i < n; i += stride) wo[i] = w[i] - lr * dw[i];
}
__global__ void __launch_bounds__(L_THREADS) sgd_momentum_kernel(const float* __restrict__ w, const float* __restrict__ dw,
                                                                 const float* __restrict__ v, float* __restrict__ wo,
                                                                 float* __restrict__ vo, float lr, float mu, int64_t n) {
  mc::pdl_sync();
  const int64_t stride = (int64_t)gridDim.x * L_THREADS;
  for (int64_t i = (int64_t)blockIdx.x * L_THREADS + threadIdx.x; i < n; i += stride) {
    const float nv = (v ? v[i] : 0.f) * mu - dw[i] * lr;  // optim.py:64
    vo[i] = nv;
    wo[i] = w[i] + nv;
  }
}
__global__ void __launch_bounds__(L_THREADS) rmsprop_kernel(const float* __restrict__ w, const float* __restrict__ dw,
                                                            const float* __restrict__ cache, float* __restrict__ wo,
                                                            float* __restrict__ co, float lr, float decay, float eps,
                                                            int64_t n) {
  mc::pdl_sync();
  const int64_t stride = (int64_t)gridDim.x * L_THREADS;
  for (int64_t i = (int64_t)blockIdx.x * L_THREADS + threadIdx.x; i < n; i += stride) {
    const float g = dw[i];
    const float c = (cache ? cache[i] : 0.f) * decay + g * g * (1.f - decay);  // optim.py:93
    co[i] = c;
    wo[i] = w[i] - lr * g / (sqrtf(c) + eps);
  }
}
__global__ void __launch_bounds__(L_THREADS) adam_kernel(const float* __restrict__ w, const float* __restrict__ dw,
                                                         const float* __restrict__ m, const float* __restrict__ v,
                                                         float* __restrict__ wo, float* __restrict__ mo, float* __restrict__ vo,
                                                         float lr, float b1, float b2, float eps, float bc1, float bc2,
                                                         int64_t n) {
  mc::pdl_sync();
  const int64_t stride = (int64_t)gridDim.x * L_THREADS;
  for (int64_t i = (int64_t)blockIdx.x * L_THREADS + threadIdx.x; i < n; i += stride) {
    const float g = dw[i];
    const float nm = b1 * (m ? m[i] : 0.f) + (1.f - b1) * g;        // optim.py:128-129
    const float nv = b2 * (v ? v[i] : 0.f) + (1.f - b2) * (g * g);
    mo[i] = nm;
    vo[i] = nv;
    wo[i] = w[i] - lr * (nm / bc1) / (sqrtf(nv / bc2) + eps);      // optim.py:130-134
  }
}

}  // namespace mc

using namespace mc;

extern "C" {

int mc_rand_uniform(float* out, int64_t n, uint64_t seed, uint64_t offset) {
  if (n <= 0) return 0;
  mc::launch(rand_uniform_kernel, grid_for((n + 3) / 4, L_THREADS), L_THREADS, 0, out, n, seed, offset, graph_epoch());
  MC_LAUNCHED();
  return 0;
}
int mc_rand_normal(float* out, int64_t n, uint64_t seed, uint64_t offset) {
  if (n <= 0) return 0;
  mc::launch(rand_normal_kernel, grid_for((n + 3) / 4, L_THREADS), L_THREADS, 0, out, n, seed, offset, graph_epoch());
  MC_LAUNCHED();
  return 0;
}
int mc_dropout(const float* x, float* out, int64_t n, float p, uint64_t seed, uint64_t offset) {
  if (n <= 0) return 0;
  const int vec_ok = aligned16(x) && aligned16(out);
  mc::launch(dropout_kernel, grid_for((n + 3) / 4, L_THREADS), L_THREADS, 0, x, out, n, p, seed, offset, vec_ok, graph_epoch());
  MC_LAUNCHED();
  return 0;
}

int mc_batchnorm_train_fwd(const float* x, const float* gamma, const float* beta, const float* running_mean,
                           const float* running_var, float* out, float* new_running_mean, float* new_running_var,
                           float* save_mean, float* save_invstd, int64_t N, int64_t D, float eps, float momentum) {
  MC_REQUIRE(N > 0 && D > 0, "empty batch");
  if (bn_slab_ok(N, D, aligned16(x) && aligned16(out))) {
    if (bn_slab_halves(D) == 2)
      mc::launch(bn_slab_fwd_kernel<false, 2>, (unsigned)(D / 8), L_THREADS, 0, x, gamma, beta, running_mean, running_var, out,
                 new_running_mean, new_running_var, save_mean, save_invstd, (int)N, D, eps, momentum, 0.f, (uint64_t)0,
                 (uint64_t)0, (const unsigned long long*)nullptr);
    else
      mc::launch(bn_slab_fwd_kernel<false, 1>, (unsigned)(D / 4), L_THREADS, 0, x, gamma, beta, running_mean, running_var, out,
                 new_running_mean, new_running_var, save_mean, save_invstd, (int)N, D, eps, momentum, 0.f, (uint64_t)0,
                 (uint64_t)0, (const unsigned long long*)nullptr);
    MC_LAUNCHED();
    return 0;
  }
  const BnTiling t = bn_tiling(aligned16(x) && aligned16(out) && aligned16(gamma) && aligned16(beta), N, D);
  ScratchFrame scratch;
  float* part = scratch.floats((size_t)(t.chunks * 2 * D));
  if (!part) return 1;
  dim3 grid((unsigned)t.col_tiles, (unsigned)t.chunks);
  if (t.vec == 4) {
    mc::launch(bn_partial_kernel<4, false>, grid, L_THREADS, 0, x, nullptr, nullptr, nullptr, part, N, D, t.rows_per_chunk);
    MC_LAUNCHED();
    mc::launch(bn_fwd_apply_kernel<4>, grid, L_THREADS, 0, x, gamma, beta, running_mean, running_var, part, (int)t.chunks,
                                                            out, new_running_mean, new_running_var, save_mean, save_invstd,
                                                            N, D, t.rows_per_chunk, eps, momentum);
  } else {
    mc::launch(bn_partial_kernel<1, false>, grid, L_THREADS, 0, x, nullptr, nullptr, nullptr, part, N, D, t.rows_per_chunk);
    MC_LAUNCHED();
    mc::launch(bn_fwd_apply_kernel<1>, grid, L_THREADS, 0, x, gamma, beta, running_mean, running_var, part, (int)t.chunks,
                                                            out, new_running_mean, new_running_var, save_mean, save_invstd,
                                                            N, D, t.rows_per_chunk, eps, momentum);
  }
  MC_LAUNCHED();
  return 0;
}

int mc_bn_fused_supported(int64_t N, int64_t D, int* ok) {
  *ok = bn_slab_ok(N, D, true) ? 1 : 0;
  return 0;
}

int mc_bn_dropout_fwd(const float* x, const float* gamma, const float* beta, const float* running_mean,
                      const float* running_var, float* out, float* new_running_mean, float* new_running_var,
                      float* save_mean, float* save_invstd, int64_t N, int64_t D, float eps, float momentum, float p,
                      uint64_t seed, uint64_t offset) {
  MC_REQUIRE(N > 0 && D > 0, "empty batch");
  MC_REQUIRE(bn_slab_ok(N, D, aligned16(x) && aligned16(out)), "shape outside the fused kernel (mc_bn_fused_supported)");
  if (bn_slab_halves(D) == 2)
    mc::launch(bn_slab_fwd_kernel<true, 2>, (unsigned)(D / 8), L_THREADS, 0, x, gamma, beta, running_mean, running_var, out,
               new_running_mean, new_running_var, save_mean, save_invstd, (int)N, D, eps, momentum, p, seed, offset,
               (const unsigned long long*)graph_epoch());
  else
    mc::launch(bn_slab_fwd_kernel<true, 1>, (unsigned)(D / 4), L_THREADS, 0, x, gamma, beta, running_mean, running_var, out,
               new_running_mean, new_running_var, save_mean, save_invstd, (int)N, D, eps, momentum, p, seed, offset,
               (const unsigned long long*)graph_epoch());
  MC_LAUNCHED();
  return 0;
}

int mc_bn_dropout_relu_bwd(const float* x, const float* dy, const float* gamma, const float* save_mean,
                           const float* save_invstd, float* dx, float* dgamma, float* dbeta, float* dbias, int64_t N,
                           int64_t D, int dropout, float p, uint64_t seed, uint64_t offset, int relu_signed) {
  MC_REQUIRE(N > 0 && D > 0, "empty batch");
  MC_REQUIRE(bn_slab_ok(N, D, aligned16(x) && aligned16(dy) && aligned16(dx) && aligned16(save_mean) && aligned16(save_invstd)),
             "shape outside the fused kernel (mc_bn_fused_supported)");
  MC_REQUIRE(!relu_signed || dbias != nullptr, "relu_signed needs dbias");
  const unsigned long long* ep = (const unsigned long long*)graph_epoch();
  const int halves = bn_slab_halves(D);
  const unsigned grid = (unsigned)(D / (4 * halves));
#define MC_BN_BWD(DR, RE)                                                                                            \
  if (halves == 2)                                                                                                   \
    mc::launch(bn_slab_bwd_kernel<DR, RE, 2>, grid, L_THREADS, 0, x, dy, gamma, save_mean, save_invstd, dx, dgamma, dbeta, \
               dbias, (int)N, D, p, seed, offset, ep);                                                               \
  else                                                                                                               \
    mc::launch(bn_slab_bwd_kernel<DR, RE, 1>, grid, L_THREADS, 0, x, dy, gamma, save_mean, save_invstd, dx, dgamma, dbeta, \
               dbias, (int)N, D, p, seed, offset, ep)
  if (dropout && relu_signed) { MC_BN_BWD(true, true); }
  else if (dropout) { MC_BN_BWD(true, false); }
  else if (relu_signed) { MC_BN_BWD(false, true); }
  else { MC_BN_BWD(false, false); }
#undef MC_BN_BWD
  MC_LAUNCHED();
  return 0;
}

int mc_batchnorm_train_bwd(const float* x, const float* dy, const float* gamma, const float* save_mean,
                           const float* save_invstd, float* dx, float* dgamma, float* dbeta, int64_t N, int64_t D) {
  MC_REQUIRE(N > 0 && D > 0, "empty batch");
  if (bn_slab_ok(N, D, aligned16(x) && aligned16(dy) && aligned16(dx) && aligned16(save_mean) && aligned16(save_invstd))) {
    if (bn_slab_halves(D) == 2)
      mc::launch(bn_slab_bwd_kernel<false, false, 2>, (unsigned)(D / 8), L_THREADS, 0, x, dy, gamma, save_mean, save_invstd, dx,
                 dgamma, dbeta, (float*)nullptr, (int)N, D, 0.f, (uint64_t)0, (uint64_t)0, (const unsigned long long*)nullptr);
    else
      mc::launch(bn_slab_bwd_kernel<false, false, 1>, (unsigned)(D / 4), L_THREADS, 0, x, dy, gamma, save_mean, save_invstd, dx,
                 dgamma, dbeta, (float*)nullptr, (int)N, D, 0.f, (uint64_t)0, (uint64_t)0, (const unsigned long long*)nullptr);
    MC_LAUNCHED();
    return 0;
  }
  const BnTiling t = bn_tiling(aligned16(x) && aligned16(dy) && aligned16(dx) && aligned16(gamma), N, D);
  ScratchFrame scratch;
  float* part = scratch.floats((size_t)(t.chunks * 2 * D));
  if (!part) return 1;
  dim3 grid((unsigned)t.col_tiles, (unsigned)t.chunks);
  if (t.vec == 4) {
    mc::launch(bn_partial_kernel<4, true>, grid, L_THREADS, 0, x, dy, save_mean, save_invstd, part, N, D, t.rows_per_chunk);
    MC_LAUNCHED();
    mc::launch(bn_bwd_apply_kernel<4>, grid, L_THREADS, 0, x, dy, gamma, save_mean, save_invstd, part, (int)t.chunks, dx,
                                                            dgamma, dbeta, N, D, t.rows_per_chunk);
  } else {
    mc::launch(bn_partial_kernel<1, true>, grid, L_THREADS, 0, x, dy, save_mean, save_invstd, part, N, D, t.rows_per_chunk);
    MC_LAUNCHED();
    mc::launch(bn_bwd_apply_kernel<1>, grid, L_THREADS, 0, x, dy, gamma, save_mean, save_invstd, part, (int)t.chunks, dx,
                                                            dgamma, dbeta, N, D, t.rows_per_chunk);
  }
  MC_LAUNCHED();
  return 0;
}

int mc_batchnorm_test_fwd(const float* x, const float* gamma, const float* beta, const float* running_mean,
                          const float* running_var, float* out, int64_t N, int64_t D, float eps) {
  MC_REQUIRE(N > 0 && D > 0, "empty batch");
  const BnTiling t = bn_tiling(aligned16(x) && aligned16(out) && aligned16(gamma) && aligned16(beta), N, D);
  dim3 grid((unsigned)t.col_tiles, (unsigned)t.chunks);
  if (t.vec == 4) mc::launch(bn_test_kernel<4>, grid, L_THREADS, 0, x, gamma, beta, running_mean, running_var, out, N, D, t.rows_per_chunk, eps);
  else mc::launch(bn_test_kernel<1>, grid, L_THREADS, 0, x, gamma, beta, running_mean, running_var, out, N, D, t.rows_per_chunk, eps);
  MC_LAUNCHED();
  return 0;
}

int mc_lstm_cell_fwd(const float* a, const float* prev_c, float* next_h, float* next_c, float* gates, int64_t N, int64_t H) {
  if (N * H <= 0) return 0;
  mc::launch(lstm_fwd_kernel, grid_for(N * H, L_THREADS), L_THREADS, 0, a, prev_c, next_h, next_c, gates, N, H);
  MC_LAUNCHED();
  return 0;
}
int mc_lstm_cell_bwd(const float* gates, const float* prev_c, const float* next_c, const float* dnext_h,
                     const float* dnext_c, float* da, float* dprev_c, int64_t N, int64_t H) {
  if (N * H <= 0) return 0;
  mc::launch(lstm_bwd_kernel, grid_for(N * H, L_THREADS), L_THREADS, 0, gates, prev_c, next_c, dnext_h, dnext_c, da, dprev_c, N, H);
  MC_LAUNCHED();
  return 0;
}

int mc_sgd(const float* w, const float* dw, float* w_out, float lr, int64_t n) {
  if (n <= 0) return 0;
  mc::launch(sgd_kernel, grid_for(n, L_THREADS), L_THREADS, 0, w, dw, w_out, lr, n);
  MC_LAUNCHED();
  return 0;
}
int mc_sgd_momentum(const float* w, const float* dw, const float* v, float* w_out, float* v_out, float lr, float momentum,
                    int64_t n) {
  if (n <= 0) return 0;
  mc::launch(sgd_momentum_kernel, grid_for(n, L_THREADS), L_THREADS, 0, w, dw, v, w_out, v_out, lr, momentum, n);
  MC_LAUNCHED();
  return 0;
}
int mc_rmsprop(const float* w, const float* dw, const float* cache, float* w_out, float* cache_out, float lr, float decay,
               float eps, int64_t n) {
  if (n <= 0) return 0;
  mc::launch(rmsprop_kernel, grid_for(n, L_THREADS), L_THREADS, 0, w, dw, cache, w_out, cache_out, lr, decay, eps, n);
  MC_LAUNCHED();
  return 0;
}
int mc_adam(const float* w, const float* dw, const float* m, const float* v, float* w_out, float* m_out, float* v_out,
            float lr, float beta1, float beta2, float eps, int64_t t, int64_t n) {
  if (n <= 0) return 0;
  const float bc1 = (float)(1.0 - pow((double)beta1, (double)t));
  const float bc2 = (float)(1.0 - pow((double)beta2, (double)t));
  mc::launch(adam_kernel, grid_for(n, L_THREADS), L_THREADS, 0, w, dw, m, v, w_out, m_out, v_out, lr, beta1, beta2, eps, bc1,
                                                                bc2, n);
  MC_LAUNCHED();
  return 0;
}

}  // extern "C"

// ---------------------------------------------------------------- multi-tensor update ------
// One launch updates up to MU_MAX parameter tensors (the per-parameter loop of
// nn/solver.py:190-195 made one kernel): block b works on tile (b - first[t]) of tensor t.
namespace mc {
constexpr int MU_MAX = 32;
constexpr int MU_TILE = L_THREADS * 8;
struct MultiUpdateDesc {
  int count, rule;
  float h[6];                   // lr, (momentum | decay | beta1), (eps | beta2), eps, bc1, bc2
  int first[MU_MAX + 1];
  int64_t n[MU_MAX];
  const float* w[MU_MAX];
  const float* dw[MU_MAX];
  const float* s1[MU_MAX];      // velocity / cache / m   (nullptr = zeros)
  const float* s2[MU_MAX];      // v (adam)
  float* wo[MU_MAX];
  float* s1o[MU_MAX];
  float* s2o[MU_MAX];
  const unsigned long long* epoch;   // graph replay counter (adam: t = t0 + epoch - 1) or nullptr
  int64_t t0;
};

__global__ void __launch_bounds__(L_THREADS) multi_update_kernel(const __grid_constant__ MultiUpdateDesc d) {
  mc::pdl_sync();
  int t = 0;
  while (t + 1 < d.count && (int)blockIdx.x >= d.first[t + 1]) ++t;
  const float* __restrict__ w = d.w[t];
  const float* __restrict__ dw = d.dw[t];
  const float* __restrict__ s1 = d.s1[t];
  const float* __restrict__ s2 = d.s2[t];
  float* __restrict__ wo = d.wo[t];
  float* __restrict__ s1o = d.s1o[t];
  float* __restrict__ s2o = d.s2o[t];
  const int64_t lo = (int64_t)(blockIdx.x - d.first[t]) * MU_TILE;
  const int64_t hi = (lo + MU_TILE < d.n[t]) ? lo + MU_TILE : d.n[t];
  const float lr = d.h[0];
  float bc1 = d.h[4], bc2 = d.h[5];
  if (d.rule == 3 && d.epoch) {
    const double t = (double)(d.t0 + (int64_t)(*d.epoch) - 1);
    bc1 = (float)(1.0 - pow((double)d.h[1], t));
    bc2 = (float)(1.0 - pow((double)d.h[2], t));
  }
  // whole tile, every stream 16-byte aligned: two float4 per thread and stream, all loads issued first
  const uintptr_t align_bits = (uintptr_t)w | (uintptr_t)dw | (uintptr_t)s1 | (uintptr_t)s2 | (uintptr_t)wo | (uintptr_t)s1o |
                               (uintptr_t)s2o;
  if (hi - lo == MU_TILE && (align_bits & 15u) == 0 && d.rule <= 1) {
    const int64_t i0 = lo + 4 * threadIdx.x, i1 = i0 + 4 * L_THREADS;
    const float4 g0 = *reinterpret_cast<const float4*>(dw + i0), g1 = *reinterpret_cast<const float4*>(dw + i1);
    const float4 w0 = *reinterpret_cast<const float4*>(w + i0), w1 = *reinterpret_cast<const float4*>(w + i1);
    if (d.rule == 0) {
      *reinterpret_cast<float4*>(wo + i0) = make_float4(w0.x - lr * g0.x, w0.y - lr * g0.y, w0.z - lr * g0.z, w0.w - lr * g0.w);
      *reinterpret_cast<float4*>(wo + i1) = make_float4(w1.x - lr * g1.x, w1.y - lr * g1.y, w1.z - lr * g1.z, w1.w - lr * g1.w);
    } else {
      float4 v0 = make_float4(0.f, 0.f, 0.f, 0.f), v1 = v0;
      if (s1) { v0 = *reinterpret_cast<const float4*>(s1 + i0); v1 = *reinterpret_cast<const float4*>(s1 + i1); }
      const float mu = d.h[1];
      v0 = make_float4(v0.x * mu - g0.x * lr, v0.y * mu - g0.y * lr, v0.z * mu - g0.z * lr, v0.w * mu - g0.w * lr);
      v1 = make_float4(v1.x * mu - g1.x * lr, v1.y * mu - g1.y * lr, v1.z * mu - g1.z * lr, v1.w * mu - g1.w * lr);
      *reinterpret_cast<float4*>(s1o + i0) = v0;
      *reinterpret_cast<float4*>(s1o + i1) = v1;
      *reinterpret_cast<float4*>(wo + i0) = make_float4(w0.x + v0.x, w0.y + v0.y, w0.z + v0.z, w0.w + v0.w);
      *reinterpret_cast<float4*>(wo + i1) = make_float4(w1.x + v1.x, w1.y + v1.y, w1.z + v1.z, w1.w + v1.w);
    }
    return;
  }
  // rmsprop / adam, whole aligned tile: the same two-float4 pattern, the same per-element expressions as the scalar
  // loop below (the scalar loop ran the C5 adam update, 4.85 M parameters, at 2.4 TB/s: 58 us)
  if (hi - lo == MU_TILE && (align_bits & 15u) == 0) {
    const int64_t idx[2] = {lo + 4 * threadIdx.x, lo + 4 * threadIdx.x + 4 * L_THREADS};
    float4 g4[2], w4[2], a4[2], b4[2];
    const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      g4[u] = *reinterpret_cast<const float4*>(dw + idx[u]);
      w4[u] = *reinterpret_cast<const float4*>(w + idx[u]);
      a4[u] = s1 ? *reinterpret_cast<const float4*>(s1 + idx[u]) : z4;
      b4[u] = (d.rule == 3 && s2) ? *reinterpret_cast<const float4*>(s2 + idx[u]) : z4;
    }
    const float h1 = d.h[1], h2 = d.h[2], h3 = d.h[3];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      float4 wn, an, bn = z4;
#define MU_CELL(m)                                                                      \
      if (d.rule == 2) {                                                                \
        const float c = a4[u].m * h1 + g4[u].m * g4[u].m * (1.f - h1);                  \
        an.m = c;                                                                       \
        wn.m = w4[u].m - lr * g4[u].m / (sqrtf(c) + h2);                                \
      } else {                                                                          \
        const float nm = h1 * a4[u].m + (1.f - h1) * g4[u].m;                           \
        const float nv = h2 * b4[u].m + (1.f - h2) * (g4[u].m * g4[u].m);               \
        an.m = nm; bn.m = nv;                                                           \
        wn.m = w4[u].m - lr * (nm / bc1) / (sqrtf(nv / bc2) + h3);                      \
      }
      MU_CELL(x) MU_CELL(y) MU_CELL(z) MU_CELL(w)
#undef MU_CELL
      *reinterpret_cast<float4*>(s1o + idx[u]) = an;
      if (d.rule == 3) *reinterpret_cast<float4*>(s2o + idx[u]) = bn;
      *reinterpret_cast<float4*>(wo + idx[u]) = wn;
    }
    return;
  }
  for (int64_t i = lo + threadIdx.x; i < hi; i += L_THREADS) {
    const float g = dw[i];
    if (d.rule == 0) {
      wo[i] = w[i] - lr * g;                                                     // optim.py:33-46
    } else if (d.rule == 1) {
      const float nv = (s1 ? s1[i] : 0.f) * d.h[1] - g * lr;                     // optim.py:64
      s1o[i] = nv;
      wo[i] = w[i] + nv;
    } else if (d.rule == 2) {
      const float c = (s1 ? s1[i] : 0.f) * d.h[1] + g * g * (1.f - d.h[1]);      // optim.py:93
      s1o[i] = c;
      wo[i] = w[i] - lr * g / (sqrtf(c) + d.h[2]);
    } else {
      const float nm = d.h[1] * (s1 ? s1[i] : 0.f) + (1.f - d.h[1]) * g;         // optim.py:128-134
      const float nv = d.h[2] * (s2 ? s2[i] : 0.f) + (1.f - d.h[2]) * (g * g);
      s1o[i] = nm;
      s2o[i] = nv;
      wo[i] = w[i] - lr * (nm / bc1) / (sqrtf(nv / bc2) + d.h[3]);
    }
  }
}
}  // namespace mc

extern "C" {

int mc_multi_update(int rule, int count, const void* const* w, const void* const* dw, const void* const* s1,
                    const void* const* s2, void* const* w_out, void* const* s1_out, void* const* s2_out, const int64_t* ns,
                    const float* hyper, int64_t t) {
  MC_REQUIRE(rule >= 0 && rule <= 3 && count >= 0, "bad rule / count");
  for (int base = 0; base < count; base += mc::MU_MAX) {
    mc::MultiUpdateDesc d;
    d.count = 0;
    d.rule = rule;
    d.epoch = mc::graph_epoch();
    d.t0 = t;
    for (int k = 0; k < 4; ++k) d.h[k] = hyper[k];
    d.h[4] = d.h[5] = 1.f;
    if (rule == 3) {
      d.h[4] = (float)(1.0 - pow((double)hyper[1], (double)t));
      d.h[5] = (float)(1.0 - pow((double)hyper[2], (double)t));
    }
    int blocks = 0;
    for (int k = base; k < count && d.count < mc::MU_MAX; ++k) {
      if (ns[k] <= 0) continue;
      const int j = d.count++;
      d.first[j] = blocks;
      d.n[j] = ns[k];
      d.w[j] = (const float*)w[k]; d.dw[j] = (const float*)dw[k];
      d.s1[j] = s1 ? (const float*)s1[k] : nullptr; d.s2[j] = s2 ? (const float*)s2[k] : nullptr;
      d.wo[j] = (float*)w_out[k];
      d.s1o[j] = s1_out ? (float*)s1_out[k] : nullptr; d.s2o[j] = s2_out ? (float*)s2_out[k] : nullptr;
      blocks += (int)((ns[k] + mc::MU_TILE - 1) / mc::MU_TILE);
    }
    if (d.count == 0) continue;
    d.first[d.count] = blocks;
    mc::launch(mc::multi_update_kernel, blocks, mc::L_THREADS, 0, d);
    MC_LAUNCHED();
  }
  return 0;
}

}  // extern "C"
