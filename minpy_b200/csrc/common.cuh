// Shared helpers for libminpy_cuda.so (sm_100a).  Internal -- the public surface is
// include/minpy_cuda.h.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include "../../include/minpy_cuda.h"

namespace mc {

// ---- per-thread error string + library-wide stream / counters (runtime.cu) ----------
void set_error(const char* fmt, ...);
cudaStream_t stream();
int sm_count();
int device();
int side_index();        // 1 while library calls are routed to the side stream (mc_side_begin .. mc_side_end), else 0
// true exactly once per (mask, current device): per-device one-time set-up such as
// cudaFuncSetAttribute, which applies to the current device only
bool first_use_on_device(uint32_t* mask);
void count_launch(int n = 1);
// Non-null while a step graph is being captured (mc_graph_begin .. mc_graph_end): device counter
// that the graph increments once per replay (1 on the first).  Kernels whose host-side
// arguments change every step (Philox offsets, adam's t) read it so that a replay is a NEW
// step, not a repetition of the captured one.
const unsigned long long* graph_epoch();

#define MC_CHECK(expr)                                                              \
  do {                                                                              \
    cudaError_t _e = (expr);                                                        \
    if (_e != cudaSuccess) {                                                        \
      mc::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
      return 1;                                                                     \
    }                                                                               \
  } while (0)

#define MC_REQUIRE(cond, msg)                                                       \
  do {                                                                              \
    if (!(cond)) {                                                                  \
      mc::set_error("%s: requirement failed: %s", __func__, msg);                   \
      return 2;                                                                     \
    }                                                                               \
  } while (0)

// After a kernel launch: catch launch-configuration errors, count the launch.
#define MC_LAUNCHED()                                                               \
  do {                                                                              \
    cudaError_t _e = cudaGetLastError();                                            \
    if (_e != cudaSuccess) {                                                        \
      mc::set_error("kernel launch failed: %s (%s:%d)", cudaGetErrorString(_e), __FILE__, __LINE__); \
      return 1;                                                                     \
    }                                                                               \
    mc::count_launch();                                                             \
  } while (0)

// ---- programmatic dependent launch (PDL) ------------------------------------------------
// A training step is 30-450 small dependent kernels; between two of them the GPU drains, the next
// grid is launched, its CTAs are scheduled and run their prologue (barrier init, TMEM allocation,
// descriptor prefetch) -- 1-3 us during which the SMs idle.  Every kernel of this library is launched
// with programmaticStreamSerializationAllowed and starts with pdl_sync(): `launch_dependents` lets the
// NEXT grid's CTAs become resident as soon as all of this grid's CTAs have started, and `wait` holds
// them until the previous grid has completed and flushed its memory, so the launch ramp and prologue
// of kernel i+1 overlap the tail of kernel i.  Correctness rule: no global memory access before
// pdl_sync() (kernel parameters live in the constant bank and are fine).  Works in eager streams and
// inside captured graphs (programmatic dependency edges).  MC_PDL=0 turns the attribute off.
// MEASURED (profiles/r2d): with the attribute on EVERY kernel the mlp_c3 step got 3 % SLOWER (0.287 -> 0.296 ms)
// and lstm_c5 2 % slower, while a chain of GEMMs alone got 8 % faster (7.6 -> 7.0 us per 1024x512x512 call).
// A dependent grid is placed when it is launched early: the blocks of a small multi-CTA-per-SM kernel
// (batchnorm, dropout, reductions: 64-600 blocks) pile onto the few SMs the running GEMM leaves idle instead of
// spreading over all 148 once it finishes.  The tcgen05 kernels hold one CTA per SM, so their early CTAs can
// only take SMs that are really free.  Default therefore: the attribute only on launches that ask for it
// (`launch_pdl`: GEMM and convolution kernels).  MC_PDL=0: never, MC_PDL=2: every kernel (the experiment).
int pdl_mode();
inline bool pdl_enabled() { return pdl_mode() == 2; }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_sync() { pdl_trigger(); pdl_wait(); }

template <class... KArgs, class... Args>
inline cudaError_t launch(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream();
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// The same with the programmatic-launch attribute unless MC_PDL=0: for kernels that keep one CTA per SM.
template <class... KArgs, class... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream();
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_mode() != 0 ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// ---- kernel-internal scratch memory -------------------------------------------------------
// One persistent arena per device, handed out LIFO: `ScratchFrame f; float* t = f.floats(n);` and the
// space returns when `f` goes out of scope.  Every kernel of the library runs on one stream (and a PDL
// kernel touches global memory only after the previous grid completed), so a later call may reuse the
// bytes an earlier call's kernels used: no cudaMallocAsync / cudaFreeAsync pair per call, which inside a
// captured step graph were two memory nodes each.  If the arena is too small a larger one is allocated;
// old arenas stay alive (captured graphs and pointers handed out earlier in the same call refer to them).
void* scratch_push(size_t bytes);          // nullptr (and the error string set) when out of memory
size_t scratch_mark();
void scratch_release(size_t mark);
struct ScratchFrame {
  size_t mark;
  ScratchFrame() : mark(scratch_mark()) {}
  ~ScratchFrame() { scratch_release(mark); }
  ScratchFrame(const ScratchFrame&) = delete;
  ScratchFrame& operator=(const ScratchFrame&) = delete;
  float* floats(size_t n) { return static_cast<float*>(scratch_push(sizeof(float) * (n ? n : 1))); }
};

// Grid sizing for bandwidth-bound grid-stride kernels: enough CTAs to fill every SM
// `waves` times over, never more than the work needs.
inline int grid_for(int64_t work_items, int per_block, int ctas_per_sm = 8) {
  int64_t need = (work_items + per_block - 1) / per_block;
  int64_t cap = (int64_t)sm_count() * ctas_per_sm;
  if (need < 1) need = 1;
  return (int)(need < cap ? need : cap);
}

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }
__device__ __forceinline__ bool aligned16_dev(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// ---- device helpers ------------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_min(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// NaN-propagating max/min (NumPy's np.max / np.maximum propagate NaN; fmaxf does not).
__device__ __forceinline__ float nanmax(float a, float b) {
  return (a != a) ? a : ((b != b) ? b : fmaxf(a, b));
}
__device__ __forceinline__ float nanmin(float a, float b) {
  return (a != a) ? a : ((b != b) ? b : fminf(a, b));
}

// GEMM epilogue relu.  mode 1: max(0, v), NaN propagates.  mode 2 ("signed zero"): the same VALUES, but a
// negative pre-activation is stored as -0.0f and a zero one as +0.0f, so the sign bit of the output still
// tells x < 0 from x >= 0: layers.relu passes the gradient where x >= 0, INCLUDING x == 0
// (numpy_core.py:284-286), and a fused affine+relu never materialises x.  -0.0f == 0.0f in every
// comparison and sum downstream.
__device__ __forceinline__ float relu_apply(float v, int mode) {
  if (v != v) return v;
  if (v < 0.f) return mode == 2 ? -0.0f : 0.f;
  return v == 0.f ? 0.f : v;
}

// Block-wide reduction through shared memory; `Op` is a functor with ::identity() and
// operator()(a,b).  Result valid in every thread.  Requires blockDim.x % 32 == 0, <= 1024.
template <class Op>
__device__ __forceinline__ float block_reduce(float v, Op op, float* smem32) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = op(v, __shfl_xor_sync(0xffffffffu, v, o));
  if (lane == 0) smem32[warp] = v;
  __syncthreads();
  const int nwarps = (blockDim.x + 31) >> 5;
  float r = (lane < nwarps) ? smem32[lane] : Op::identity();
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) r = op(r, __shfl_xor_sync(0xffffffffu, r, o));
  __syncthreads();
  return r;
}

struct SumOp {
  __device__ static float identity() { return 0.f; }
  __device__ float operator()(float a, float b) const { return a + b; }
};
struct MaxOp {
  __device__ static float identity() { return -INFINITY; }
  __device__ float operator()(float a, float b) const { return nanmax(a, b); }
};
struct MinOp {
  __device__ static float identity() { return INFINITY; }
  __device__ float operator()(float a, float b) const { return nanmin(a, b); }
};

}  // namespace mc

// In-kernel phase timing (tools/probe.sh builds a second library with -DMC_PROBE; nothing of this is compiled into
// libminpy_cuda.so).  MC_PROBE_DEFINE(tag) at file scope declares the stamp buffer of a kernel and the entry point
// mc_probe_read_<tag>(out[1024 * 16]); MC_STAMP(tag, i), i < 16, makes thread 0 of every block (the first 1024) store
// %globaltimer (ns).  tools/probe.py prints per-phase medians over blocks.
#ifdef MC_PROBE
#define MC_PROBE_DEFINE(tag)                                                                           \
  __device__ unsigned long long g_probe_##tag[1024 * 16];                                               \
  extern "C" int mc_probe_read_##tag(unsigned long long* out) {                                        \
    return cudaMemcpyFromSymbol(out, g_probe_##tag, sizeof(g_probe_##tag)) == cudaSuccess ? 0 : 1;     \
  }
#define MC_STAMP(tag, i)                                                                               \
  do {                                                                                                 \
    const unsigned lin_ = blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z);              \
    if (threadIdx.x == 0 && lin_ < 1024) {                                                             \
      unsigned long long t_;                                                                           \
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));                                           \
      g_probe_##tag[lin_ * 16 + (i)] = t_;                                                              \
    }                                                                                                  \
  } while (0)
#else
#define MC_PROBE_DEFINE(tag)
#define MC_STAMP(tag, i)
#endif
