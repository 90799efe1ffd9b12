// The one exchange step of data-parallel training: an in-place average of the
// flat gradient buffer over all ranks, NCCL over NVLink / NVSwitch, enqueued on the LIBRARY's
// stream.  Being on that stream is the point: the all-reduce is then part of the captured step
// graph (NCCL collectives are graph-capturable), so a data-parallel step is still ONE
// cudaGraphLaunch with no host synchronisation between backward, exchange and update.
// The reference has no counterpart (no multi-device training loop, minpy/context.py:92-98).
//
// NCCL is bound at run time (dlopen "libnccl.so.2"): in a process that already imported torch
// this resolves to the copy torch loaded, so there is exactly one NCCL in the address space; the
// library itself has no link-time dependency and single-GPU use never touches NCCL.
#include <dlfcn.h>
#include "common.cuh"

namespace mc {
namespace {

struct NcclUniqueId { char internal[128]; };
typedef void* NcclComm;
typedef int (*GetUniqueIdFn)(NcclUniqueId*);
typedef int (*CommInitRankFn)(NcclComm*, int, NcclUniqueId, int);
typedef int (*AllReduceFn)(const void*, void*, size_t, int, int, NcclComm, cudaStream_t);
typedef int (*BroadcastFn)(const void*, void*, size_t, int, int, NcclComm, cudaStream_t);
typedef int (*CommDestroyFn)(NcclComm);
typedef const char* (*GetErrorStringFn)(int);
typedef int (*GetVersionFn)(int*);

constexpr int kNcclFloat32 = 7;   // ncclDataType_t
constexpr int kNcclSum = 0, kNcclAvg = 4;   // ncclRedOp_t

struct Api {
  void* handle = nullptr;
  GetUniqueIdFn get_unique_id = nullptr;
  CommInitRankFn comm_init_rank = nullptr;
  AllReduceFn all_reduce = nullptr;
  BroadcastFn broadcast = nullptr;
  CommDestroyFn comm_destroy = nullptr;
  GetErrorStringFn error_string = nullptr;
  GetVersionFn get_version = nullptr;
} g_api;

NcclComm g_comm = nullptr;
int g_world = 1, g_rank = 0;

int load_api() {
  if (g_api.handle) return 0;
  void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
  if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
  if (!h) { set_error("NCCL not found: %s", dlerror()); return 1; }
  g_api.get_unique_id = (GetUniqueIdFn)dlsym(h, "ncclGetUniqueId");
  g_api.comm_init_rank = (CommInitRankFn)dlsym(h, "ncclCommInitRank");
  g_api.all_reduce = (AllReduceFn)dlsym(h, "ncclAllReduce");
  g_api.broadcast = (BroadcastFn)dlsym(h, "ncclBroadcast");
  g_api.comm_destroy = (CommDestroyFn)dlsym(h, "ncclCommDestroy");
  g_api.error_string = (GetErrorStringFn)dlsym(h, "ncclGetErrorString");
  g_api.get_version = (GetVersionFn)dlsym(h, "ncclGetVersion");
  if (!g_api.get_unique_id || !g_api.comm_init_rank || !g_api.all_reduce || !g_api.broadcast || !g_api.comm_destroy) {
    set_error("libnccl.so.2 lacks a required symbol");
    return 1;
  }
  g_api.handle = h;
  return 0;
}

int check(int rc, const char* what) {
  if (rc == 0) return 0;
  set_error("%s failed: %s", what, g_api.error_string ? g_api.error_string(rc) : "NCCL error");
  return 1;
}


// ---------------------------------------------------------------------------------------------
// Two-shot all-reduce over NVLink peer memory (one node, W <= 8), used instead of NCCL when the
// gradient bucket lives in the library's IPC window (mc_p2p_create / mc_p2p_connect).  The bucket is
// 0.5-20 MB: NCCL's ring moves it at 150-300 GB/s algbw because the call is latency-bound (48 us for
// 9.5 MB at W = 2, profiles/r1z/n2/comm_default.log); here every GPU pulls what it needs straight
// from its peers' buffers:
//   barrier                      all ranks have packed their gradients
//   reduce-scatter   rank r sums slice r of every rank's buffer (fixed rank order: the result is
//                    bit-identical on all ranks and from run to run), scales by 1/W, stores it in its own slice
//   barrier
//   all-gather       rank r copies slice q from rank q, for every q != r
//   barrier                      nobody reads my buffer any more: the next step may overwrite it
// The barriers are PER BLOCK: block b only ever reads what block b of a peer wrote (same index
// pattern in both phases), so block b of rank r signals / waits for block b of the peers -- no
// grid-wide barrier, no co-residency requirement beyond "the grid fits the GPU".  Flags carry a
// per-block counter kept in device memory, so the kernel replays unchanged inside a CUDA graph.
constexpr int P2P_MAX_WORLD = 8;
constexpr int P2P_MAX_BLOCKS = 128;
constexpr int P2P_THREADS = 512;
constexpr unsigned long long P2P_TIMEOUT_NS = 20ull * 1000ull * 1000ull * 1000ull;

struct P2PParams {
  float* bufs[P2P_MAX_WORLD];        // every rank's window (own entry = local pointer)
  uint32_t* flags[P2P_MAX_WORLD];    // every rank's flag array [P2P_MAX_BLOCKS][P2P_MAX_WORLD]
  uint32_t* counter;                 // local: per-block barrier count
  int* err;                          // mapped host memory: 1 after a barrier timed out
  int rank, world;
};

__device__ __forceinline__ unsigned long long p2p_now() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

__device__ __forceinline__ void p2p_block_barrier(const P2PParams& p) {
  // Every thread reads the count before the CTA barrier, thread 0 bumps it after: no race, no smem.
  const uint32_t val = p.counter[blockIdx.x] + 1u;
  __syncthreads();                   // the block's stores of the phase before are ordered before the signal:
  if (threadIdx.x == 0) p.counter[blockIdx.x] = val;
  if ((int)threadIdx.x < p.world && (int)threadIdx.x != p.rank) {
    const int q = threadIdx.x;
    // ONE system-scope fence per signalling thread (cumulative over what the CTA barrier ordered) --
    // a fence in all 512 threads of all blocks cost ~10 us per barrier (profiles/r1z/n2/comm_p2p_v1.log)
    __threadfence_system();
    volatile uint32_t* theirs = p.flags[q] + blockIdx.x * P2P_MAX_WORLD + p.rank;
    *theirs = val;
    const volatile uint32_t* mine = p.flags[p.rank] + blockIdx.x * P2P_MAX_WORLD + q;
    unsigned long long t0 = 0;
    for (uint32_t spins = 1;; ++spins) {
      if ((int32_t)(*mine - val) >= 0) break;
      if ((spins & 4095u) == 0) {     // watchdog off the fast path (the error word lives in host memory)
        if (t0 == 0) t0 = p2p_now();
        if (*(volatile int*)p.err) break;
        if (p2p_now() - t0 > P2P_TIMEOUT_NS) { *(volatile int*)p.err = 1; break; }   // a peer never arrived
      }
    }
    __threadfence_system();          // acquire: the peers' stores before their signal are visible after this
  }
  __syncthreads();
}

__device__ __forceinline__ float4 p2p_ld(const float* base, int64_t i4) {
  return __ldcv(reinterpret_cast<const float4*>(base) + i4);     // never from a stale cache line
}

__global__ void __launch_bounds__(P2P_THREADS) p2p_allreduce_kernel(P2PParams p, int64_t off4, int64_t n4, float scale) {
  const int W = p.world;
  const int64_t chunk = (n4 + W - 1) / W;
  const int64_t lane = (int64_t)blockIdx.x * P2P_THREADS + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * P2P_THREADS;
  float* mine = p.bufs[p.rank];
  p2p_block_barrier(p);
  constexpr int U = 4;               // float4 loads in flight per thread and peer (NVLink latency ~2-3 us)
  {
    const int64_t lo = (int64_t)p.rank * chunk, hi = min(n4, lo + chunk);
    int64_t i = lo + lane;
    for (; i + (U - 1) * stride < hi; i += U * stride) {
      float4 acc[U];
#pragma unroll
      for (int u = 0; u < U; ++u) acc[u] = p2p_ld(p.bufs[0], off4 + i + u * stride);
      for (int q = 1; q < W; ++q) {
        float4 v[U];
#pragma unroll
        for (int u = 0; u < U; ++u) v[u] = p2p_ld(p.bufs[q], off4 + i + u * stride);
#pragma unroll
        for (int u = 0; u < U; ++u) { acc[u].x += v[u].x; acc[u].y += v[u].y; acc[u].z += v[u].z; acc[u].w += v[u].w; }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        acc[u].x *= scale; acc[u].y *= scale; acc[u].z *= scale; acc[u].w *= scale;
        reinterpret_cast<float4*>(mine)[off4 + i + u * stride] = acc[u];
      }
    }
    for (; i < hi; i += stride) {
      float4 acc = p2p_ld(p.bufs[0], off4 + i);
      for (int q = 1; q < W; ++q) {
        const float4 v = p2p_ld(p.bufs[q], off4 + i);
        acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
      }
      acc.x *= scale; acc.y *= scale; acc.z *= scale; acc.w *= scale;
      reinterpret_cast<float4*>(mine)[off4 + i] = acc;
    }
  }
  p2p_block_barrier(p);
  for (int d = 1; d < W; ++d) {
    const int q = (p.rank + d) % W;          // start at a different peer on every rank
    const int64_t lo = (int64_t)q * chunk, hi = min(n4, lo + chunk);
    int64_t i = lo + lane;
    for (; i + (U - 1) * stride < hi; i += U * stride) {
      float4 v[U];
#pragma unroll
      for (int u = 0; u < U; ++u) v[u] = p2p_ld(p.bufs[q], off4 + i + u * stride);
#pragma unroll
      for (int u = 0; u < U; ++u) reinterpret_cast<float4*>(mine)[off4 + i + u * stride] = v[u];
    }
    for (; i < hi; i += stride) reinterpret_cast<float4*>(mine)[off4 + i] = p2p_ld(p.bufs[q], off4 + i);
  }
  p2p_block_barrier(p);
}

// PUSH version (default; MINPY_P2P_PUSH=0 selects the pull kernel above).  NVLink stores are posted -- they cost
// bandwidth, not a round trip -- while every peer LOAD of the pull version is a 2-3 us round trip, and the pull
// version needs three barriers (data ready / slices reduced / everybody done reading).  Here:
//   reduce-scatter   rank r WRITES slice q of its gradients into rank q's receive area (slot r)
//   barrier          (flag after a system fence: the pushed data is visible before the flag)
//   reduce + gather  rank r sums its slice over the W slots in rank order (own slot = its own buffer), scales,
//                    and WRITES the result into its own window and into every peer's window
//   barrier
// Two barriers, no peer loads.  No trailing barrier is needed: a peer can only start pushing step k+1 after it has
// seen this rank's second flag of step k, which is sent after this rank finished reading its receive area.
// The receive area is the second half of the IPC allocation (mc_p2p_create allocates 2 x bytes).
__global__ void __launch_bounds__(P2P_THREADS) p2p_allreduce_push_kernel(P2PParams p, int64_t off4, int64_t n4, int64_t recv4,
                                                                         float scale) {
  const int W = p.world, R = p.rank;
  const int64_t chunk = (n4 + W - 1) / W;
  const int64_t lane = (int64_t)blockIdx.x * P2P_THREADS + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * P2P_THREADS;
  float4* mine = reinterpret_cast<float4*>(p.bufs[R]) + off4;
  constexpr int U = 4;
  for (int d = 1; d < W; ++d) {
    const int q = (R + d) % W;                 // start at a different peer on every rank
    const int64_t lo = (int64_t)q * chunk, hi = min(n4, lo + chunk);
    float4* dst = reinterpret_cast<float4*>(p.bufs[q]) + recv4 + off4 + (int64_t)R * chunk - lo;
    int64_t i = lo + lane;
    for (; i + (U - 1) * stride < hi; i += U * stride) {
      float4 v[U];
#pragma unroll
      for (int u = 0; u < U; ++u) v[u] = mine[i + u * stride];
#pragma unroll
      for (int u = 0; u < U; ++u) dst[i + u * stride] = v[u];
    }
    for (; i < hi; i += stride) dst[i] = mine[i];
  }
  p2p_block_barrier(p);
  {
    const int64_t lo = (int64_t)R * chunk, hi = min(n4, lo + chunk);
    const float4* recv = reinterpret_cast<const float4*>(p.bufs[R]) + recv4 + off4 - lo;
    for (int64_t i = lo + lane; i < hi; i += stride) {
      float4 acc = (R == 0) ? mine[i] : __ldcv(recv + i);
      for (int s = 1; s < W; ++s) {
        const float4 v = (s == R) ? mine[i] : __ldcv(recv + (int64_t)s * chunk + i);
        acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
      }
      acc.x *= scale; acc.y *= scale; acc.z *= scale; acc.w *= scale;
      mine[i] = acc;
      for (int d = 1; d < W; ++d) {
        const int q = (R + d) % W;
        (reinterpret_cast<float4*>(p.bufs[q]) + off4)[i] = acc;
      }
    }
  }
  p2p_block_barrier(p);
}

struct P2PState {
  bool active = false;
  int rank = 0, world = 1;
  size_t bytes = 0;
  float* base = nullptr;
  uint32_t* flags = nullptr;
  void* peer_data[P2P_MAX_WORLD] = {};
  void* peer_flags[P2P_MAX_WORLD] = {};
  size_t recv_offset = 0;            // bytes from the window base to the receive area of the push kernel
  P2PParams params;
  int* err_host = nullptr;
} g_p2p;

int p2p_allreduce(float* buf, int64_t n, float scale) {
  if (*(volatile int*)g_p2p.err_host) { set_error("peer-memory all-reduce: a barrier timed out earlier (a rank never arrived)"); return 1; }
  const int64_t off4 = (buf - g_p2p.base) / 4, n4 = n / 4;
  const int64_t per_rank = (n4 + g_p2p.world - 1) / g_p2p.world;
  int64_t blocks = (per_rank + P2P_THREADS - 1) / P2P_THREADS;
  if (blocks < 1) blocks = 1;
  const int64_t cap = sm_count() < P2P_MAX_BLOCKS ? sm_count() : P2P_MAX_BLOCKS;
  if (blocks > cap) blocks = cap;
  static int push = -1;
  if (push < 0) { const char* e = getenv("MINPY_P2P_PUSH"); push = (e && e[0] == '0') ? 0 : 1; }
  if (push) p2p_allreduce_push_kernel<<<(unsigned)blocks, P2P_THREADS, 0, stream()>>>(g_p2p.params, off4, n4,
                                                                                  (int64_t)(g_p2p.recv_offset / 16), scale);
  else p2p_allreduce_kernel<<<(unsigned)blocks, P2P_THREADS, 0, stream()>>>(g_p2p.params, off4, n4, scale);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_error("p2p_allreduce_kernel launch failed: %s", cudaGetErrorString(e)); return 1; }
  count_launch();
  return 0;
}

bool p2p_covers(const float* buf, int64_t n) {
  if (!g_p2p.active || (n & 3) || !aligned16(buf)) return false;
  const char* b = (const char*)buf;
  return b >= (const char*)g_p2p.base && b + 4 * n <= (const char*)g_p2p.base + g_p2p.bytes;
}

}  // namespace
}  // namespace mc

using namespace mc;

extern "C" {

int mc_comm_unique_id(void* id128) {
  if (load_api()) return 1;
  return check(g_api.get_unique_id((NcclUniqueId*)id128), "ncclGetUniqueId");
}

int mc_comm_init(int rank, int world, const void* id128) {
  MC_REQUIRE(world >= 1 && rank >= 0 && rank < world, "bad rank / world size");
  if (load_api()) return 1;
  if (g_comm) { g_api.comm_destroy(g_comm); g_comm = nullptr; }
  NcclUniqueId id;
  memcpy(&id, id128, sizeof(id));
  if (check(g_api.comm_init_rank(&g_comm, world, id, rank), "ncclCommInitRank")) return 1;
  g_world = world;
  g_rank = rank;
  return 0;
}

int mc_comm_info(int* rank, int* world, int* nccl_version) {
  *rank = g_rank;
  *world = g_comm ? g_world : 1;
  *nccl_version = 0;
  if (g_api.get_version) g_api.get_version(nccl_version);
  return 0;
}

int mc_allreduce_avg(float* buf, int64_t n) {
  if (n <= 0) return 0;
  if (p2p_covers(buf, n)) return p2p_allreduce(buf, n, 1.0f / (float)g_p2p.world);
  MC_REQUIRE(g_comm != nullptr, "mc_comm_init has not been called");
  if (check(g_api.all_reduce(buf, buf, (size_t)n, kNcclFloat32, kNcclAvg, g_comm, stream()), "ncclAllReduce")) return 1;
  count_launch();
  return 0;
}

int mc_allreduce_sum(float* buf, int64_t n) {
  if (n <= 0) return 0;
  MC_REQUIRE(g_comm != nullptr, "mc_comm_init has not been called");
  if (check(g_api.all_reduce(buf, buf, (size_t)n, kNcclFloat32, kNcclSum, g_comm, stream()), "ncclAllReduce")) return 1;
  count_launch();
  return 0;
}

int mc_broadcast(float* buf, int64_t n, int root) {
  if (n <= 0) return 0;
  MC_REQUIRE(g_comm != nullptr, "mc_comm_init has not been called");
  if (check(g_api.broadcast(buf, buf, (size_t)n, kNcclFloat32, root, g_comm, stream()), "ncclBroadcast")) return 1;
  count_launch();
  return 0;
}

/* Peer-memory window for the gradient bucket (see p2p_allreduce_kernel). */
int mc_p2p_create(int rank, int world, size_t bytes, void* handles128) {
  MC_REQUIRE(world >= 2 && world <= P2P_MAX_WORLD && rank >= 0 && rank < world, "peer-memory all-reduce needs 2..8 ranks");
  MC_REQUIRE(!g_p2p.base, "mc_p2p_create called twice");
  MC_REQUIRE(sizeof(cudaIpcMemHandle_t) == 64, "unexpected IPC handle size");
  bytes = (bytes + 4095) & ~(size_t)4095;
  // [window: bytes][receive area of the push kernel: bytes + slack for the rounded-up slices]
  g_p2p.recv_offset = bytes;
  MC_CHECK(cudaMalloc((void**)&g_p2p.base, 2 * bytes + 4096));
  MC_CHECK(cudaMemset(g_p2p.base, 0, 2 * bytes + 4096));
  const size_t flag_bytes = sizeof(uint32_t) * P2P_MAX_BLOCKS * (P2P_MAX_WORLD + 1);   // flags + local counters
  MC_CHECK(cudaMalloc((void**)&g_p2p.flags, flag_bytes));
  MC_CHECK(cudaMemset(g_p2p.flags, 0, flag_bytes));
  MC_CHECK(cudaHostAlloc((void**)&g_p2p.err_host, sizeof(int), cudaHostAllocMapped));
  *g_p2p.err_host = 0;
  MC_CHECK(cudaDeviceSynchronize());
  cudaIpcMemHandle_t h[2];
  MC_CHECK(cudaIpcGetMemHandle(&h[0], g_p2p.base));
  MC_CHECK(cudaIpcGetMemHandle(&h[1], g_p2p.flags));
  memcpy(handles128, h, 128);
  g_p2p.rank = rank; g_p2p.world = world; g_p2p.bytes = bytes;
  return 0;
}
/* all_handles: world x 128 bytes, rank-major, as gathered from every rank's mc_p2p_create. */
int mc_p2p_connect(const void* all_handles) {
  MC_REQUIRE(g_p2p.base != nullptr, "mc_p2p_create first");
  P2PParams& p = g_p2p.params;
  for (int q = 0; q < g_p2p.world; ++q) {
    if (q == g_p2p.rank) { p.bufs[q] = g_p2p.base; p.flags[q] = g_p2p.flags; continue; }
    cudaIpcMemHandle_t h[2];
    memcpy(h, (const char*)all_handles + (size_t)q * 128, 128);
    MC_CHECK(cudaIpcOpenMemHandle(&g_p2p.peer_data[q], h[0], cudaIpcMemLazyEnablePeerAccess));
    MC_CHECK(cudaIpcOpenMemHandle(&g_p2p.peer_flags[q], h[1], cudaIpcMemLazyEnablePeerAccess));
    p.bufs[q] = (float*)g_p2p.peer_data[q];
    p.flags[q] = (uint32_t*)g_p2p.peer_flags[q];
  }
  p.counter = g_p2p.flags + P2P_MAX_BLOCKS * P2P_MAX_WORLD;
  int* err_dev = nullptr;
  MC_CHECK(cudaHostGetDevicePointer((void**)&err_dev, g_p2p.err_host, 0));
  p.err = err_dev;
  p.rank = g_p2p.rank; p.world = g_p2p.world;
  g_p2p.active = true;
  return 0;
}
/* The local window (ptr NULL while inactive) and whether a barrier ever timed out. */
int mc_p2p_info(void** ptr, size_t* bytes, int* timed_out) {
  *ptr = g_p2p.active ? (void*)g_p2p.base : nullptr;
  *bytes = g_p2p.active ? g_p2p.bytes : 0;
  *timed_out = g_p2p.err_host ? *(volatile int*)g_p2p.err_host : 0;
  return 0;
}

int mc_comm_destroy(void) {
  if (g_comm && g_api.comm_destroy) g_api.comm_destroy(g_comm);
  g_comm = nullptr;
  g_world = 1;
  g_rank = 0;
  return 0;
}

}  // extern "C"
