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

int mc_comm_destroy(void) {
  if (g_comm && g_api.comm_destroy) g_api.comm_destroy(g_comm);
  g_comm = nullptr;
  g_world = 1;
  g_rank = 0;
  return 0;
}

}  // extern "C"
