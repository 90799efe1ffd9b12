// Runtime plumbing of libminpy_cuda.so: device / stream selection, the stream-ordered
// caching allocator, copies, events, CUDA-graph capture and the launch counter.
#include <atomic>
#include <stdarg.h>
#include <stdlib.h>
#include "common.cuh"

namespace mc {

static thread_local char g_err[512] = "";
#define MC_MAX_DEVICES 16
static cudaStream_t g_stream = nullptr;                  // the CURRENT device's compute stream
static cudaStream_t g_streams[MC_MAX_DEVICES] = {};      // one blocking compute stream per device ever selected
static int g_device = -1;
static int g_sm_count = 148;
static std::atomic<uint64_t> g_launches{0};
static unsigned long long* g_epoch = nullptr;     // counter of the graph being captured
static bool g_capturing = false;
static cudaStream_t g_copy_stream = nullptr;     // input prefetch (host -> device) beside the compute stream
static cudaEvent_t g_copy_done = nullptr, g_main_done = nullptr;

__global__ void graph_epoch_inc_kernel(unsigned long long* e) { *e += 1ull; }

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
cudaStream_t stream() { return g_stream; }
int sm_count() { return g_sm_count; }
int device() { return g_device; }
int pdl_mode() {
  static int mode = -1;
  if (mode < 0) { const char* e = getenv("MC_PDL"); mode = e ? atoi(e) : 1; if (mode < 0 || mode > 2) mode = 1; }
  return mode;
}
bool first_use_on_device(uint32_t* mask) {
  const uint32_t bit = 1u << (g_device < 0 ? 0 : g_device);
  if (*mask & bit) return false;
  *mask |= bit;
  return true;
}
void count_launch(int n) { g_launches.fetch_add((uint64_t)n, std::memory_order_relaxed); }
const unsigned long long* graph_epoch() { return g_capturing ? g_epoch : nullptr; }

// ---- side stream (weight-gradient contractions beside the main backward chain) ------------------------
static cudaStream_t g_side_streams[MC_MAX_DEVICES] = {};
static cudaEvent_t g_fork_ev[MC_MAX_DEVICES] = {}, g_join_ev[MC_MAX_DEVICES] = {};
static bool g_in_side = false, g_side_pending = false;
int side_index() { return g_in_side ? 1 : 0; }

struct Arena { char* base; size_t cap, top; };
static Arena g_arena[MC_MAX_DEVICES][2] = {};       // [device][main / side stream]: the two run concurrently
size_t scratch_mark() { return g_arena[g_device < 0 ? 0 : g_device][side_index()].top; }
void scratch_release(size_t mark) { g_arena[g_device < 0 ? 0 : g_device][side_index()].top = mark; }
void* scratch_push(size_t bytes) {
  Arena& a = g_arena[g_device < 0 ? 0 : g_device][side_index()];
  bytes = (bytes + 255) & ~(size_t)255;
  if (a.top + bytes > a.cap) {
    size_t want = a.cap ? a.cap * 2 : ((size_t)64 << 20);
    while (want < a.top + bytes) want *= 2;
    char* fresh = nullptr;
    if (cudaMalloc((void**)&fresh, want) != cudaSuccess) {
      cudaGetLastError();
      set_error("scratch arena: cudaMalloc of %zu bytes failed", want);
      return nullptr;
    }
    a.base = fresh;              // the previous arena is kept alive on purpose (see common.cuh)
    a.cap = want;
  }
  void* p = a.base + a.top;
  a.top += bytes;
  return p;
}

static int configure_device(int device) {
  MC_CHECK(cudaSetDevice(device));
  cudaDeviceProp prop;
  MC_CHECK(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10) {
    set_error("libminpy_cuda is built for sm_100a only; device %d is sm_%d%d", device, prop.major, prop.minor);
    return 3;
  }
  if (device < 0 || device >= MC_MAX_DEVICES) {
    set_error("device %d out of range (0..%d)", device, MC_MAX_DEVICES - 1);
    return 2;
  }
  if (g_capturing && device != g_device) {
    set_error("cannot change the device while a step graph is being captured");
    return 2;
  }
  g_sm_count = prop.multiProcessorCount;
  g_device = device;
  // The library's own stream.  A BLOCKING stream: it keeps the implicit ordering with the legacy
  // default stream (where torch.distributed enqueues the NCCL all-reduce), and unlike the legacy
  // stream it can be captured into a CUDA graph (mc_graph_begin).
  // One such stream PER DEVICE (minpy/context.py:92-98 `with gpu(i):` places work on device i).
  if (g_streams[device] == nullptr) MC_CHECK(cudaStreamCreateWithFlags(&g_streams[device], cudaStreamDefault));
  g_stream = g_streams[device];
  // Keep freed blocks cached in the default mempool instead of returning them to the
  // driver at every synchronisation: training steps reuse the same sizes every step.
  cudaMemPool_t pool;
  MC_CHECK(cudaDeviceGetDefaultMemPool(&pool, device));
  uint64_t threshold = UINT64_MAX;
  MC_CHECK(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &threshold));
  return 0;
}

}  // namespace mc

using namespace mc;

extern "C" {

const char* mc_last_error(void) { return g_err; }

int mc_device_count(int* count) {
  MC_CHECK(cudaGetDeviceCount(count));
  return 0;
}
int mc_init(int device) { return configure_device(device); }
int mc_set_device(int device) { return configure_device(device); }
int mc_get_device(int* device) {
  *device = g_device;
  return 0;
}
int mc_sm_count(int* count) {
  *count = g_sm_count;
  return 0;
}
int mc_set_stream(void* s) {
  g_stream = (cudaStream_t)s;
  return 0;
}
int mc_get_stream(void** s) {
  *s = (void*)g_stream;
  return 0;
}
/* Side stream: work that nothing on the main chain waits for until the end of the backward pass -- the
 * weight-gradient GEMMs / convolutions (dW = x^T . g is read only by the update rule, while dX = g . W^T feeds
 * the next VJP) -- runs on a second stream, forked from and joined back into the compute stream with events,
 * inside a captured step graph too (parallel branches).  Between mc_side_begin and mc_side_end every library
 * call is enqueued on the side stream; mc_side_join makes the compute stream wait for all of it.  The caller
 * keeps the operands of side work alive until the join (the host-side free lists assume one stream). */
int mc_side_begin(void) {
  MC_REQUIRE(!g_in_side, "mc_side_begin: already on the side stream");
  MC_REQUIRE(g_device >= 0 && g_stream != nullptr, "mc_init first");
  const int d = g_device;
  if (g_side_streams[d] == nullptr) {
    MC_CHECK(cudaStreamCreateWithFlags(&g_side_streams[d], cudaStreamNonBlocking));
    MC_CHECK(cudaEventCreateWithFlags(&g_fork_ev[d], cudaEventDisableTiming));
    MC_CHECK(cudaEventCreateWithFlags(&g_join_ev[d], cudaEventDisableTiming));
  }
  MC_CHECK(cudaEventRecord(g_fork_ev[d], g_streams[d]));
  MC_CHECK(cudaStreamWaitEvent(g_side_streams[d], g_fork_ev[d], 0));
  g_stream = g_side_streams[d];
  g_in_side = true;
  return 0;
}
int mc_side_end(void) {
  MC_REQUIRE(g_in_side, "mc_side_end without mc_side_begin");
  const int d = g_device;
  MC_CHECK(cudaEventRecord(g_join_ev[d], g_side_streams[d]));
  g_stream = g_streams[d];
  g_in_side = false;
  g_side_pending = true;
  return 0;
}
int mc_side_join(void) {
  if (g_in_side) { if (mc_side_end()) return 1; }
  if (!g_side_pending) return 0;
  MC_CHECK(cudaStreamWaitEvent(g_streams[g_device], g_join_ev[g_device], 0));
  g_side_pending = false;
  return 0;
}
int mc_sync(void) {
  if (g_side_pending || g_in_side) { if (mc_side_join()) return 1; }
  MC_CHECK(cudaStreamSynchronize(g_stream));
  return 0;
}
int mc_launch_count(uint64_t* count) {
  *count = g_launches.load();
  return 0;
}

int mc_alloc(void** ptr, size_t nbytes) {
  if (nbytes == 0) nbytes = 16;
  // During capture a stream-ordered allocation would become a graph-owned node whose address
  // dies with the replay, while the caller keeps the pointer (host-side free list): take the
  // memory from the driver instead (legal in relaxed capture mode; cudaFreeAsync accepts it later).
  if (g_capturing) MC_CHECK(cudaMalloc(ptr, nbytes));
  else MC_CHECK(cudaMallocAsync(ptr, nbytes, g_stream));
  return 0;
}
int mc_free(void* ptr) {
  if (!ptr) return 0;
  MC_CHECK(cudaFreeAsync(ptr, g_stream));
  return 0;
}
int mc_host_alloc(void** ptr, size_t nbytes) {
  MC_CHECK(cudaMallocHost(ptr, nbytes ? nbytes : 16));
  return 0;
}
int mc_host_free(void* ptr) {
  MC_CHECK(cudaFreeHost(ptr));
  return 0;
}
int mc_mem_info(size_t* free_bytes, size_t* total_bytes) {
  MC_CHECK(cudaMemGetInfo(free_bytes, total_bytes));
  return 0;
}
int mc_memcpy_h2d(void* dst, const void* src, size_t n) {
  if (n == 0) return 0;
  // a captured memcpy node would re-read `src` (usually a temporary pageable NumPy buffer) on every
  // replay, long after it was freed: refuse, so that the step graph falls back to eager execution
  MC_REQUIRE(!g_capturing, "mc_memcpy_h2d inside a graph capture (a host array was created inside the step)");
  MC_CHECK(cudaMemcpyAsync(dst, src, n, cudaMemcpyHostToDevice, g_stream));
  return 0;
}
int mc_memcpy_d2h(void* dst, const void* src, size_t n) {
  if (n) MC_CHECK(cudaMemcpyAsync(dst, src, n, cudaMemcpyDeviceToHost, g_stream));
  MC_CHECK(cudaStreamSynchronize(g_stream));
  return 0;
}
int mc_memcpy_d2h_async(void* dst, const void* src, size_t n) {
  if (n) MC_CHECK(cudaMemcpyAsync(dst, src, n, cudaMemcpyDeviceToHost, g_stream));
  return 0;
}
int mc_memcpy_d2d(void* dst, const void* src, size_t n) {
  if (n) MC_CHECK(cudaMemcpyAsync(dst, src, n, cudaMemcpyDeviceToDevice, g_stream));
  return 0;
}
/* Copy between two devices of this process (`with gpu(i):` operands that live on another device,
 * minpy/context.py:92-98).  The source device's stream is drained first -- the producer of `src` ran
 * there --, then the copy is enqueued on the CURRENT device's stream. */
int mc_memcpy_peer(void* dst, int dst_device, const void* src, int src_device, size_t n) {
  MC_REQUIRE(!g_capturing, "mc_memcpy_peer inside a graph capture");
  MC_REQUIRE(src_device >= 0 && src_device < MC_MAX_DEVICES && dst_device >= 0 && dst_device < MC_MAX_DEVICES,
             "device out of range");
  if (n == 0) return 0;
  if (g_streams[src_device]) MC_CHECK(cudaStreamSynchronize(g_streams[src_device]));
  MC_CHECK(cudaMemcpyPeerAsync(dst, dst_device, src, src_device, n, g_stream));
  return 0;
}
int mc_memset(void* dst, int byte, size_t n) {
  if (n) MC_CHECK(cudaMemsetAsync(dst, byte, n, g_stream));
  return 0;
}

/* Input pipeline (minpy/nn/io.py NDArrayIter prefetch): host->device copies on a second,
 * non-blocking stream so that batch i+1 travels while step i computes. */
/* Small tensors of a batch (the labels: 4 KB) travel on a stream of their own: behind the 12.6 MB of the images on one
 * stream their ~2.5 us of copy latency were added to every batch of a PCIe-bound step. */
static cudaStream_t g_copy_small = nullptr;
static cudaEvent_t g_copy_small_done = nullptr;
static bool g_copy_small_used = false;
#define MC_PREFETCH_SMALL (64u << 10)
static int ensure_copy_stream() {
  if (g_copy_stream) return 0;
  MC_CHECK(cudaStreamCreateWithFlags(&g_copy_stream, cudaStreamNonBlocking));
  MC_CHECK(cudaStreamCreateWithFlags(&g_copy_small, cudaStreamNonBlocking));
  MC_CHECK(cudaEventCreateWithFlags(&g_copy_done, cudaEventDisableTiming));
  MC_CHECK(cudaEventCreateWithFlags(&g_copy_small_done, cudaEventDisableTiming));
  MC_CHECK(cudaEventCreateWithFlags(&g_main_done, cudaEventDisableTiming));
  return 0;
}
int mc_prefetch_h2d(void* dst, const void* src, size_t n) {
  MC_REQUIRE(!g_capturing, "mc_prefetch_h2d inside a graph capture");
  if (ensure_copy_stream()) return 1;
  if (!n) return 0;
  static int small_stream = -1;
  if (small_stream < 0) { const char* e = getenv("MC_PREFETCH_SMALL_STREAM"); small_stream = (e && e[0] == '0') ? 0 : 1; }
  // MEASURED (same box, A/B): mlp_c1 (3.1 MB of images per batch) 0.0715 -> 0.0684 ms/step end to end with the labels on
  // their own stream; mlp_c3 (12.6 MB) 0.2345 -> 0.2358: a second DMA stream disturbs a long copy more than its latency
  // costs behind it.  Hence: only beside batches whose large tensor is below 8 MB.
  static size_t last_big = 0;
  if (small_stream && n <= MC_PREFETCH_SMALL && last_big < ((size_t)8 << 20)) {
    MC_CHECK(cudaMemcpyAsync(dst, src, n, cudaMemcpyHostToDevice, g_copy_small));
    g_copy_small_used = true;
  } else {
    if (n > MC_PREFETCH_SMALL) last_big = n;
    MC_CHECK(cudaMemcpyAsync(dst, src, n, cudaMemcpyHostToDevice, g_copy_stream));
  }
  return 0;
}
/* The compute stream waits for every prefetch copy enqueued so far. */
int mc_prefetch_acquire(void) {
  if (ensure_copy_stream()) return 1;
  MC_CHECK(cudaEventRecord(g_copy_done, g_copy_stream));
  MC_CHECK(cudaStreamWaitEvent(g_stream, g_copy_done, 0));
  if (g_copy_small_used) {
    MC_CHECK(cudaEventRecord(g_copy_small_done, g_copy_small));
    MC_CHECK(cudaStreamWaitEvent(g_stream, g_copy_small_done, 0));
  }
  return 0;
}
/* Later prefetch copies wait for everything enqueued on the compute stream so far (the
 * consumer of the staging buffer they are about to overwrite). */
int mc_prefetch_release(void) {
  if (ensure_copy_stream()) return 1;
  MC_CHECK(cudaEventRecord(g_main_done, g_stream));
  MC_CHECK(cudaStreamWaitEvent(g_copy_stream, g_main_done, 0));
  MC_CHECK(cudaStreamWaitEvent(g_copy_small, g_main_done, 0));
  return 0;
}

/* Deferred read of a small device value (the loss of a step, minpy/nn/solver.py:187 reads it with a
 * blocking asnumpy()): the device->host copy runs on a third stream behind an event of the compute
 * stream and lands in a pinned slot of a ring, so waiting for step i's loss does not wait for
 * step i+1 that was enqueued after it.  A ticket stays valid until MC_READBACK_SLOTS later begins. */
#define MC_READBACK_SLOTS 256
#define MC_READBACK_BYTES 64
static cudaStream_t g_rb_stream = nullptr;
static char* g_rb_host = nullptr;
static cudaEvent_t g_rb_ready[MC_READBACK_SLOTS], g_rb_done[MC_READBACK_SLOTS];
static uint64_t g_rb_next = 0;

int mc_readback_begin(const void* src_dev, size_t nbytes, uint64_t* ticket) {
  MC_REQUIRE(nbytes > 0 && nbytes <= MC_READBACK_BYTES, "mc_readback_begin: 1..64 bytes");
  MC_REQUIRE(!g_capturing, "mc_readback_begin inside a graph capture");
  if (g_rb_stream == nullptr) {
    MC_CHECK(cudaStreamCreateWithFlags(&g_rb_stream, cudaStreamNonBlocking));
    MC_CHECK(cudaMallocHost((void**)&g_rb_host, (size_t)MC_READBACK_SLOTS * MC_READBACK_BYTES));
    for (int i = 0; i < MC_READBACK_SLOTS; ++i) {
      MC_CHECK(cudaEventCreateWithFlags(&g_rb_ready[i], cudaEventDisableTiming));
      MC_CHECK(cudaEventCreateWithFlags(&g_rb_done[i], cudaEventDisableTiming));
    }
  }
  const uint64_t t = g_rb_next++;
  const int slot = (int)(t % MC_READBACK_SLOTS);
  MC_CHECK(cudaEventRecord(g_rb_ready[slot], g_stream));
  MC_CHECK(cudaStreamWaitEvent(g_rb_stream, g_rb_ready[slot], 0));
  MC_CHECK(cudaMemcpyAsync(g_rb_host + (size_t)slot * MC_READBACK_BYTES, src_dev, nbytes, cudaMemcpyDeviceToHost,
                           g_rb_stream));
  MC_CHECK(cudaEventRecord(g_rb_done[slot], g_rb_stream));
  *ticket = t;
  return 0;
}
/* Blocks until the copy of `ticket` has landed, then copies it out.  Status 4: the ticket's slot has
 * been reused (the caller falls back to a blocking read of the device value). */
int mc_readback_wait(uint64_t ticket, void* dst_host, size_t nbytes) {
  MC_REQUIRE(nbytes > 0 && nbytes <= MC_READBACK_BYTES, "mc_readback_wait: 1..64 bytes");
  if (ticket >= g_rb_next || g_rb_next - ticket > MC_READBACK_SLOTS) {
    set_error("mc_readback_wait: ticket %llu is no longer in the ring", (unsigned long long)ticket);
    return 4;
  }
  const int slot = (int)(ticket % MC_READBACK_SLOTS);
  MC_CHECK(cudaEventSynchronize(g_rb_done[slot]));
  if (g_rb_next - ticket > MC_READBACK_SLOTS) {      // overwritten while we waited
    set_error("mc_readback_wait: ticket %llu was overwritten", (unsigned long long)ticket);
    return 4;
  }
  memcpy(dst_host, g_rb_host + (size_t)slot * MC_READBACK_BYTES, nbytes);
  return 0;
}

int mc_event_create(void** ev) {
  cudaEvent_t e;
  MC_CHECK(cudaEventCreate(&e));
  *ev = (void*)e;
  return 0;
}
int mc_event_record(void* ev) {
  MC_CHECK(cudaEventRecord((cudaEvent_t)ev, g_stream));
  return 0;
}
int mc_event_sync(void* ev) {
  MC_CHECK(cudaEventSynchronize((cudaEvent_t)ev));
  return 0;
}
int mc_event_elapsed_ms(void* start, void* stop, float* ms) {
  MC_CHECK(cudaEventElapsedTime(ms, (cudaEvent_t)start, (cudaEvent_t)stop));
  return 0;
}
int mc_event_destroy(void* ev) {
  MC_CHECK(cudaEventDestroy((cudaEvent_t)ev));
  return 0;
}

int mc_graph_begin(void) {
  MC_REQUIRE(g_stream != nullptr, "graph capture needs the library stream (mc_init first)");
  MC_REQUIRE(!g_capturing, "a capture is already in progress");
  // one replay counter per graph; it lives as long as the process (8 bytes per captured graph)
  MC_CHECK(cudaMalloc((void**)&g_epoch, sizeof(unsigned long long)));
  MC_CHECK(cudaMemsetAsync(g_epoch, 0, sizeof(unsigned long long), g_stream));
  MC_CHECK(cudaStreamSynchronize(g_stream));
  MC_CHECK(cudaStreamBeginCapture(g_stream, cudaStreamCaptureModeRelaxed));
  g_capturing = true;
  graph_epoch_inc_kernel<<<1, 1, 0, g_stream>>>(g_epoch);
  MC_LAUNCHED();
  return 0;
}
int mc_graph_end(void** graph_exec) {
  MC_REQUIRE(g_capturing, "no capture in progress");
  if (g_side_pending || g_in_side) { if (mc_side_join()) return 1; }     // every forked branch rejoins the capture
  g_capturing = false;
  cudaGraph_t graph = nullptr;
  MC_CHECK(cudaStreamEndCapture(g_stream, &graph));
  MC_REQUIRE(graph != nullptr, "capture was invalidated");
  cudaGraphExec_t exec;
  MC_CHECK(cudaGraphInstantiate(&exec, graph, 0));
  MC_CHECK(cudaGraphDestroy(graph));
  *graph_exec = (void*)exec;
  return 0;
}
/* Give up a capture after an error inside it: ends the capture, discards the partial graph and
 * clears the sticky capture error so that eager work can continue. */
int mc_graph_abort(void) {
  if (g_in_side) { g_stream = g_streams[g_device]; g_in_side = false; }
  g_side_pending = false;
  if (!g_capturing) return 0;
  g_capturing = false;
  cudaGraph_t graph = nullptr;
  cudaStreamEndCapture(g_stream, &graph);
  if (graph) cudaGraphDestroy(graph);
  cudaGetLastError();
  return 0;
}
int mc_graph_capturing(int* flag) {
  *flag = g_capturing ? 1 : 0;
  return 0;
}
int mc_graph_launch(void* graph_exec) {
  MC_CHECK(cudaGraphLaunch((cudaGraphExec_t)graph_exec, g_stream));
  return 0;
}
int mc_graph_destroy(void* graph_exec) {
  MC_CHECK(cudaGraphExecDestroy((cudaGraphExec_t)graph_exec));
  return 0;
}

}  // extern "C"
