// engine.cu — C ABI (include/emrf.h) of the B200-native random-forest engine.
//
// Host side of RandomForest.run (ml/tree/impl/RandomForest.scala, Spark 2.4.6; dependency pin
// pom.xml:56-61): the sequential node scheduler (selectNodesToSplit: LIFO stack, memory-bounded
// groups, per-node feature subsets drawn from the driver RNG) stays on the host; everything
// row-sized runs in the kernels of kernels.cuh.  Per group the GPU runs histogram+scan, the decide
// kernels and the partition kernels back to back; the host gets a 16-byte scheduling record per node
// first (what it needs to number the children and select the next group; stored by the deciding kernel
// straight into page-locked host memory) and the full split records (~0.5 KB per node) behind them on a
// second stream, and builds the tree nodes one group later, under GPU work.  Several GPUs: rows sharded
// (reduce-scatter of the node histograms keyed by node, light nodes packed to 16 bits, groups tiled on
// the device) or trees sharded; see DESIGN.md §5.
//
// No PyTorch, no CPU fallback: every compute entry point needs a CUDA device.
#include <cuda_runtime.h>
#include <dlfcn.h>

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <limits>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <utility>
#include <vector>

#include "../../include/emrf.h"
#include "host_logic.hpp"
#include "kernels.cuh"

using namespace emrf;
namespace hl = emrf::hostlogic;

// ---------------------------------------------------------------------------------------------
// errors
// ---------------------------------------------------------------------------------------------
static thread_local std::string g_err;
static int fail(int code, const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  g_err = buf;
  return code;
}
#define CU(expr)                                                                                   \
  do {                                                                                             \
    cudaError_t _e = (expr);                                                                       \
    if (_e != cudaSuccess)                                                                         \
      return fail(EMRF_ERR_CUDA, "CUDA error %s at %s:%d: %s", cudaGetErrorName(_e), __FILE__, __LINE__, \
                  cudaGetErrorString(_e));                                                         \
  } while (0)
#define TRY(expr)           \
  do {                      \
    int _rc = (expr);       \
    if (_rc != EMRF_OK) return _rc; \
  } while (0)

extern "C" const char* emrf_last_error(void) { return g_err.c_str(); }
extern "C" const char* emrf_version(void) { return "emrf 0.1 (sm_100a; Spark MLlib 2.4.6 RandomForestClassifier semantics)"; }

// ---------------------------------------------------------------------------------------------
// small RAII helpers
// ---------------------------------------------------------------------------------------------
// Device allocations go through a small caching layer: a released block is parked in a per-device free
// list (size classes: powers of two up to 64 MB, 2 MB multiples above) instead of going back to the
// driver, because cudaFree synchronises the whole device and cudaMalloc/cudaFree of many buffers per
// fit showed up as 10-100 ms outliers.  Every release in this file happens after the stream that used
// the block has been synchronised.  emrf_ctx_destroy trims the cache of its device.
class DevCache {
 public:
  static DevCache& get() { static DevCache c; return c; }
  static size_t round_up(size_t n) {
    if (n < 256) n = 256;
    if (n <= ((size_t)64 << 20)) { size_t p = 256; while (p < n) p <<= 1; return p; }
    return (n + ((size_t)2 << 20) - 1) & ~(((size_t)2 << 20) - 1);
  }
  cudaError_t take(size_t n, void** out, size_t* cap, int* dev_out) {   // on the CURRENT device; *dev_out = that device
    const size_t want = round_up(n);
    int dev = 0;
    cudaGetDevice(&dev);
    *dev_out = dev;
    {
      std::lock_guard<std::mutex> lk(mu_);
      auto it = free_.find(std::make_pair(dev, want));
      if (it != free_.end()) { *out = it->second; *cap = want; free_.erase(it); return cudaSuccess; }
    }
    cudaError_t e = cudaMalloc(out, want);
    if (e != cudaSuccess) {       // out of memory: give cached blocks back and retry once
      cudaGetLastError();
      trim(dev);
      e = cudaMalloc(out, want);
    }
    if (e == cudaSuccess) *cap = want;
    return e;
  }
  void give(void* p, size_t cap, int dev) {   // dev: the device the block was allocated on (not the caller's current one)
    std::lock_guard<std::mutex> lk(mu_);
    free_.insert(std::make_pair(std::make_pair(dev, cap), p));
  }
  void trim(int dev) {
    std::vector<void*> blocks;
    {
      std::lock_guard<std::mutex> lk(mu_);
      for (auto it = free_.begin(); it != free_.end();) {
        if (it->first.first == dev) { blocks.push_back(it->second); it = free_.erase(it); } else ++it;
      }
    }
    for (void* p : blocks) cudaFree(p);
  }

 private:
  std::mutex mu_;
  std::multimap<std::pair<int, size_t>, void*> free_;
};

struct DevBuf {
  void* p = nullptr;
  size_t bytes = 0;   // capacity
  int dev = -1;       // the device that owns the block: it goes back to that device's free list whoever releases it
  DevBuf() {}
  DevBuf(const DevBuf&) = delete;
  DevBuf& operator=(const DevBuf&) = delete;
  DevBuf(DevBuf&& o) noexcept : p(o.p), bytes(o.bytes), dev(o.dev) { o.p = nullptr; o.bytes = 0; o.dev = -1; }
  ~DevBuf() { release(); }
  void release() {
    if (p) DevCache::get().give(p, bytes, dev);
    p = nullptr;
    bytes = 0;
    dev = -1;
  }
  cudaError_t alloc(size_t n) {   // on the current device; a block that lives on another device is released, not reused
    int cur = 0;
    cudaGetDevice(&cur);
    if (p && n <= bytes && dev == cur) return cudaSuccess;
    release();
    return DevCache::get().take(n, &p, &bytes, &dev);
  }
  cudaError_t ensure(size_t n) {  // grow-only (contexts own their buffers: one device for life)
    if (n <= bytes) return cudaSuccess;
    return alloc(n + n / 4);
  }
  cudaError_t ensure_exact(size_t n) {  // grow-only, no slack (GB-sized buffers)
    if (n <= bytes) return cudaSuccess;
    return alloc(n);
  }
  template <class T>
  T* as() const { return reinterpret_cast<T*>(p); }
};
struct PinnedBuf {
  void* p = nullptr;
  size_t bytes = 0;
  ~PinnedBuf() { if (p) cudaFreeHost(p); }
  cudaError_t ensure(size_t n) {
    if (n <= bytes) return cudaSuccess;
    if (p) cudaFreeHost(p);
    p = nullptr;
    bytes = 0;
    n += n / 4;
    cudaError_t e = cudaHostAlloc(&p, n, cudaHostAllocDefault);
    if (e == cudaSuccess) bytes = n;
    return e;
  }
  template <class T>
  T* as() const { return reinterpret_cast<T*>(p); }
};
static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }
static inline double now_ms() {
  return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}
static bool env_flag(const char* name) {   // set and not empty / "0"
  const char* v = getenv(name);
  return v && *v && strcmp(v, "0") != 0;
}
static bool is_device_ptr(const void* p) {
  if (!p) return false;
  cudaPointerAttributes a;
  cudaError_t e = cudaPointerGetAttributes(&a, p);
  if (e != cudaSuccess) { cudaGetLastError(); return false; }
  return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

// ---------------------------------------------------------------------------------------------
// context
// ---------------------------------------------------------------------------------------------
struct EventPool {
  std::vector<cudaEvent_t> ev;
  size_t used = 0;
  cudaEvent_t get() {
    if (used == ev.size()) {
      cudaEvent_t e;
      cudaEventCreate(&e);
      ev.push_back(e);
    }
    return ev[used++];
  }
  void reset() { used = 0; }
  ~EventPool() { for (auto e : ev) cudaEventDestroy(e); }
};

// host-side tree under construction (LearningNode)
struct HNode {
  int32_t id = 1;
  int32_t left = -1, right = -1;
  bool is_leaf = false, valid = true;
  int32_t pred = 0;   // first index of the largest class count (from the scan kernel)
  int32_t feature = -1, split_bin = -1;
  double threshold = 0.0, gain = std::numeric_limits<double>::quiet_NaN(), impurity = 0.0;
  // class counts = cnt_a[c] - (cnt_b ? cnt_b[c] : 0); both point into the fit's result arena (the
  // split records copied back from the device): a left child shares its parent's "left" array, a
  // right child is parent total minus parent left, a node that went through a group owns its totals.
  const uint32_t* cnt_a = nullptr;
  const uint32_t* cnt_b = nullptr;
};
// node pool of one tree: grows without constructing (the scheduler's parallel pass builds the children in place)
class NodePool {
 public:
  NodePool() = default;
  NodePool(NodePool&& o) noexcept : p_(o.p_), n_(o.n_), cap_(o.cap_) { o.p_ = nullptr; o.n_ = o.cap_ = 0; }
  NodePool& operator=(NodePool&& o) noexcept { std::swap(p_, o.p_); std::swap(n_, o.n_); std::swap(cap_, o.cap_); return *this; }
  NodePool(const NodePool&) = delete;
  NodePool& operator=(const NodePool&) = delete;
  ~NodePool() { free(p_); }
  size_t size() const { return n_; }
  void clear() { n_ = 0; }
  void reserve(size_t cap) {
    if (cap <= cap_) return;
    HNode* q = (HNode*)realloc((void*)p_, cap * sizeof(HNode));   // HNode is trivially copyable
    if (!q) throw std::bad_alloc();
    p_ = q; cap_ = cap;
  }
  void resize_uninit(size_t n) { if (n > cap_) reserve(std::max(n, cap_ * 2)); n_ = n; }
  HNode& emplace_back() { resize_uninit(n_ + 1); return *new (&p_[n_ - 1]) HNode(); }
  HNode& operator[](size_t i) { return p_[i]; }
  const HNode& operator[](size_t i) const { return p_[i]; }
 private:
  HNode* p_ = nullptr;
  size_t n_ = 0, cap_ = 0;
};
struct HTree {
  NodePool pool;
};
struct emrf_ctx {
  int device = 0;
  int rank = 0, world = 1;
  cudaStream_t stream = nullptr;
  cudaStream_t stream2 = nullptr;   // bagging runs here, underneath input copy / findSplits / binning
  cudaStream_t stream3 = nullptr;   // chunked host -> device copy of the feature matrix (binning follows chunk by chunk)
  cudaStream_t stream_comm = nullptr;   // rows sharded: per-piece reduce-scatter + owner-side scan, under the next piece's histogram kernel
  int sm_count = 148;
  int host_threads = 1;   // OpenMP threads for the host scheduler's parallel pieces (cores / ranks on this node)
  size_t smem_optin = 0;
  emrf_timings_t tm;
  std::mutex fit_mu;
  EventPool events;
  // NCCL (loaded lazily with dlopen; only multi-GPU contexts touch it)
  void* nccl_lib = nullptr;
  void* nccl_comm = nullptr;
  // grow-only scratch reused across groups
  DevBuf d_group, d_results, d_ghist, d_tickets;
  PinnedBuf h_group, h_part, h_seeds, h_table, h_tsum, h_sample, h_sched, h_seg;
  size_t last_nodes_split = 0, last_trace_len = 0;   // sizes the previous fit ended with (reserve hints)
  PinnedBuf h_export;   // world > 4: this rank's trees, exported, on their way to the other ranks
  DevBuf d_export;
  DevBuf d_split, d_part;   // partition work of the current group (written by decide_kernel)
  std::vector<HTree> tree_pools;   // LearningNode pools of the last fit (memory reused by the next)
  DevBuf d_seeds, d_table, d_tsum;
  // fit-sized buffers kept between fits (grow-only): cudaMalloc/cudaFree of GB-sized buffers costs
  // milliseconds and cudaFree synchronises the device; they are released by emrf_ctx_destroy
  DevBuf p_w, p_X, p_y, p_labels, p_rm, p_cm, p_ent0, p_ent1, p_counts;
  // row-sharded push mode: this rank's staging buffer and the peers' buffers mapped through CUDA IPC
  DevBuf p_stage, p_ipc;
  DevBuf p_segtab;   // device segment table (kern::Seg per node slot)
  void* peer_stage[8] = {nullptr};
  size_t peer_bytes = 0;
  bool peer_ok = false;
  // pinned arena that keeps every group's split records for the whole fit (nodes point into it)
  std::vector<std::unique_ptr<PinnedBuf>> arena;
  size_t arena_chunk = 0, arena_used = 0;
  void arena_reset() { arena_chunk = 0; arena_used = 0; }
  void* arena_alloc(size_t bytes) {
    bytes = (bytes + 255) & ~(size_t)255;
    while (true) {
      if (arena_chunk < arena.size()) {
        PinnedBuf& b = *arena[arena_chunk];
        if (arena_used + bytes <= b.bytes) { void* p = (char*)b.p + arena_used; arena_used += bytes; return p; }
        arena_chunk++; arena_used = 0;
        continue;
      }
      std::unique_ptr<PinnedBuf> nb(new PinnedBuf());
      if (nb->ensure(std::max<size_t>(bytes, (size_t)48 << 20)) != cudaSuccess) { cudaGetLastError(); return nullptr; }
      arena.push_back(std::move(nb));
      arena_used = 0;
    }
  }
};

static int ctx_init(emrf_ctx* c, int device) {
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0) {
    cudaGetLastError();
    return fail(EMRF_ERR_CUDA, "no CUDA device available (%s): this engine has no CPU fallback",
                e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
  }
  if (device < 0 || device >= n) return fail(EMRF_ERR_ARG, "device %d out of range (have %d)", device, n);
  c->device = device;
  CU(cudaSetDevice(device));
  cudaDeviceProp prop;
  CU(cudaGetDeviceProperties(&prop, device));
  if (prop.major < 10)
    return fail(EMRF_ERR_CUDA, "device %d is sm_%d%d; this library is built for sm_100a only", device, prop.major,
                prop.minor);
  c->sm_count = prop.multiProcessorCount;
  c->smem_optin = prop.sharedMemPerBlockOptin;
  CU(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
  CU(cudaStreamCreateWithFlags(&c->stream2, cudaStreamNonBlocking));
  CU(cudaStreamCreateWithFlags(&c->stream3, cudaStreamNonBlocking));
  CU(cudaStreamCreateWithFlags(&c->stream_comm, cudaStreamNonBlocking));
  memset(&c->tm, 0, sizeof(c->tm));
  c->host_threads = std::max(1u, std::min(32u, std::thread::hardware_concurrency()));
  if (const char* e = getenv("EMRF_HOST_THREADS")) c->host_threads = std::max(1, atoi(e));   // experiments
  return EMRF_OK;
}

extern "C" int emrf_ctx_create(int32_t device, emrf_ctx_t** out) {
  if (!out) return fail(EMRF_ERR_ARG, "out is NULL");
  emrf_ctx* c = new emrf_ctx();
  int rc = ctx_init(c, device);
  if (rc != EMRF_OK) { delete c; return rc; }
  *out = c;
  return EMRF_OK;
}

// ---- NCCL via dlopen (kept out of the single-GPU link) ----------------------------------------
typedef struct { char internal[128]; } nccl_uid_t;
typedef int (*ncclGetUniqueId_t)(nccl_uid_t*);
typedef int (*ncclCommInitRank_t)(void**, int, nccl_uid_t, int);
typedef int (*ncclCommDestroy_t)(void*);
typedef int (*ncclAllReduce_t)(const void*, void*, size_t, int, int, void*, cudaStream_t);
typedef const char* (*ncclGetErrorString_t)(int);
typedef int (*ncclReduceScatter_t)(const void*, void*, size_t, int, int, void*, cudaStream_t);
static void* nccl_open() {
  static void* lib = nullptr;
  if (!lib) lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
  if (!lib) lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
  return lib;
}
extern "C" int emrf_nccl_unique_id(uint8_t* uid128) {
  void* lib = nccl_open();
  if (!lib) return fail(EMRF_ERR_CUDA, "libnccl.so.2 not found: %s", dlerror());
  auto f = (ncclGetUniqueId_t)dlsym(lib, "ncclGetUniqueId");
  if (!f) return fail(EMRF_ERR_CUDA, "ncclGetUniqueId missing");
  nccl_uid_t u;
  int rc = f(&u);
  if (rc) return fail(EMRF_ERR_CUDA, "ncclGetUniqueId failed (%d)", rc);
  memcpy(uid128, &u, 128);
  return EMRF_OK;
}
extern "C" int emrf_ctx_create_dist(int32_t device, int32_t rank, int32_t world, const uint8_t* nccl_uid,
                                    emrf_ctx_t** out) {
  if (!out) return fail(EMRF_ERR_ARG, "out is NULL");
  if (world < 1 || rank < 0 || rank >= world) return fail(EMRF_ERR_ARG, "bad rank/world %d/%d", rank, world);
  emrf_ctx* c = new emrf_ctx();
  int rc = ctx_init(c, device);
  if (rc != EMRF_OK) { delete c; return rc; }
  c->rank = rank;
  c->world = world;
  // launchers such as torchrun export OMP_NUM_THREADS=1; size the host pool explicitly instead: this node's cores shared
  // by the ranks (EMRF_HOST_THREADS, when set, is already the per-rank count)
  if (!getenv("EMRF_HOST_THREADS")) c->host_threads = std::max(1, c->host_threads / std::max(1, world));
  if (world > 1) {
    if (!nccl_uid) { delete c; return fail(EMRF_ERR_ARG, "nccl_uid is NULL"); }
    void* lib = nccl_open();
    if (!lib) { delete c; return fail(EMRF_ERR_CUDA, "libnccl.so.2 not found: %s", dlerror()); }
    c->nccl_lib = lib;
    auto init = (ncclCommInitRank_t)dlsym(lib, "ncclCommInitRank");
    nccl_uid_t u;
    memcpy(&u, nccl_uid, 128);
    int nrc = init(&c->nccl_comm, world, u, rank);
    if (nrc) { delete c; return fail(EMRF_ERR_CUDA, "ncclCommInitRank failed (%d)", nrc); }
  }
  *out = c;
  return EMRF_OK;
}
static int nccl_allreduce(emrf_ctx* c, void* buf, size_t count, bool int32_max, cudaStream_t stream);
extern "C" void emrf_ctx_destroy(emrf_ctx_t* c) {
  if (!c) return;
  cudaSetDevice(c->device);
  for (int r = 0; r < 8; r++)
    if (r != c->rank && c->peer_stage[r]) cudaIpcCloseMemHandle(c->peer_stage[r]);
  // EMRF_PEER_PUSH: this rank's staging buffer is mapped by its peers.  It may be freed only when every rank has
  // finished with it and closed its mappings: a 1-word all-reduce is the barrier (contexts of one communicator are
  // destroyed together, as they are created)
  if (c->peer_ok && c->nccl_comm && c->stream && c->p_ipc.p) {
    cudaStreamSynchronize(c->stream);
    cudaMemsetAsync(c->p_ipc.p, 0, 4, c->stream);
    nccl_allreduce(c, c->p_ipc.p, 1, false, c->stream);
    cudaStreamSynchronize(c->stream);
  }
  if (c->nccl_comm) {
    auto d = (ncclCommDestroy_t)dlsym(c->nccl_lib, "ncclCommDestroy");
    if (d) d(c->nccl_comm);
  }
  if (c->stream) { cudaStreamSynchronize(c->stream); cudaStreamDestroy(c->stream); }
  if (c->stream2) { cudaStreamSynchronize(c->stream2); cudaStreamDestroy(c->stream2); }
  if (c->stream3) { cudaStreamSynchronize(c->stream3); cudaStreamDestroy(c->stream3); }
  if (c->stream_comm) { cudaStreamSynchronize(c->stream_comm); cudaStreamDestroy(c->stream_comm); }
  const int dev = c->device;
  delete c;
  DevCache::get().trim(dev);
}
extern "C" int emrf_ctx_release_buffers(emrf_ctx_t* c, int64_t* bytes_released) {
  if (!c) return fail(EMRF_ERR_ARG, "ctx is NULL");
  std::lock_guard<std::mutex> lk(c->fit_mu);
  CU(cudaSetDevice(c->device));
  for (cudaStream_t s_ : {c->stream, c->stream2, c->stream3, c->stream_comm})
    if (s_) CU(cudaStreamSynchronize(s_));
  size_t before = 0, after = 0, total = 0;
  CU(cudaMemGetInfo(&before, &total));
  // (the peer-mapped staging buffer of EMRF_PEER_PUSH stays: its peers have it mapped)
  for (DevBuf* b : {&c->d_group, &c->d_results, &c->d_ghist, &c->d_tickets, &c->d_export, &c->d_split, &c->d_part, &c->d_seeds,
                    &c->d_table, &c->d_tsum, &c->p_w, &c->p_X, &c->p_y, &c->p_labels, &c->p_rm, &c->p_cm, &c->p_ent0, &c->p_ent1,
                    &c->p_counts, &c->p_segtab})
    b->release();
  DevCache::get().trim(c->device);
  CU(cudaMemGetInfo(&after, &total));
  if (bytes_released) *bytes_released = (int64_t)after - (int64_t)before;
  return EMRF_OK;
}
extern "C" int emrf_get_timings(const emrf_ctx_t* c, emrf_timings_t* out) {
  if (!c || !out) return fail(EMRF_ERR_ARG, "NULL argument");
  *out = c->tm;
  return EMRF_OK;
}
// in-place all-reduce over the ctx communicator: uint32 sum (histograms, split records, raw bit
// patterns against zeros) or int32 max (label flags)
static int nccl_allreduce(emrf_ctx* c, void* buf, size_t count, bool int32_max, cudaStream_t stream) {
  if (c->world == 1) return EMRF_OK;
  static ncclAllReduce_t ar = nullptr;
  if (!ar) ar = (ncclAllReduce_t)dlsym(c->nccl_lib, "ncclAllReduce");
  int rc = ar(buf, buf, count, int32_max ? /*ncclInt32*/ 2 : /*ncclUint32*/ 3, int32_max ? /*ncclMax*/ 2 : /*ncclSum*/ 0,
              c->nccl_comm, stream ? stream : c->stream);
  if (rc) return fail(EMRF_ERR_CUDA, "ncclAllReduce failed (%d)", rc);
  return EMRF_OK;
}
static int nccl_allreduce_u32(emrf_ctx* c, uint32_t* buf, size_t count, cudaStream_t stream = nullptr) { return nccl_allreduce(c, buf, count, false, stream); }
// in-place reduce-scatter (uint32 sum): rank r ends up with the sum of words [r*per_rank, (r+1)*per_rank)
static int nccl_reduce_scatter_u32(emrf_ctx* c, uint32_t* buf, size_t per_rank, cudaStream_t stream = nullptr) {
  static ncclReduceScatter_t rs = nullptr;
  if (!rs) rs = (ncclReduceScatter_t)dlsym(c->nccl_lib, "ncclReduceScatter");
  if (!rs) return fail(EMRF_ERR_CUDA, "ncclReduceScatter missing");
  int rc = rs(buf, buf + (size_t)c->rank * per_rank, per_rank, /*ncclUint32*/ 3, /*ncclSum*/ 0, c->nccl_comm, stream ? stream : c->stream);
  if (rc) return fail(EMRF_ERR_CUDA, "ncclReduceScatter failed (%d)", rc);
  return EMRF_OK;
}
// in-place all-gather of bytes: rank r's block is [r*per_rank, (r+1)*per_rank)
static int nccl_all_gather_bytes(emrf_ctx* c, void* buf, size_t per_rank, cudaStream_t stream = nullptr) {
  typedef int (*ncclAllGather_t)(const void*, void*, size_t, int, void*, cudaStream_t);
  static ncclAllGather_t ag = nullptr;
  if (!ag) ag = (ncclAllGather_t)dlsym(c->nccl_lib, "ncclAllGather");
  if (!ag) return fail(EMRF_ERR_CUDA, "ncclAllGather missing");
  int rc = ag((const char*)buf + (size_t)c->rank * per_rank, buf, per_rank, /*ncclInt8*/ 0, c->nccl_comm, stream ? stream : c->stream);
  if (rc) return fail(EMRF_ERR_CUDA, "ncclAllGather failed (%d)", rc);
  return EMRF_OK;
}

// Staging buffers of the fused histogram + reduce-scatter path: every rank allocates `bytes`, the CUDA IPC
// handles are exchanged with an all-reduce over zero-filled slots, and each rank maps its peers' buffers.
// Collective: every rank must call it with the same `bytes`.
static int peer_stage_setup(emrf_ctx* c, size_t bytes) {
  if (c->world > 8) return fail(EMRF_ERR_LIMIT, "peer staging supports up to 8 ranks");
  if (c->peer_ok && c->peer_bytes >= bytes) return EMRF_OK;
  cudaStream_t st = c->stream;
  for (int r = 0; r < c->world; r++)
    if (r != c->rank && c->peer_stage[r]) { cudaIpcCloseMemHandle(c->peer_stage[r]); c->peer_stage[r] = nullptr; }
  c->peer_ok = false;
  CU(c->p_ipc.ensure((size_t)c->world * 64 + 64));
  {  // nobody may free a buffer a peer still has mapped: everybody has closed before anybody re-allocates
    CU(cudaMemsetAsync(c->p_ipc.p, 0, 4, st));
    TRY(nccl_allreduce_u32(c, c->p_ipc.as<uint32_t>(), 1));
    CU(cudaStreamSynchronize(st));
  }
  CU(c->p_stage.ensure_exact(bytes));
  cudaIpcMemHandle_t mine;
  CU(cudaIpcGetMemHandle(&mine, c->p_stage.p));
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  std::vector<uint8_t> all((size_t)c->world * 64, 0);
  memcpy(&all[(size_t)c->rank * 64], &mine, 64);
  CU(cudaMemcpyAsync(c->p_ipc.p, all.data(), all.size(), cudaMemcpyHostToDevice, st));
  TRY(nccl_allreduce_u32(c, c->p_ipc.as<uint32_t>(), all.size() / 4));
  CU(cudaMemcpyAsync(all.data(), c->p_ipc.p, all.size(), cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  uint32_t failed = 0;
  std::string why;
  for (int r = 0; r < c->world && !failed; r++) {
    if (r == c->rank) { c->peer_stage[r] = c->p_stage.p; continue; }
    cudaIpcMemHandle_t h;
    memcpy(&h, &all[(size_t)r * 64], 64);
    cudaError_t e = cudaIpcOpenMemHandle(&c->peer_stage[r], h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
      cudaGetLastError();
      c->peer_stage[r] = nullptr;
      failed = 1;
      why = std::string("cudaIpcOpenMemHandle(rank ") + std::to_string(r) + ") failed: " + cudaGetErrorString(e);
    }
  }
  // every rank proceeds or every rank fails: a rank that returned alone would leave the others in the next collective
  CU(cudaMemcpyAsync(c->p_ipc.p, &failed, 4, cudaMemcpyHostToDevice, st));
  TRY(nccl_allreduce_u32(c, c->p_ipc.as<uint32_t>(), 1));
  uint32_t any_failed = 0;
  CU(cudaMemcpyAsync(&any_failed, c->p_ipc.p, 4, cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  if (any_failed) {
    for (int r = 0; r < c->world; r++)
      if (r != c->rank && c->peer_stage[r]) { cudaIpcCloseMemHandle(c->peer_stage[r]); c->peer_stage[r] = nullptr; }
    return fail(EMRF_ERR_CUDA, "%s", failed ? why.c_str() : "a peer rank could not map the staging buffers");
  }
  c->peer_bytes = c->p_stage.bytes >= bytes ? bytes : 0;
  c->peer_ok = true;
  return EMRF_OK;
}

// Row-sharded fit (shard_mode 1): this rank holds the rows of partitions
// [first_part, first_part + n_parts_local) of a dataset of n_rows_global rows in n_parts_global partitions.
struct ShardSpec {
  int first_part = 0;
  int64_t n_rows_global = 0;
  const int64_t* goff = nullptr;   // global partition offsets, n_parts_global + 1
  int n_parts_global = 0;
  int64_t row0 = 0;                // global index of the first local row
};

extern "C" int emrf_dev_alloc(emrf_ctx_t* c, int64_t bytes, void** out) {
  if (!c || !out || bytes < 0) return fail(EMRF_ERR_ARG, "bad argument");
  CU(cudaSetDevice(c->device));
  CU(cudaMalloc(out, (size_t)std::max<int64_t>(bytes, 16)));
  return EMRF_OK;
}
extern "C" int emrf_dev_free(emrf_ctx_t* c, void* p) {
  if (!c) return fail(EMRF_ERR_ARG, "ctx is NULL");
  CU(cudaSetDevice(c->device));
  CU(cudaFree(p));
  return EMRF_OK;
}
extern "C" int emrf_host_alloc_pinned(int64_t bytes, void** out) {
  if (!out || bytes < 0) return fail(EMRF_ERR_ARG, "bad argument");
  CU(cudaHostAlloc(out, (size_t)std::max<int64_t>(bytes, 16), cudaHostAllocDefault));
  return EMRF_OK;
}
extern "C" int emrf_host_free_pinned(void* p) {
  CU(cudaFreeHost(p));
  return EMRF_OK;
}
extern "C" int emrf_memcpy_h2d(emrf_ctx_t* c, void* dst, const void* src, int64_t bytes) {
  if (!c) return fail(EMRF_ERR_ARG, "ctx is NULL");
  CU(cudaSetDevice(c->device));
  CU(cudaMemcpyAsync(dst, src, (size_t)bytes, cudaMemcpyHostToDevice, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return EMRF_OK;
}
extern "C" int emrf_memcpy_d2h(emrf_ctx_t* c, void* dst, const void* src, int64_t bytes) {
  if (!c) return fail(EMRF_ERR_ARG, "ctx is NULL");
  CU(cudaSetDevice(c->device));
  CU(cudaMemcpyAsync(dst, src, (size_t)bytes, cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return EMRF_OK;
}

// ---------------------------------------------------------------------------------------------
// parameter checks (the `require`s of ml/tree/treeParams.scala and Strategy.assertValid)
// ---------------------------------------------------------------------------------------------
static int check_params(const emrf_params_t* p) {
  if (!p) return fail(EMRF_ERR_ARG, "params is NULL");
  if (p->num_trees < 1) return fail(EMRF_ERR_ARG, "numTrees must be >= 1 (got %d)", p->num_trees);
  if (p->max_depth < 0 || p->max_depth > 30) return fail(EMRF_ERR_ARG, "maxDepth must be in [0,30] (got %d)", p->max_depth);
  if (p->max_bins < 2) return fail(EMRF_ERR_ARG, "maxBins must be >= 2 (got %d)", p->max_bins);
  if (p->max_bins > 256) return fail(EMRF_ERR_LIMIT, "maxBins > 256 is not supported by the uint8 TreePoint matrix");
  if (p->min_instances_per_node < 1) return fail(EMRF_ERR_ARG, "minInstancesPerNode must be >= 1");
  if (!(p->subsampling_rate > 0.0 && p->subsampling_rate <= 1.0)) return fail(EMRF_ERR_ARG, "subsamplingRate must be in (0,1]");
  if (p->impurity != 0 && p->impurity != 1) return fail(EMRF_ERR_ARG, "impurity must be gini(0) or entropy(1)");
  if (p->feature_subset_kind < 0 || p->feature_subset_kind > 6) return fail(EMRF_ERR_ARG, "bad featureSubsetStrategy");
  if (p->feature_subset_kind == 5 && p->feature_subset_value < 1) return fail(EMRF_ERR_ARG, "featureSubsetStrategy n must be >= 1");
  if (p->feature_subset_kind == 6 && !(p->feature_subset_value > 0 && p->feature_subset_value <= 1.0))
    return fail(EMRF_ERR_ARG, "featureSubsetStrategy fraction must be in (0,1]");
  if (p->max_memory_mb < 0) return fail(EMRF_ERR_ARG, "maxMemoryInMB must be >= 0");
  if (p->num_classes < 0 || p->num_classes > 256) return fail(EMRF_ERR_LIMIT, "numClasses must be <= 256");
  return EMRF_OK;
}
static int check_partitions(int64_t n_rows, const int64_t* off, int n_parts) {
  if (!off || n_parts < 1) return fail(EMRF_ERR_ARG, "partition_offsets missing");
  if (off[0] != 0 || off[n_parts] != n_rows) return fail(EMRF_ERR_ARG, "partition_offsets must span [0, n_rows]");
  for (int i = 0; i < n_parts; i++)
    if (off[i + 1] < off[i]) return fail(EMRF_ERR_ARG, "partition_offsets must be non-decreasing");
  return EMRF_OK;
}

// Bring a host-or-device input onto the device.  `owned` keeps the copy alive.
static int stage_input(emrf_ctx* c, const void* src_hd, size_t bytes, DevBuf& owned, const void** dev_out, bool* copied) {
  if (is_device_ptr(src_hd)) {
    *dev_out = src_hd;
    if (copied) *copied = false;
    return EMRF_OK;
  }
  CU(owned.ensure_exact(bytes));
  CU(cudaMemcpyAsync(owned.p, src_hd, bytes, cudaMemcpyHostToDevice, c->stream));
  c->tm.h2d_bytes += (int64_t)bytes;
  *dev_out = owned.p;
  if (copied) *copied = true;
  return EMRF_OK;
}

// ---------------------------------------------------------------------------------------------
// findSplits on the device
// ---------------------------------------------------------------------------------------------
// A feature matrix that is still on its way to the device: findSplits then takes its (few) sampled rows from
// the host copy instead of waiting for the transfer; `ready` fires when the whole matrix has arrived.
struct HostX {
  const double* h;
  cudaEvent_t ready;
};
static int run_find_splits(emrf_ctx* c, const emrf_params_t& P, const double* dX, int64_t N_local, int F, const int64_t* part_off_local,
                           int n_parts_local, DevBuf& d_thr, DevBuf& d_ns, std::vector<double>& h_thr, std::vector<int>& h_ns,
                           const ShardSpec* sh = nullptr, std::vector<DevBuf>* keep = nullptr, const HostX* hx = nullptr) {
  // `keep`: where temporaries are parked instead of being freed here — cudaFree synchronises the whole
  // device and would stall this stream behind the bagging kernel running on the other one.
  // metadata and sampling are defined on the WHOLE dataset
  const int64_t N = sh ? sh->n_rows_global : N_local;
  const int64_t* part_off = sh ? sh->goff : part_off_local;
  const int n_parts = sh ? sh->n_parts_global : n_parts_local;
  const int TS = P.max_bins;
  const int max_possible_bins = (int)std::min<int64_t>(P.max_bins, N);
  const int num_splits = max_possible_bins - 1;
  CU(d_thr.alloc((size_t)F * TS * 8));
  CU(d_ns.alloc((size_t)F * 4));
  CU(cudaMemsetAsync(d_thr.p, 0, (size_t)F * TS * 8, c->stream));
  CU(cudaMemsetAsync(d_ns.p, 0, (size_t)F * 4, c->stream));
  h_thr.assign((size_t)F * TS, 0.0);
  h_ns.assign(F, 0);
  if (num_splits >= 1) {
    std::vector<int64_t> rows;
    const double t_s0 = now_ms();
    const bool sampled = hl::find_splits_rows(P, N, part_off, n_parts, &rows);
    if (env_flag("EMRF_DEBUG_TIMERS")) fprintf(stderr, "[emrf] findSplits row sampler: %.3f ms (%zu rows)\n", now_ms() - t_s0, rows.size());
    const int64_t n_s64 = sampled ? (int64_t)rows.size() : N;
    if (n_s64 > (1 << 22)) return fail(EMRF_ERR_LIMIT, "findSplits sample too large (%lld)", (long long)n_s64);
    const int n_s = (int)n_s64;
    const double fraction = hl::find_splits_fraction(P, N);
    const int64_t expected = (int64_t)(int32_t)(fraction * (double)N);
    if (n_s > 0) {
      int n_pad = 2;
      while (n_pad < n_s) n_pad <<= 1;
      DevBuf d_rows, d_S;
      if (sampled && !hx) {
        CU(d_rows.alloc((size_t)n_s * 8));
        CU(cudaMemcpyAsync(d_rows.p, rows.data(), (size_t)n_s * 8, cudaMemcpyHostToDevice, c->stream));
      }
      CU(d_S.alloc((size_t)F * n_pad * 8));
      const int64_t total = (int64_t)n_pad * F;
      // rows gathered on the host (matrix still in flight) into a pinned, device-mapped block that the gather
      // kernel reads straight over PCIe: a copy would queue behind the matrix chunks in the copy engine
      const double* d_comp = nullptr;
      auto host_gather = [&](const int64_t* local_rows, int n) -> int {
        CU(c->h_sample.ensure((size_t)std::max(n, 1) * F * 8));
        double* hb = c->h_sample.as<double>();
#pragma omp parallel for num_threads(c->host_threads) schedule(static)
        for (int j = 0; j < n; j++) memcpy(hb + (size_t)j * F, hx->h + (size_t)local_rows[j] * F, (size_t)F * 8);
        void* dp = nullptr;
        CU(cudaHostGetDevicePointer(&dp, hb, 0));
        d_comp = (const double*)dp;
        c->tm.h2d_bytes += (int64_t)n * F * 8;
        return EMRF_OK;
      };
      if (!sh) {
        if (hx && sampled) {
          TRY(host_gather(rows.data(), n_s));
          kern::fs_gather_kernel<<<(unsigned)((total + 255) / 256), 256, 0, c->stream>>>(
              d_comp, F, nullptr, n_s, n_pad, 1, P.find_splits_policy == 0 ? 1 : 0, d_S.as<double>());
        } else {
          if (hx) CU(cudaStreamWaitEvent(c->stream, hx->ready, 0));   // every row is a sample: the whole matrix is needed
          kern::fs_gather_kernel<<<(unsigned)((total + 255) / 256), 256, 0, c->stream>>>(
              dX, F, d_rows.as<int64_t>(), n_s, n_pad, sampled ? 0 : 1, P.find_splits_policy == 0 ? 1 : 0, d_S.as<double>());
        }
      } else {
        // own samples → global sample positions, all-reduce the bit patterns, then mark sentinels
        std::vector<int64_t> own_rows;
        std::vector<int32_t> own_pos;
        const int64_t lo = sh->row0, hi = sh->row0 + N_local;
        if (sampled) {
          for (int j = 0; j < n_s; j++)
            if (rows[j] >= lo && rows[j] < hi) { own_rows.push_back(rows[j] - lo); own_pos.push_back(j); }
        } else {
          for (int64_t r = lo; r < hi; r++) { own_rows.push_back(r - lo); own_pos.push_back((int32_t)r); }
        }
        const int n_own = (int)own_rows.size();
        const double* gX = dX;
        if (hx && sampled) {
          TRY(host_gather(own_rows.data(), n_own));
          for (int j = 0; j < n_own; j++) own_rows[j] = j;
          gX = d_comp;
        } else if (hx) {
          CU(cudaStreamWaitEvent(c->stream, hx->ready, 0));
        }
        DevBuf d_or, d_op;
        CU(d_or.alloc((size_t)n_own * 8));
        CU(d_op.alloc((size_t)n_own * 4));
        CU(cudaMemcpyAsync(d_or.p, own_rows.data(), (size_t)n_own * 8, cudaMemcpyHostToDevice, c->stream));
        CU(cudaMemcpyAsync(d_op.p, own_pos.data(), (size_t)n_own * 4, cudaMemcpyHostToDevice, c->stream));
        CU(cudaMemsetAsync(d_S.p, 0, (size_t)F * n_pad * 8, c->stream));
        if (n_own)
          kern::fs_gather_sharded_kernel<<<(unsigned)(((int64_t)n_own * F + 255) / 256), 256, 0, c->stream>>>(
              gX, F, d_or.as<int64_t>(), d_op.as<int32_t>(), n_own, n_pad, d_S.as<double>());
        const double t_ar0 = now_ms();
        if (env_flag("EMRF_DEBUG_TIMERS")) CU(cudaStreamSynchronize(c->stream));
        const double t_ar1 = now_ms();
        TRY(nccl_allreduce_u32(c, d_S.as<uint32_t>(), (size_t)F * n_pad * 2));
        if (env_flag("EMRF_DEBUG_TIMERS")) {
          CU(cudaStreamSynchronize(c->stream));
          fprintf(stderr, "[emrf] sharded findSplits: gather %.3f ms, all-reduce of %zu B %.3f ms\n", t_ar1 - t_ar0,
                  (size_t)F * n_pad * 8, now_ms() - t_ar1);
        }
        kern::fs_mark_kernel<<<(unsigned)((total + 255) / 256), 256, 0, c->stream>>>(d_S.as<double>(), F, n_s, n_pad,
                                                                                    P.find_splits_policy == 0 ? 1 : 0);
        c->tm.kernel_launches += 1;
        CU(cudaGetLastError());
        CU(cudaStreamSynchronize(c->stream));  // own_rows/own_pos die here
        if (keep) { keep->push_back(std::move(d_or)); keep->push_back(std::move(d_op)); }
      }
      const int use_smem = ((size_t)n_pad * 8 <= 128 * 1024) ? 1 : 0;
      const size_t smem = use_smem ? (size_t)n_pad * 8 : 0;
      if (smem > 48 * 1024)
        CU(cudaFuncSetAttribute(kern::fs_sort_walk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      kern::fs_sort_walk_kernel<<<F, 1024, smem, c->stream>>>(d_S.as<double>(), n_pad, n_s, use_smem, num_splits, expected,
                                                             P.find_splits_policy, TS, d_thr.as<double>(), d_ns.as<int>());
      c->tm.kernel_launches += 2;
      CU(cudaGetLastError());
      c->tm.d2h_bytes += (int64_t)F * TS * 8 + (int64_t)F * 4;
      c->tm.h2d_bytes += sampled ? (int64_t)n_s * 8 : 0;
      CU(cudaMemcpyAsync(h_thr.data(), d_thr.p, (size_t)F * TS * 8, cudaMemcpyDeviceToHost, c->stream));
      CU(cudaMemcpyAsync(h_ns.data(), d_ns.p, (size_t)F * 4, cudaMemcpyDeviceToHost, c->stream));
      CU(cudaStreamSynchronize(c->stream));  // rows dies here
      if (keep) { keep->push_back(std::move(d_rows)); keep->push_back(std::move(d_S)); }
    }
  }
  for (int f = 0; f < F; f++)
    if (h_ns[f] > TS || h_ns[f] >= 256)
      return fail(EMRF_ERR_LIMIT, "feature %d produced %d splits (> maxBins); not representable", f, h_ns[f]);
  return EMRF_OK;
}

// bin_kernel with the thresholds in shared memory, instantiated for the key table's row length
template <bool kNumeric, int KROW>
static int bin_launch_smem_t(unsigned grid, size_t smem, cudaStream_t st, const double* dX, int64_t N, int F, int Fp, int64_t Np,
                             const double* d_thr, const int* d_ns, int TS, int max_n, uint8_t* d_rm, uint8_t* d_cm) {
  if (smem > 48 * 1024) CU(cudaFuncSetAttribute(kern::bin_kernel<true, kNumeric, KROW>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kern::bin_kernel<true, kNumeric, KROW><<<grid, 256, smem, st>>>(dX, N, F, Fp, Np, d_thr, d_ns, TS, max_n, d_rm, d_cm);
  return EMRF_OK;
}
template <bool kNumeric>
static int bin_launch_smem(int krow, unsigned grid, size_t smem, cudaStream_t st, const double* dX, int64_t N, int F, int Fp, int64_t Np,
                           const double* d_thr, const int* d_ns, int TS, int max_n, uint8_t* d_rm, uint8_t* d_cm) {
  switch (krow) {
    case 16: return bin_launch_smem_t<kNumeric, 16>(grid, smem, st, dX, N, F, Fp, Np, d_thr, d_ns, TS, max_n, d_rm, d_cm);
    case 32: return bin_launch_smem_t<kNumeric, 32>(grid, smem, st, dX, N, F, Fp, Np, d_thr, d_ns, TS, max_n, d_rm, d_cm);
    case 64: return bin_launch_smem_t<kNumeric, 64>(grid, smem, st, dX, N, F, Fp, Np, d_thr, d_ns, TS, max_n, d_rm, d_cm);
    case 128: return bin_launch_smem_t<kNumeric, 128>(grid, smem, st, dX, N, F, Fp, Np, d_thr, d_ns, TS, max_n, d_rm, d_cm);
    default: return bin_launch_smem_t<kNumeric, 0>(grid, smem, st, dX, N, F, Fp, Np, d_thr, d_ns, TS, max_n, d_rm, d_cm);
  }
}
static int run_bin(emrf_ctx* c, const double* dX, int64_t N, int F, const double* d_thr, const int* d_ns, int TS, int max_n, int Fp,
                   int64_t Np, uint8_t* d_rm, uint8_t* d_cm) {   // max_n: upper bound on the features' split counts (<= TS)
  const size_t tile_bytes = kern::bin_tile_bytes(Fp);
  const size_t thr_bytes = kern::bin_thr_smem_bytes(F, max_n);
  // persistent CTAs: each stages the thresholds once and walks tiles with the grid's stride
  const unsigned grid = (unsigned)std::min<int64_t>((N + kern::kBinRows - 1) / kern::kBinRows, (int64_t)c->sm_count * 8);
  if (tile_bytes + thr_bytes <= 160 * 1024) {
    const size_t smem = tile_bytes + thr_bytes;
    TRY((bin_launch_smem<false>(kern::bin_key_row(F), grid, smem, c->stream, dX, N, F, Fp, Np, d_thr, d_ns, TS, max_n, d_rm, d_cm)));
  } else {
    if (tile_bytes > 200 * 1024) return fail(EMRF_ERR_LIMIT, "too many features for the binning tile (%d)", F);
    if (tile_bytes > 48 * 1024) CU(cudaFuncSetAttribute(kern::bin_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tile_bytes));
    kern::bin_kernel<false><<<grid, 256, tile_bytes, c->stream>>>(dX, N, F, Fp, Np, d_thr, d_ns, TS, max_n, d_rm, d_cm);
  }
  c->tm.kernel_launches += 1;
  CU(cudaGetLastError());
  return EMRF_OK;
}

// Bootstrap weights.  bag_launch only enqueues (on `stream`): it depends on nothing but the partition
// layout, so emrf_fit runs it on a second stream underneath the input copy, findSplits and binning.
struct BagJob {
  DevBuf d_off, d_max;
};
static int bag_launch(emrf_ctx* c, cudaStream_t stream, const emrf_params_t& P, int64_t N, const int64_t* part_off, int n_parts,
                      uint8_t* d_w, BagJob& job) {
  const int T = P.num_trees;
  CU(job.d_off.alloc((size_t)(n_parts + 1) * 8));
  CU(job.d_max.alloc(4));
  CU(cudaMemcpyAsync(job.d_off.p, part_off, (size_t)(n_parts + 1) * 8, cudaMemcpyHostToDevice, stream));
  CU(cudaMemsetAsync(job.d_max.p, 0, 4, stream));
  if (T > 1) {
    const double mean = P.subsampling_rate;
    const double p_exp = hl::poisson_exp_policy(mean);
    static const bool serial = env_flag("EMRF_BAG_SERIAL");  // one thread per partition (cross-check)
    if (serial) {
      CU(cudaFuncSetAttribute(kern::bag_poisson_serial_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kern::kBagSmemBytes));
      kern::bag_poisson_serial_kernel<<<(n_parts + 31) / 32, 32, kern::kBagSmemBytes, stream>>>(
          P.seed, job.d_off.as<int64_t>(), n_parts, T, mean, p_exp, d_w, job.d_max.as<int>());
    } else {
      const size_t smem = sizeof(kern::BagSmem);
      CU(cudaFuncSetAttribute(kern::bag_poisson_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      kern::bag_poisson_kernel<<<(n_parts + kern::kBagPairs - 1) / kern::kBagPairs, kern::kBagWarps * 32, smem, stream>>>(
          P.seed, job.d_off.as<int64_t>(), n_parts, T, mean, p_exp, d_w, job.d_max.as<int>());
    }
  } else if (P.subsampling_rate == 1.0) {
    CU(cudaMemsetAsync(d_w, 1, (size_t)N * T, stream));
    CU(cudaMemsetAsync(job.d_max.p, 0, 4, stream));
  } else {
    kern::bag_bernoulli_kernel<<<(n_parts + 63) / 64, 64, 0, stream>>>(P.seed, job.d_off.as<int64_t>(), n_parts, T,
                                                                         P.subsampling_rate, d_w);
  }
  c->tm.kernel_launches += 1;
  CU(cudaGetLastError());
  return EMRF_OK;
}
// after the stream that ran bag_launch has been waited for / synchronised
static int bag_finish(emrf_ctx* c, cudaStream_t stream, const emrf_params_t& P, BagJob& job, int* max_w_out) {
  int mw = 0;
  CU(cudaMemcpyAsync(&mw, job.d_max.p, 4, cudaMemcpyDeviceToHost, stream));
  CU(cudaStreamSynchronize(stream));
  if (P.num_trees == 1) mw = 1;
  if (max_w_out) *max_w_out = mw;
  if (mw > 255) return fail(EMRF_ERR_LIMIT, "a bootstrap weight exceeded 255");
  return EMRF_OK;
}
static int run_bag(emrf_ctx* c, const emrf_params_t& P, int64_t N, const int64_t* part_off, int n_parts, uint8_t* d_w,
                   int* max_w_out) {
  BagJob job;
  TRY(bag_launch(c, c->stream, P, N, part_off, n_parts, d_w, job));
  return bag_finish(c, c->stream, P, job, max_w_out);
}

// ---------------------------------------------------------------------------------------------
// histogram kernel dispatch (template on 16 B chunks per TreePoint row)
// ---------------------------------------------------------------------------------------------
template <int CH, class HT>
static cudaError_t hist_launch_t(dim3 grid, size_t smem, cudaStream_t st, const kern::HistArgs& A, bool prepare_only) {
  if (prepare_only)
    return cudaFuncSetAttribute(kern::hist_scan_kernel<CH, HT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  kern::hist_scan_kernel<CH, HT><<<grid, kern::kHistThreads, smem, st>>>(A);
  return cudaGetLastError();
}
// u16: 16-bit shared-memory counters (launches whose nodes are all single-tile and weigh <= 65535)
static cudaError_t hist_launch(int ch, dim3 grid, size_t smem, cudaStream_t st, const kern::HistArgs& A, bool prepare_only,
                               bool u16 = false) {
#define EMRF_HIST_CASE(n) \
  case n: return u16 ? hist_launch_t<n, uint16_t>(grid, smem, st, A, prepare_only) : hist_launch_t<n, uint32_t>(grid, smem, st, A, prepare_only);
  switch (ch) {
    EMRF_HIST_CASE(2) EMRF_HIST_CASE(4) EMRF_HIST_CASE(8)   // row_pitch(): 32 / 64 / 128 bytes, or wider rows (generic path, 0)
    default: return u16 ? hist_launch_t<0, uint16_t>(grid, smem, st, A, prepare_only) : hist_launch_t<0, uint32_t>(grid, smem, st, A, prepare_only);
  }
#undef EMRF_HIST_CASE
}
// Row pitch of the row-major TreePoint matrix: 32/64/128 bytes for up to 128 features (the sizes a TMA gather with a
// matching swizzle mode serves: one row = one swizzle span, four rows = a 128-byte aligned destination), F rounded up
// to 16 above that (generic byte gathers).
static int row_pitch(int F) {
  if (F <= 32) return 32;
  if (F <= 64) return 64;
  if (F <= 128) return 128;
  return (int)((F + 15) / 16 * 16);
}
// cuTensorMapEncodeTiled through the runtime's driver entry point lookup (no link-time libcuda dependency)
typedef CUresult (*tmap_encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                   const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static tmap_encode_fn tmap_encoder() {
  static tmap_encode_fn fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) {
      cudaGetLastError();
      p = nullptr;
    }
    return (tmap_encode_fn)p;
  }();
  return fn;
}
// Row staging by TMA gathers is the default wherever a row is 16/32/64/128 bytes; EMRF_NO_TMA=1 keeps the LDG/STS staging
// (A/B runs and the variant parity test).
static bool hist_use_tma(int ch) { return kern::hist_tma_rows(ch) && !env_flag("EMRF_NO_TMA") && tmap_encoder() != nullptr; }
// 2-D map over this GPU's bins_rm [rows][Fp] (uint8), box {Fp, 1}: tile::gather4 then moves four whole rows per instruction
static int make_rows_tmap(CUtensorMap* out, const uint8_t* bins_rm, int64_t rows, int Fp) {
  const cuuint64_t dims[2] = {(cuuint64_t)Fp, (cuuint64_t)rows}, strides[1] = {(cuuint64_t)Fp};
  const cuuint32_t box[2] = {(cuuint32_t)Fp, 1}, estr[2] = {1, 1};
  const CUtensorMapSwizzle sw = Fp == 128 ? CU_TENSOR_MAP_SWIZZLE_128B : Fp == 64 ? CU_TENSOR_MAP_SWIZZLE_64B
                              : Fp == 32 ? CU_TENSOR_MAP_SWIZZLE_32B : CU_TENSOR_MAP_SWIZZLE_NONE;
  const CUresult r = tmap_encoder()(out, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<uint8_t*>(bins_rm), dims, strides, box, estr,
                                    CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(EMRF_ERR_CUDA, "cuTensorMapEncodeTiled failed (%d) for %lld rows of %d bytes", (int)r, (long long)rows, Fp);
  return EMRF_OK;
}
// Shared-memory geometry of hist_scan_kernel: [staging region][histogram block of k_chunk features]
static int hist_geometry(emrf_ctx* c, int Fp, int f_sub, int B, int C, int* ch_out, int* k_chunk_out, int* n_chunks_out,
                         size_t* region_out, size_t* smem_out, bool* tma_out = nullptr) {
  const int ch = kern::hist_row_chunks(Fp);
  const bool tma = hist_use_tma(ch);
  const size_t smem_cap = std::min<size_t>(c->smem_optin, 227 * 1024);
  const size_t per_feat = (size_t)B * C * 4 + 8;
  const size_t fixed = kern::hist_region_bytes(ch, tma, std::min(f_sub, 64)) + 64;
  // two CTAs per SM (each may use (228 KB - 2 * 1 KB) / 2) unless even four features would not fit
  const size_t two_per_sm = 112 * 1024;
  int k_chunk = (two_per_sm > fixed) ? (int)((two_per_sm - fixed) / per_feat) : 0;
  if (k_chunk < f_sub && k_chunk < 4) k_chunk = (int)((smem_cap - fixed) / per_feat);  // big bins x classes: one CTA per SM
  if (k_chunk < 1) return fail(EMRF_ERR_LIMIT, "one feature histogram (%d bins x %d classes) exceeds shared memory", B, C);
  k_chunk = std::min(k_chunk, f_sub);
  if (ch) k_chunk = std::min(k_chunk, 64);   // the row-staging paths keep 64 per-feature entries in shared memory
  *ch_out = ch;
  *k_chunk_out = k_chunk;
  *n_chunks_out = (f_sub + k_chunk - 1) / k_chunk;
  *region_out = kern::hist_region_bytes(ch, tma, k_chunk);
  *smem_out = *region_out + kern::hist_block_bytes(k_chunk, B, C);
  if (tma_out) *tma_out = tma;
  if (*smem_out > smem_cap) return fail(EMRF_ERR_LIMIT, "histogram tile needs %zu B of shared memory", *smem_out);
  return EMRF_OK;
}

// ---------------------------------------------------------------------------------------------
// forest
// ---------------------------------------------------------------------------------------------
struct FlatNode {
  int32_t id, left, right, feature;
  double threshold, gain, impurity, prediction;
};
// Host blocks of freed forests are kept (process-wide, bounded) and handed to the next forest: a fit's
// model is tens of MB, above glibc's mmap threshold, so a fresh block would be page-faulted in 4 KB
// steps on every fit (the host-side counterpart of DevCache).
class HostBlockCache {
 public:
  static HostBlockCache& get() { static HostBlockCache g; return g; }
  // pinned: page-locked (cudaHostAlloc) blocks, for arrays that are the target of asynchronous device -> host copies
  void* take(size_t bytes, size_t* got, bool pinned = false) {
    {
      std::lock_guard<std::mutex> lk(mu_);
      int best = -1;
      for (int i = 0; i < (int)blocks_.size(); i++)
        if (blocks_[i].pinned == pinned && blocks_[i].bytes >= bytes && blocks_[i].bytes <= 2 * bytes + 4096 &&
            (best < 0 || blocks_[i].bytes < blocks_[best].bytes)) best = i;
      if (best >= 0) {
        void* p = blocks_[best].p;
        *got = blocks_[best].bytes;
        held_ -= blocks_[best].bytes;
        blocks_.erase(blocks_.begin() + best);
        return p;
      }
    }
    if (pinned) {
      const size_t n = bytes + bytes / 8 + 4096;   // slack: the next fit's model is about, not exactly, this size
      void* p = nullptr;
      if (cudaHostAlloc(&p, n, cudaHostAllocDefault) != cudaSuccess) { cudaGetLastError(); return nullptr; }
      *got = n;
      return p;
    }
    *got = bytes;
    return malloc(bytes ? bytes : 1);
  }
  void give(void* p, size_t bytes, bool pinned = false) {
    if (!p) return;
    {
      std::lock_guard<std::mutex> lk(mu_);
      if (bytes >= ((size_t)1 << 20) && held_ + bytes <= kMaxHeld && blocks_.size() < 64) {
        blocks_.push_back(Block{p, bytes, pinned});
        held_ += bytes;
        return;
      }
    }
    if (pinned) cudaFreeHost(p);
    else free(p);
  }
 private:
  struct Block { void* p; size_t bytes; bool pinned; };
  static constexpr size_t kMaxHeld = (size_t)2 << 30;
  std::mutex mu_;
  std::vector<Block> blocks_;
  size_t held_ = 0;
};
// uninitialised array of trivially copyable T on a cached host block
template <class T>
class HostArr {
 public:
  HostArr() = default;
  HostArr(const HostArr&) = delete;
  HostArr& operator=(const HostArr&) = delete;
  ~HostArr() { HostBlockCache::get().give(p_, cap_bytes_, pinned_); }
  bool resize_uninit(size_t n, bool pinned = false) {
    if (n * sizeof(T) > cap_bytes_ || (pinned && !pinned_)) {
      HostBlockCache::get().give(p_, cap_bytes_, pinned_);
      pinned_ = pinned;
      p_ = (T*)HostBlockCache::get().take(n * sizeof(T), &cap_bytes_, pinned);
      if (!p_) { cap_bytes_ = 0; n_ = 0; return false; }
    }
    n_ = n;
    return true;
  }
  size_t size() const { return n_; }
  T* data() { return p_; }
  const T* data() const { return p_; }
  T& operator[](size_t i) { return p_[i]; }
  const T& operator[](size_t i) const { return p_[i]; }
 private:
  T* p_ = nullptr;
  size_t n_ = 0, cap_bytes_ = 0;
  bool pinned_ = false;
};
struct emrf_forest {
  int num_trees = 0, num_classes = 0, num_features = 0;
  std::vector<int64_t> tree_off;
  HostArr<FlatNode> nodes;
  std::vector<double> stats;  // nodes*C (filled lazily from counts for forests made by emrf_fit)
  HostArr<uint32_t> counts;   // nodes*C integer class counts as trained (left uninitialised until filled)
  size_t counts_len = 0;
  std::mutex stats_mu;
  std::vector<int32_t> group_sizes, trace;
  int f_sub = 0;
  // device copy for predict (built once)
  std::mutex dev_mu;
  int dev_device = -1;
  DevBuf d_nodes, d_roots, d_leaf_prob;
  // binned-walk layout (predict v2): 8-byte nodes with adjacent children, per-feature threshold tables
  bool walk_ok = false;
  int walk_ts = 0;            // threshold table stride
  DevBuf d_wnodes, d_wroots, d_pthr, d_pns;
  DevBuf d_wblocks, d_broots; // 32-byte two-level tree blocks of the compact walk and each tree's root block
  DevBuf d_leaf_rec;          // compact leaf records (counts + total + 1/total), see predict_walk_compact_kernel
  int leaf_count_bytes = 0;   // 0: not available (non-integral or huge class counts), 2 or 4
};

static inline void node_counts(const HNode& n, int C, uint32_t* out) {
  if (n.cnt_b) for (int c = 0; c < C; c++) out[c] = n.cnt_a[c] - n.cnt_b[c];
  else memcpy(out, n.cnt_a, (size_t)C * 4);
}

// LearningNode.toNode(prune = true) (ml/tree/Node.scala), two passes per tree so that trees can be
// exported in parallel into pre-sized arrays:
//   pass A (post-order): does the node come out as a leaf (own leaf, or both children leaves with
//                        the same prediction), its prediction, and the size of its exported subtree
//   pass B (pre-order):  fill the flat arrays
struct ExportInfo { int32_t size; int32_t pred; bool leaf; };
static void export_pass_a(const HTree& t, int ni, int C, std::vector<ExportInfo>& info) {
  const HNode& n = t.pool[ni];
  ExportInfo& e = info[ni];
  if (n.left >= 0) {
    export_pass_a(t, n.left, C, info);
    export_pass_a(t, n.right, C, info);
    const ExportInfo &l = info[n.left], &r = info[n.right];
    if (l.leaf && r.leaf && l.pred == r.pred) { e.leaf = true; e.pred = l.pred; e.size = 1; }
    else { e.leaf = false; e.pred = n.pred; e.size = 1 + l.size + r.size; }
  } else {
    e.leaf = true; e.pred = n.pred; e.size = 1;
  }
  (void)C;
}
static void export_pass_b(const HTree& t, int ni, int C, const std::vector<ExportInfo>& info, int32_t pos, FlatNode* nodes_out,
                          uint32_t* counts_out) {   // both outputs start at the tree's first node
  const HNode& n = t.pool[ni];
  const ExportInfo& e = info[ni];
  FlatNode& f = nodes_out[pos];
  node_counts(n, C, counts_out + (size_t)pos * C);
  f.id = n.id;
  f.prediction = (double)e.pred;
  if (e.leaf) {
    f.left = -1; f.right = -1; f.feature = -1; f.threshold = 0.0; f.gain = 0.0;
    // own leaf with invalid stats → -1.0 (old mllib behaviour kept by Spark); merged leaf → parent's impurity
    f.impurity = (n.left < 0 && !n.valid) ? -1.0 : n.impurity;
  } else {
    const int32_t lpos = pos + 1, rpos = pos + 1 + info[n.left].size;
    f.left = lpos; f.right = rpos; f.feature = n.feature; f.threshold = n.threshold; f.gain = n.gain;
    f.impurity = n.impurity;
    export_pass_b(t, n.left, C, info, lpos, nodes_out, counts_out);
    export_pass_b(t, n.right, C, info, rpos, nodes_out, counts_out);
  }
}
static bool export_forest(const std::vector<HTree>& trees, int C, emrf_forest& fo, int n_threads) {
  const int T = (int)trees.size();
  std::vector<std::vector<ExportInfo>> info(T);
#pragma omp parallel for schedule(dynamic) num_threads(n_threads)
  for (int t = 0; t < T; t++) {
    info[t].resize(trees[t].pool.size());
    export_pass_a(trees[t], 0, C, info[t]);
  }
  fo.tree_off.assign(T + 1, 0);
  for (int t = 0; t < T; t++) fo.tree_off[t + 1] = fo.tree_off[t] + info[t][0].size;
  fo.counts_len = (size_t)fo.tree_off[T] * C;
  if (!fo.nodes.resize_uninit((size_t)fo.tree_off[T]) || !fo.counts.resize_uninit(fo.counts_len)) return false;
#pragma omp parallel for schedule(dynamic) num_threads(n_threads)
  for (int t = 0; t < T; t++)
    export_pass_b(trees[t], 0, C, info[t], 0, fo.nodes.data() + fo.tree_off[t], fo.counts.data() + (size_t)fo.tree_off[t] * C);
  return true;
}
static void ensure_double_stats(emrf_forest* f) {
  std::lock_guard<std::mutex> lk(f->stats_mu);
  if (!f->stats.empty() || f->counts_len == 0) return;
  f->stats.resize(f->counts_len);
  for (size_t i = 0; i < f->counts_len; i++) f->stats[i] = (double)f->counts[i];
}

// The same with the work divided: every rank holds every tree (the scheduler is replicated), so with many ranks sharing
// the host's cores each exports only trees t % world == rank and an all-gather hands the blocks round.  Rank r's block
// is [exported size of each of its trees][its trees' nodes][their class counts], the trees laid out at the prefix sums
// of their POOL sizes (known to every rank; an exported tree is never larger than its pool).  The model arrays are
// page-locked and filled tree by tree straight from the gathered blocks.  Collective: every rank calls it.
static int export_forest_sharded(emrf_ctx* c, const std::vector<HTree>& trees, int C, emrf_forest& fo, cudaStream_t st) {
  const int T = (int)trees.size(), W = c->world;
  const int per = (T + W - 1) / W;                         // trees per rank, at most
  std::vector<size_t> at(T, 0), share(W, 0);               // node offset of tree t inside its owner's block; nodes per block
  for (int t = 0; t < T; t++) { at[t] = share[t % W]; share[t % W] += trees[t].pool.size(); }
  size_t cap = 0;
  for (int r = 0; r < W; r++) cap = std::max(cap, share[r]);
  const size_t off_nodes = align_up((size_t)per * 4, 16);
  const size_t off_counts = align_up(off_nodes + cap * sizeof(FlatNode), 16);
  const size_t block = align_up(off_counts + cap * (size_t)C * 4, 16);
  CU(c->h_export.ensure(block));
  CU(c->d_export.ensure(block * (size_t)W));
  char* hb = (char*)c->h_export.p;
  uint32_t* own_sizes = (uint32_t*)hb;
#pragma omp parallel for schedule(dynamic) num_threads(c->host_threads)
  for (int t = c->rank; t < T; t += W) {
    std::vector<ExportInfo> info(trees[t].pool.size());
    export_pass_a(trees[t], 0, C, info);
    own_sizes[t / W] = (uint32_t)info[0].size;
    export_pass_b(trees[t], 0, C, info, 0, (FlatNode*)(hb + off_nodes) + at[t], (uint32_t*)(hb + off_counts) + at[t] * (size_t)C);
  }
  char* db = (char*)c->d_export.p;
  CU(cudaMemcpyAsync(db + (size_t)c->rank * block, hb, block, cudaMemcpyHostToDevice, st));
  TRY(nccl_all_gather_bytes(c, db, block, st));
  std::vector<uint32_t> sizes((size_t)W * per, 0);
  for (int r = 0; r < W; r++)
    CU(cudaMemcpyAsync(sizes.data() + (size_t)r * per, db + (size_t)r * block, (size_t)per * 4, cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  fo.tree_off.assign(T + 1, 0);
  for (int t = 0; t < T; t++) {
    const uint32_t sz = sizes[(size_t)(t % W) * per + t / W];
    if (sz < 1 || sz > trees[t].pool.size()) return fail(EMRF_ERR_INTERNAL, "rank %d exported tree %d with %u nodes", t % W, t, sz);
    fo.tree_off[t + 1] = fo.tree_off[t] + sz;
  }
  fo.counts_len = (size_t)fo.tree_off[T] * C;
  if (!fo.nodes.resize_uninit((size_t)fo.tree_off[T], true) || !fo.counts.resize_uninit(fo.counts_len, true))
    return fail(EMRF_ERR_INTERNAL, "out of page-locked host memory for the model");
  for (int t = 0; t < T; t++) {
    const char* src = db + (size_t)(t % W) * block;
    const size_t n = (size_t)(fo.tree_off[t + 1] - fo.tree_off[t]);
    CU(cudaMemcpyAsync(fo.nodes.data() + fo.tree_off[t], src + off_nodes + at[t] * sizeof(FlatNode), n * sizeof(FlatNode),
                       cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(fo.counts.data() + (size_t)fo.tree_off[t] * C, src + off_counts + at[t] * (size_t)C * 4, n * (size_t)C * 4,
                       cudaMemcpyDeviceToHost, st));
  }
  c->tm.comm_bytes += (int64_t)block * W;
  c->tm.d2h_bytes += (int64_t)fo.tree_off[T] * (int64_t)(sizeof(FlatNode) + (size_t)C * 4);
  c->tm.h2d_bytes += (int64_t)block;
  return EMRF_OK;   // the copies complete with the fit's final stream synchronisation
}

// ---------------------------------------------------------------------------------------------
// emrf_fit
// ---------------------------------------------------------------------------------------------
// growable array of trivially copyable T that never value-initialises (the scheduler's per-group arrays are
// overwritten in full every group; zero-filling them first is a measurable part of the host's critical path)
template <class T>
class PodVec {
 public:
  PodVec() = default;
  PodVec(const PodVec&) = delete;
  PodVec& operator=(const PodVec&) = delete;
  ~PodVec() { free(p_); }
  size_t size() const { return n_; }
  bool empty() const { return n_ == 0; }
  void clear() { n_ = 0; }
  void resize(size_t n) {
    if (n > cap_) {
      const size_t cap = std::max(n, cap_ * 2);
      T* q = (T*)realloc((void*)p_, cap * sizeof(T));
      if (!q) throw std::bad_alloc();
      p_ = q; cap_ = cap;
    }
    n_ = n;
  }
  void reserve(size_t n) { const size_t keep = n_; resize(std::max(n, n_)); n_ = keep; }
  void push_back(const T& v) { resize(n_ + 1); p_[n_ - 1] = v; }
  T* data() { return p_; }
  const T* data() const { return p_; }
  T& operator[](size_t i) { return p_[i]; }
  const T& operator[](size_t i) const { return p_[i]; }
  T* begin() { return p_; }
  T* end() { return p_ + n_; }
  const T* begin() const { return p_; }
  const T* end() const { return p_ + n_; }
 private:
  T* p_ = nullptr;
  size_t n_ = 0, cap_ = 0;
};
// What the scheduler knows about a node that waits on the stack or sits in a group — everything selection and
// staging need, so that neither touches the tree pools (their nodes are filled in later, off the critical path).
struct StackEnt {
  int32_t tree, pool;   // pool index of the node inside its tree
  int32_t id, level;    // Spark node id (1 = root) and its depth
  uint32_t slot;        // index of the node's entry segment in the device segment table
  uint32_t ub;          // upper bound on the segment's length (exact for roots and when rows are sharded)
  uint32_t weight;      // sum of the node's class counts (roots: an upper bound)
};
struct GroupEntry : StackEnt {
  int feat_at;  // index into the group's feature array
};

static int fit_impl(emrf_ctx* c, const emrf_params_t& P_in, const double* features_hd, const double* labels_hd, int64_t N,
                    int F, const int64_t* part_off, int n_parts, emrf_forest_t** out, const ShardSpec* sh = nullptr) {
  emrf_params_t P = P_in;
  emrf_timings_t& tm = c->tm;
  memset(&tm, 0, sizeof(tm));
  c->events.reset();
  cudaStream_t st = c->stream;
  CU(cudaSetDevice(c->device));
  if (N >= (1ll << 32)) return fail(EMRF_ERR_LIMIT, "n_rows must be < 2^32");
  if (F >= (1 << 24)) return fail(EMRF_ERR_LIMIT, "n_features must be < 2^24");
  // row_sharded: this rank holds a slice of the rows and every tree; node histograms are all-reduced.
  // otherwise (world > 1): tree-sharded — every rank holds all rows and builds trees t % world == rank.
  const bool row_sharded = (sh != nullptr);
  if (c->world > 1 && P.shard_mode == 1 && !row_sharded)
    return fail(EMRF_ERR_ARG, "shard_mode 1 (row-sharded) needs emrf_fit_sharded: each rank passes its own rows");
  auto mine = [&](int tree) { return row_sharded || (tree % c->world == c->rank); };

  cudaEvent_t ev_start = c->events.get(), ev_in = c->events.get(), ev_fs = c->events.get(), ev_bin = c->events.get(),
              ev_bag = c->events.get(), ev_end = c->events.get();
  CU(cudaEventRecord(ev_start, st));

  // ---- bootstrap weights start now, on the second stream (they need only the partition layout)
  DevBuf& d_w = c->p_w;
  BagJob bag_job;
  cudaEvent_t ev_b0 = c->events.get(), ev_b1 = c->events.get();
  CU(d_w.ensure_exact((size_t)N * P.num_trees));
  CU(cudaEventRecord(ev_b0, c->stream2));
  {
    emrf_params_t Pb = P;   // partition p of this rank is global partition first_part + p: seed + p_global + 1
    if (row_sharded) Pb.seed = P.seed + sh->first_part;
    TRY(bag_launch(c, c->stream2, Pb, N, part_off, n_parts, d_w.as<uint8_t>(), bag_job));
  }
  CU(cudaEventRecord(ev_b1, c->stream2));

  // ---- inputs
  DevBuf &own_X = c->p_X, &own_y = c->p_y;
  const void* dXv = nullptr;
  const void* dyv = nullptr;
  bool copied = false;
  // A host-resident feature matrix crosses PCIe in chunks on its own stream: findSplits works from the sampled
  // rows (gathered on the host) and the TreePoint conversion follows the copy chunk by chunk, so that by the
  // time the last chunk lands only its own binning is left.  Bagging runs under all of it on stream2.
  struct XChunk { int64_t r0, rows; cudaEvent_t ev; };
  std::vector<XChunk> x_chunks;
  HostX hostx{nullptr, nullptr};
  TRY(stage_input(c, labels_hd, (size_t)N * 8, own_y, &dyv, nullptr));   // first: copies leave in issue order
  if (!is_device_ptr(features_hd)) {
    CU(own_X.ensure_exact((size_t)N * F * 8));
    cudaEvent_t ev_x0 = c->events.get();
    CU(cudaEventRecord(ev_x0, st));                       // the pool buffer may still be read by work queued on st
    CU(cudaStreamWaitEvent(c->stream3, ev_x0, 0));
    const int64_t per = (int64_t)align_up((size_t)((N + 11) / 12), 128);
    for (int64_t r0 = 0; r0 < N; r0 += per) {
      const int64_t rows = std::min(per, N - r0);
      CU(cudaMemcpyAsync((char*)own_X.p + (size_t)r0 * F * 8, (const char*)features_hd + (size_t)r0 * F * 8, (size_t)rows * F * 8,
                         cudaMemcpyHostToDevice, c->stream3));
      cudaEvent_t e = c->events.get();
      CU(cudaEventRecord(e, c->stream3));
      x_chunks.push_back({r0, rows, e});
    }
    tm.h2d_bytes += (int64_t)N * F * 8;
    dXv = own_X.p;
    copied = true;
    hostx.h = features_hd;
    hostx.ready = x_chunks.back().ev;
  } else {
    dXv = features_hd;
  }
  const double* dX = (const double*)dXv;
  const double* dy = (const double*)dyv;
  CU(cudaEventRecord(ev_in, st));

  // ---- labels (Predictor.extractLabeledPoints / getNumClasses)
  DevBuf& d_labels = c->p_labels;
  DevBuf d_flags;
  CU(d_labels.ensure_exact((size_t)N));
  CU(d_flags.alloc(8));
  CU(cudaMemsetAsync(d_flags.p, 0, 8, st));
  kern::labels_kernel<<<(unsigned)((N + 255) / 256), 256, 0, st>>>(dy, N, P.num_classes, d_labels.as<uint8_t>(), d_flags.as<int>());
  tm.kernel_launches++;
  int flags[2] = {0, 0};
  if (row_sharded) TRY(nccl_allreduce(c, d_flags.p, 2, true, nullptr));
  CU(cudaMemcpyAsync(flags, d_flags.p, 8, cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  if (flags[0]) return fail(EMRF_ERR_ARG, "Classifier given dataset with invalid label: labels must be integers in [0, numClasses)");
  if (P.num_classes == 0) {
    P.num_classes = flags[1] + 1;
    if (P.num_classes > 100)
      return fail(EMRF_ERR_ARG, "Classifier inferred %d classes from label values, above the 100 allowed to be inferred; set num_classes",
                  P.num_classes);
  }
  const int C = P.num_classes;
  const int T = P.num_trees;

  // ---- findSplits
  DevBuf d_thr, d_ns;
  std::vector<double> h_thr;
  std::vector<int> h_ns;
  std::vector<DevBuf> parked;   // temporaries released once nothing overlaps any more
  TRY(run_find_splits(c, P, dX, N, F, part_off, n_parts, d_thr, d_ns, h_thr, h_ns, sh, &parked, copied ? &hostx : nullptr));
  CU(cudaEventRecord(ev_fs, st));
  const int TS = P.max_bins;
  std::vector<int> num_bins(F);
  int B = 1;
  for (int f = 0; f < F; f++) { num_bins[f] = h_ns[f] + 1; B = std::max(B, num_bins[f]); }
  DevBuf d_num_bins;
  CU(d_num_bins.alloc((size_t)F * 4));

  // ---- TreePoint matrices
  const int Fp = row_pitch(F);
  const int64_t Np = (int64_t)align_up((size_t)N, 128);
  DevBuf &d_rm = c->p_rm, &d_cm = c->p_cm;
  CU(d_rm.ensure_exact((size_t)Np * Fp));
  CU(d_cm.ensure_exact((size_t)F * Np));
  if (env_flag("EMRF_DEBUG_TIMERS")) fprintf(stderr, "[emrf] binning: %d features, longest threshold table %d, table stride %d\n", F, B - 1, TS);
  if (x_chunks.empty()) {
    TRY(run_bin(c, dX, N, F, d_thr.as<double>(), d_ns.as<int>(), TS, B - 1, Fp, Np, d_rm.as<uint8_t>(), d_cm.as<uint8_t>()));
  } else {
    for (const XChunk& xc : x_chunks) {   // chunk starts are multiples of 128 rows (the binning tile)
      CU(cudaStreamWaitEvent(st, xc.ev, 0));
      TRY(run_bin(c, dX + (size_t)xc.r0 * F, xc.rows, F, d_thr.as<double>(), d_ns.as<int>(), TS, B - 1, Fp, Np,
                  d_rm.as<uint8_t>() + (size_t)xc.r0 * Fp, d_cm.as<uint8_t>() + xc.r0));
    }
  }
  CU(cudaEventRecord(ev_bin, st));
  CU(cudaMemcpyAsync(d_num_bins.p, num_bins.data(), (size_t)F * 4, cudaMemcpyHostToDevice, st));   // behind the matrix chunks

  // ---- metadata (DecisionTreeMetadata.buildMetadata) and the driver RNG (RandomForest.run)
  const int f_sub = hl::features_per_node(P, F);
  if (f_sub < 1 || f_sub > F) return fail(EMRF_ERR_ARG, "featureSubsetStrategy selects %d of %d features", f_sub, F);
  const bool subsampling = (f_sub != F);
  hl::JavaUtilRandom rng(P.seed);
  // ---- feature-subset table: draw k of the driver RNG → (subset, sum of numBins), tabulated on the GPU
  // in chunks (kernels.cuh reservoir_table_kernel).  selectNodesToSplit then only counts draws.
  const int64_t draws_bound = 2 * (int64_t)T * ((int64_t)1 << std::min(P.max_depth, 24)) + 1024;
  const int table_cap = (int)std::min<int64_t>(draws_bound, (int64_t)1 << 20);
  DevBuf &d_seeds = c->d_seeds, &d_table = c->d_table, &d_tsum = c->d_tsum;
  PinnedBuf &h_seeds = c->h_seeds, &h_table = c->h_table, &h_tsum = c->h_tsum;
  int64_t table_base = -1;        // first draw index held by the table (-1: nothing yet)
  int64_t draw_idx = 0;           // draws consumed so far (accepted nodes + rejected ones)
  if (subsampling) {
    CU(d_seeds.ensure((size_t)table_cap * 8));
    CU(d_table.ensure((size_t)table_cap * f_sub * 4));
    CU(d_tsum.ensure((size_t)table_cap * 4));
    CU(h_seeds.ensure((size_t)table_cap * 8));
    CU(h_table.ensure((size_t)table_cap * f_sub * 4));
    CU(h_tsum.ensure((size_t)table_cap * 4));
  }
  auto table_fill = [&](int64_t base, bool sync = true) -> int {   // rng is positioned at draw `base`
    int64_t* hs = h_seeds.as<int64_t>();
    for (int k = 0; k < table_cap; k++) hs[k] = rng.next_long();
    CU(cudaMemcpyAsync(d_seeds.p, hs, (size_t)table_cap * 8, cudaMemcpyHostToDevice, st));
    kern::reservoir_table_kernel<<<(table_cap + 127) / 128, 128, 0, st>>>(d_seeds.as<int64_t>(), table_cap, F, f_sub,
                                                                         d_num_bins.as<int32_t>(), d_table.as<int32_t>(),
                                                                         d_tsum.as<int32_t>());
    tm.kernel_launches++;
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(h_table.p, d_table.p, (size_t)table_cap * f_sub * 4, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(h_tsum.p, d_tsum.p, (size_t)table_cap * 4, cudaMemcpyDeviceToHost, st));
    if (sync) CU(cudaStreamSynchronize(st));
    table_base = base;
    return EMRF_OK;
  };
  // the first chunk is tabulated now, underneath bagging; the host waits for it with the entry lists
  if (subsampling) TRY(table_fill(0, false));


  // ---- BaggedPoint weights → per-tree entry lists
  // tree-sharded multi-GPU: every rank schedules all trees, but only builds lists (and
  // histograms) for trees t % world == rank.
  DevBuf* d_ent[2] = {&c->p_ent0, &c->p_ent1};
  int bag_maxw = 255;
  std::vector<uint64_t> tree_base(T + 1, 0), tree_total(T, 0);
  {
    // on the bagging stream, right behind the weights: with host-resident features (or a sharded findSplits) the
    // lists are ready before the last rows are binned; the main stream joins below
    cudaStream_t sl = c->stream2;
    int maxw = 0;
    TRY(bag_finish(c, sl, P, bag_job, &maxw));
    bag_maxw = maxw;
    const int n_tiles = (int)((N + kern::kListTile - 1) / kern::kListTile);
    DevBuf& d_counts = c->p_counts;
    DevBuf d_totals, d_base;
    CU(d_counts.ensure_exact((size_t)n_tiles * T * 4));
    CU(d_totals.alloc((size_t)T * 8));
    CU(d_base.alloc((size_t)T * 8));
    const int bt = std::min(256, (int)align_up(T, 32));
    // slab of rows staged in shared memory by the tiled kernels (0: T too large, thread-per-tree kernels)
    int slab_rows = (int)std::min<int64_t>(512, (48 * 1024 / T) / 32 * 32);
    if (env_flag("EMRF_LIST_SIMPLE")) slab_rows = 0;
    if (slab_rows >= 32) {
      kern::list_count_tiled_kernel<<<n_tiles, 256, (size_t)slab_rows * T, sl>>>(d_w.as<uint8_t>(), N, T, slab_rows, d_counts.as<uint32_t>());
      kern::list_scan_warp_kernel<<<(T * 32 + 127) / 128, 128, 0, sl>>>(d_counts.as<uint32_t>(), n_tiles, T, d_totals.as<uint64_t>());
    } else {
      kern::list_count_kernel<<<n_tiles, bt, 0, sl>>>(d_w.as<uint8_t>(), N, T, d_counts.as<uint32_t>());
      kern::list_scan_kernel<<<(T + 127) / 128, 128, 0, sl>>>(d_counts.as<uint32_t>(), n_tiles, T, d_totals.as<uint64_t>());
    }
    tm.kernel_launches += 2;
    CU(cudaMemcpyAsync(tree_total.data(), d_totals.p, (size_t)T * 8, cudaMemcpyDeviceToHost, sl));
    CU(cudaStreamSynchronize(sl));
    for (int t = 0; t < T; t++) tree_base[t + 1] = tree_base[t] + tree_total[t];
    CU(d_ent[0]->ensure((size_t)tree_base[T] * 8));
    CU(d_ent[1]->ensure((size_t)tree_base[T] * 8));
    CU(cudaMemcpyAsync(d_base.p, tree_base.data(), (size_t)T * 8, cudaMemcpyHostToDevice, sl));
    if (slab_rows >= 32)
      kern::list_fill_tiled_kernel<<<n_tiles, 256, (size_t)slab_rows * T, sl>>>(d_w.as<uint8_t>(), d_labels.as<uint8_t>(), N, T, slab_rows,
                                                                              d_counts.as<uint32_t>(), d_base.as<uint64_t>(),
                                                                              d_ent[0]->as<uint64_t>());
    else
      kern::list_fill_kernel<<<n_tiles, bt, 0, sl>>>(d_w.as<uint8_t>(), d_labels.as<uint8_t>(), N, T, d_counts.as<uint32_t>(),
                                                     d_base.as<uint64_t>(), d_ent[0]->as<uint64_t>());
    tm.kernel_launches += 1;
    CU(cudaGetLastError());
    cudaEvent_t ev_lists = c->events.get();
    CU(cudaEventRecord(ev_lists, sl));
    CU(cudaStreamWaitEvent(st, ev_lists, 0));
    CU(cudaEventRecord(ev_bag, st));
    CU(cudaStreamSynchronize(sl));  // the small tables die here; d_w stays in the ctx pool
    CU(cudaStreamSynchronize(st));
    parked.clear();
  }

  // ---- metadata (DecisionTreeMetadata.buildMetadata)
  int row_ch, k_chunk, n_chunks;
  size_t hist_region, hist_smem;
  bool hist_tma = false;
  TRY(hist_geometry(c, Fp, f_sub, B, C, &row_ch, &k_chunk, &n_chunks, &hist_region, &hist_smem, &hist_tma));
  CU(hist_launch(row_ch, dim3(1), hist_smem, st, kern::HistArgs(), true));
  const size_t hist_bytes = kern::hist_smem_bytes(k_chunk, B, C);   // the stand-alone scan kernels (rows sharded)
  CUtensorMap rows_tmap;
  memset(&rows_tmap, 0, sizeof(rows_tmap));
  if (hist_tma) TRY(make_rows_tmap(&rows_tmap, d_rm.as<uint8_t>(), Np, Fp));
  // 16-bit counters for launches of small single-tile nodes (EMRF_HIST_U32=1 turns them off)
  const size_t hist_smem16 = hist_region + kern::hist_block_bytes(k_chunk, B, C, 2);
  const bool allow_u16 = !row_sharded && !env_flag("EMRF_HIST_U32");
  // ... also for groups with multi-tile nodes (their last tile then scans the 32-bit sums k_chunk / 2 features at a time)
  const bool u16_mixed_ok = allow_u16 && n_chunks == 1 && k_chunk >= 2 && !env_flag("EMRF_NO_U16_MIXED");
  // nodes holding at least 1/EMRF_CM_DIV of the rows gather from the column-major matrix (0: never)
  uint32_t cm_min_len = 0;
  {
    const char* e = getenv("EMRF_CM_DIV");
    const long div = e ? atol(e) : 4;
    if (div > 0) cm_min_len = (uint32_t)std::max<int64_t>(N / div, 1024);
  }
  if (allow_u16) CU(hist_launch(row_ch, dim3(1), hist_smem16, st, kern::HistArgs(), true, true));
  if (row_sharded) {
    CU(cudaFuncSetAttribute(kern::scan_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)hist_bytes));
    CU(cudaFuncSetAttribute(kern::scan_owned_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)hist_bytes));
  }
  const size_t slot_words = (size_t)f_sub * B * C;
  // L2 blocking of the row gathers (shapes whose row-major TreePoint matrix does not fit L2, e.g. configs[2]);
  // opt-in (EMRF_L2_BLOCK=1): measured on 10M x 128, depth 12, 50 trees it shortens the histogram kernel by
  // only 6 % (217 -> 203 ms: the deep levels, where a group spans < 2 trees, have no reuse to win) and the
  // host spends more than that building the per-tile tables
  const size_t rm_bytes = (size_t)Np * Fp;
  const bool l2_blocked = rm_bytes > ((size_t)80 << 20) && env_flag("EMRF_L2_BLOCK");
  const int l2_ranges = l2_blocked ? (int)std::min<size_t>(256, (rm_bytes + ((size_t)24 << 20) - 1) / ((size_t)24 << 20)) : 1;
  // row-sharded, EMRF_PEER_PUSH=1: partial histograms are pushed straight into the owner rank's staging
  // buffer over NVLink (fused histogram + reduce-scatter).  Measured on 8 B200s it is ~4 % SLOWER per fit
  // than local flush + ncclReduceScatter (the remote stores lengthen the histogram kernel by more than the
  // collective costs), so the NCCL path stays the default; see DESIGN.md §5.
  bool push_mode = false;
  if (row_sharded && c->world > 1 && c->world <= 8 && env_flag("EMRF_PEER_PUSH")) {
    int64_t sb = 0;   // fewest bins a node's aggregate can have → most nodes a group can hold
    {
      std::vector<int> nbs(num_bins.begin(), num_bins.end());
      std::sort(nbs.begin(), nbs.end());
      for (int k = 0; k < f_sub; k++) sb += nbs[k];
    }
    const int64_t max_mem_b = (int64_t)P.max_memory_mb * 1024ll * 1024ll;
    const int64_t max_nodes = max_mem_b / std::max<int64_t>((int64_t)C * sb * 8, 1) + 1;
    const int64_t npr_max = (max_nodes + c->world - 1) / c->world + 1;
    const size_t stage_bytes = (size_t)c->world * (size_t)npr_max * slot_words * 4;
    if (stage_bytes <= ((size_t)8 << 30)) {
      TRY(peer_stage_setup(c, stage_bytes));
      CU(cudaFuncSetAttribute(kern::scan_sum_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)hist_bytes));
      push_mode = true;
    }
  }
  // rows sharded over NCCL: the device tiles every group from the exact segment lengths (tile_map_kernel), the host only
  // provisions the launches from bounds — no stream synchronisation per group to read the children's sizes back
  // (EMRF_RS_HOST_SIZES=1: the host reads them and tiles, as on one GPU)
  static const int rs_pieces_want = [] { const char* e = getenv("EMRF_RS_PIECES"); return e ? std::min(8, std::max(1, atoi(e))) : 1; }();
  const bool dev_tiles = row_sharded && c->world > 1 && !push_mode && !l2_blocked && rs_pieces_want == 1 && !env_flag("EMRF_RS_HOST_SIZES");
  std::vector<uint64_t> tree_acc(dev_tiles ? T : 0, 0);
  std::vector<uint32_t> bound_ub;
  std::vector<int32_t> bound_tree;
  // rows sharded over NCCL: light nodes cross the links as packed 16-bit counters (EMRF_NO_RS_PACK=1: all 32-bit)
  const bool rs_pack_ok = row_sharded && c->world > 1 && !push_mode && (B * C) % 2 == 0 && !env_flag("EMRF_NO_RS_PACK");

  // ---- trees, stack, driver RNG (RandomForest.run)
  // node pools live in the context: a fit reuses the (already faulted-in) memory of the previous one
  std::vector<HTree>& trees = c->tree_pools;
  trees.resize(T);
  const size_t pool_reserve = (size_t)1 << std::min(P.max_depth + 1, 12);
  for (int t = 0; t < T; t++) {
    trees[t].pool.clear();
    trees[t].pool.reserve(pool_reserve);
    trees[t].pool.emplace_back();
  }
  // device segment table: roots now, children written by the partition kernels
  size_t seg_cap = std::max<size_t>((size_t)1 << 18, (size_t)4 * T);
  uint32_t next_slot = (uint32_t)T;
  CU(c->p_segtab.ensure(seg_cap * sizeof(kern::Seg)));
  seg_cap = c->p_segtab.bytes / sizeof(kern::Seg);
  {
    std::vector<kern::Seg> roots(T);
    for (int t = 0; t < T; t++) { roots[t].start = tree_base[t]; roots[t].len = (uint32_t)tree_total[t]; roots[t].buf = 0; }
    CU(cudaMemcpyAsync(c->p_segtab.p, roots.data(), (size_t)T * sizeof(kern::Seg), cudaMemcpyHostToDevice, st));
    CU(cudaStreamSynchronize(st));
  }
  std::vector<uint32_t> hist_slots;   // slot of every node this rank built a histogram for (exact stats at the end)
  hist_slots.reserve(c->last_nodes_split + 1024);   // sized like the previous fit's: no regrowth inside the group loop
  PodVec<StackEnt> stack;
  stack.reserve(1024);
  for (int t = 0; t < T; t++) {
    StackEnt r;
    r.tree = t; r.pool = 0; r.id = 1; r.level = 0;
    r.slot = (uint32_t)t;
    r.ub = (uint32_t)tree_total[t];
    // upper bound; rows sharded: the local count says nothing about the other ranks' rows (and the packed transport
    // must class a node the same way on every rank), so roots count as heavy
    r.weight = row_sharded ? 0xffffffffu : (uint32_t)std::min<uint64_t>(tree_total[t] * (uint64_t)std::max(bag_maxw, 1), 0xffffffffull);
    stack.push_back(r);
  }
  const int64_t max_mem = (int64_t)P.max_memory_mb * 1024ll * 1024ll;
  int64_t all_feat_bins = 0;
  for (int f = 0; f < F; f++) all_feat_bins += num_bins[f];

  emrf_forest* fo = new emrf_forest();
  std::unique_ptr<emrf_forest> fo_guard(fo);
  fo->num_trees = T; fo->num_classes = C; fo->num_features = F; fo->f_sub = f_sub;
  fo->trace.reserve(c->last_trace_len + 4096);
  fo->group_sizes.reserve(256);

  // A group as selectNodesToSplit leaves it: popped nodes in pop order, their feature subsets, the
  // trees in first-seen order (= mutable.HashMap insertion order) and each tree's nodes.
  struct GroupSel {
    PodVec<GroupEntry> group;
    PodVec<int32_t> feat;
    PodVec<int32_t> tree_of;     // group[i].tree, compact (the runs below are found in it)
    std::vector<int> trees_seen;
    struct Run { int tree, start, len; };
    std::vector<Run> runs;       // maximal runs of consecutive nodes of one tree, in pop order (a tree's children
                                 // are pushed together, so a group is a handful of runs per tree)
    uint64_t total_ub = 0;       // entry bounds of this rank's nodes, summed; the heaviest of them
    uint32_t max_w = 0;
    void clear() { group.clear(); feat.clear(); tree_of.clear(); trees_seen.clear(); runs.clear(); total_ub = 0; max_w = 0; }
  };
  GroupSel sel_a, sel_b;
  GroupSel* cur = &sel_a;
  GroupSel* nxt = &sel_b;
  std::vector<int> tree_slot(T, -1);
  std::vector<std::pair<cudaEvent_t, cudaEvent_t>> hist_ev, part_ev, comm_ev;
  double host_ms = 0.0, host_sel = 0.0, host_upd = 0.0;
  static const bool verify_gains = env_flag("EMRF_VERIFY_GAINS");
  c->arena_reset();

  // ---- selectNodesToSplit.  The driver RNG hands each popped node one nextLong() (drawn BEFORE the
  // memory check, so the node that does not fit consumes a draw too); the reservoir sampling a seed
  // implies is independent per node, so a batch of candidate nodes is drawn sequentially (cheap LCG
  // steps, state saved after every draw), sampled in parallel, and then admitted sequentially exactly
  // as Spark does; the RNG is rewound to just after the rejected node's draw.  Needs only the stack,
  // so it runs on the host while the GPU is still partitioning the previous group.
  int select_rc = EMRF_OK;
  auto select_group = [&](GroupSel& G) {
    G.clear();
    int64_t mem = 0;
    bool rejected = false;     // `rejected`: a node did not fit (it has spent its draw); `full`: the older loop form stopped first
    bool full = false;
    const int32_t* tab = h_table.as<int32_t>();
    const int32_t* tsum_all = h_tsum.as<int32_t>();
    while (!stack.empty() && !rejected && !full) {
      // featureSubset = reservoirSampleAndCount(Range(0, F), F_sub, rng.nextLong()): the i-th node examined takes
      // draw number draw_idx + i, so a run of stack entries is served by the current window of the table
      if (subsampling && draw_idx >= table_base + table_cap) {
        select_rc = table_fill(table_base + table_cap);
        if (select_rc != EMRF_OK) return;
      }
      const size_t s_top = stack.size();
      const size_t at0 = subsampling ? (size_t)(draw_idx - table_base) : 0;
      const size_t m = subsampling ? std::min<size_t>(s_top, (size_t)(table_base + table_cap - draw_idx)) : s_top;
      // how many of the next m entries fit: a running sum over the table's bin totals (sequential, tiny)
      size_t k = 0;
      for (; k < m; k++) {
        if (!hl::select_nodes_examines_next(P.policy_flags, mem, max_mem)) { full = true; break; }
        const int64_t node_mem = (int64_t)C * (subsampling ? (int64_t)tsum_all[at0 + k] : all_feat_bins) * 8;
        if (mem + node_mem <= max_mem || mem == 0) mem += node_mem;
        else { rejected = true; break; }
      }
      // the admitted nodes, in pop order; the node that did not fit has consumed its draw too
      const size_t g0 = G.group.size();
      G.group.resize(g0 + k);
      G.feat.resize((g0 + k) * (size_t)f_sub);
      G.tree_of.resize(g0 + k);
      const int nt = k >= 1024 ? c->host_threads : 1;
      (void)nt;
      uint64_t tot = 0;
      uint32_t mw = 0;
#pragma omp parallel for num_threads(nt) schedule(static) reduction(+ : tot) reduction(max : mw)
      for (size_t i = 0; i < k; i++) {
        GroupEntry& ge = G.group[g0 + i];
        static_cast<StackEnt&>(ge) = stack[s_top - 1 - i];
        ge.feat_at = (int)((g0 + i) * (size_t)f_sub);
        G.tree_of[g0 + i] = ge.tree;
        if (mine(ge.tree)) { tot += ge.ub; mw = std::max(mw, ge.weight); }
        int32_t* fo_ = G.feat.data() + (g0 + i) * (size_t)f_sub;
        if (subsampling) memcpy(fo_, tab + (at0 + i) * (size_t)f_sub, (size_t)f_sub * 4);
        else for (int q = 0; q < F; q++) fo_[q] = q;
      }
      G.total_ub += tot;
      G.max_w = std::max(G.max_w, mw);
      stack.resize(s_top - k);
      if (subsampling) draw_idx += (int64_t)k + (rejected ? 1 : 0);
    }
    // trees in first-seen order (= mutable.HashMap insertion order) and the runs of each tree's nodes in pop order
    const int n = (int)G.group.size();
    const int32_t* tof = G.tree_of.data();
    for (int i = 0; i < n;) {
      const int t = tof[i];
      int j = i + 1;
      while (j < n && tof[j] == t) j++;
      G.runs.push_back({t, i, j - i});
      if (tree_slot[t] < 0) { tree_slot[t] = (int)G.trees_seen.size(); G.trees_seen.push_back(t); }
      i = j;
    }
    for (int t : G.trees_seen) tree_slot[t] = -1;   // slots are recomputed by whoever consumes the group
  };

  {
    const double t0 = now_ms();
    select_group(*cur);
    const double d_ = now_ms() - t0; host_ms += d_; host_sel += d_;
  }
  TRY(select_rc);
  // Per group the GPU runs histogram+scan, decide_kernel and the partition kernels back to back; the host reads
  // the records and decision words on the second stream, builds the tree nodes and children under the
  // partition kernels, and must only have the next group's launch ready by the time they end.
  if (n_chunks > 255) return fail(EMRF_ERR_LIMIT, "a node's histogram needs %d feature chunks (at most 255)", n_chunks);
  struct ChildPos { int l, r; uint32_t parent_ub; };   // rows sharded: where a split node's children went on the stack
  std::vector<ChildPos> child_pos;
  struct Upd { uint32_t dec, slot_l, split_idx; int32_t left_pool; };
  std::vector<Upd> upd_vec;
  std::vector<int> pool_next(T, 0);
  PodVec<uint32_t> nt_tmp, tb_tmp, push_off;   // per-group scratch of the staging and push passes
  PodVec<int32_t> slot_tmp;
  std::vector<uint32_t> own_nodes, pack_w;     // rows sharded, packed transport: the group nodes this rank scans
  std::vector<int> pack_owner;
  std::vector<int> run_order;
  bool tickets_zeroed = false, ghist_zeroed = false;   // per fit: another entry point may have used the buffers
  // Node statistics of a processed group (impurities, gain, split, class-count pointers, predictions): nothing in
  // the scheduling depends on them, so they are filled from the full records one group later, under GPU work.
  struct PendingStats {
    GroupSel* sel = nullptr;
    std::vector<Upd> upd;
    const char* h_res = nullptr;
    size_t res_left = 0, res_total = 0;
    int n_nodes = 0, nt = 1;
    cudaEvent_t rec_done = nullptr;
  } pending;
  auto fill_stats = [&](PendingStats& pd) -> int {
    if (!pd.sel) return EMRF_OK;
    CU(cudaEventSynchronize(pd.rec_done));
    const kern::ScanResult* res = (const kern::ScanResult*)pd.h_res;
    const uint32_t* res_leftc = (const uint32_t*)(pd.h_res + pd.res_left);
    const uint32_t* res_totalc = (const uint32_t*)(pd.h_res + pd.res_total);
    const PodVec<GroupEntry>& group = pd.sel->group;
    const PodVec<int32_t>& group_feat = pd.sel->feat;
    const Upd* upd = pd.upd.data();
    int64_t mism = 0, verified = 0;
#pragma omp parallel for num_threads(pd.nt) schedule(static) reduction(+ : mism, verified)
    for (int gi = 0; gi < pd.n_nodes; gi++) {
      const GroupEntry& g = group[gi];
      const uint32_t dw = upd[gi].dec;
      HTree& tr = trees[g.tree];
      const size_t rec = (size_t)gi * n_chunks + (dw & 0xffu);
      const kern::ScanResult& r = res[rec];
      const uint32_t* totc = res_totalc + rec * C;
      const uint32_t* leftc = res_leftc + rec * C;
      const bool valid = (dw & kern::kDecValid) != 0;
      // impurities and gain come from the scan kernel (bit-exact IEEE, Spark's operation order);
      // EMRF_VERIFY_GAINS=1 recomputes them here.
      const double pimp = r.impurity;
      const double gain = valid ? r.gain : kern::kInvalidGain;
      const double limp = valid ? r.left_impurity : 0.0, rimp = valid ? r.right_impurity : 0.0;
      HNode& n = tr.pool[g.pool];
      const bool is_leaf = !(dw & kern::kDecNonLeaf);
      if (verify_gains) {
        verified++;
        if (valid) {
          std::vector<uint32_t> rightc(C);
          uint64_t lc = 0, rc = 0;
          for (int cc = 0; cc < C; cc++) { rightc[cc] = totc[cc] - leftc[cc]; lc += leftc[cc]; rc += rightc[cc]; }
          const double hl_ = hl::impurity_of(P.impurity, leftc, C), hr_ = hl::impurity_of(P.impurity, rightc.data(), C);
          const double hp_ = hl::impurity_of(P.impurity, totc, C);
          const double tc = (double)(lc + rc);
          const double lw = (double)lc / tc, rw = (double)rc / tc;
          const double hg = hp_ - lw * hl_ - rw * hr_;
          if (hg != gain || hl_ != limp || hr_ != rimp || hp_ != pimp) mism++;
        }
        const int level = 31 - __builtin_clz((unsigned)n.id);   // the device's leaf rule against the host's
        if (is_leaf != ((gain <= 0) || (level == P.max_depth))) mism++;
      }
      n.is_leaf = is_leaf; n.gain = gain; n.impurity = pimp; n.valid = valid;
      n.cnt_a = totc; n.cnt_b = nullptr; n.pred = r.pred;
      if (is_leaf) continue;
      const int feat = group_feat[g.feat_at + r.k];
      n.feature = feat; n.split_bin = r.split; n.threshold = h_thr[(size_t)feat * TS + r.split];
      const int li = upd[gi].left_pool;        // the pool was grown when the group was numbered
      n.left = li; n.right = li + 1;
      HNode& l = *new (&tr.pool[li]) HNode();
      HNode& rn = *new (&tr.pool[li + 1]) HNode();
      l.id = 2 * n.id; l.is_leaf = (dw & kern::kDecLLeaf) != 0;
      rn.id = 2 * n.id + 1; rn.is_leaf = (dw & kern::kDecRLeaf) != 0;
      l.impurity = limp; l.cnt_a = leftc; l.pred = r.left_pred;
      rn.impurity = rimp; rn.cnt_a = totc; rn.cnt_b = leftc; rn.pred = r.right_pred;
    }
    tm.gain_mismatches += mism;
    tm.gains_verified += verified;
    pd.sel = nullptr;
    return EMRF_OK;
  };
  double ph[12] = {0};   // EMRF_DEBUG_TIMERS: where the host's time goes, per phase
  double tp = now_ms();
  auto lap = [&](int i) { const double n_ = now_ms(); ph[i] += n_ - tp; tp = n_; };
  while (!cur->group.empty()) {
    double t_host0 = now_ms();
    lap(11);
    PodVec<GroupEntry>& group = cur->group;
    PodVec<int32_t>& group_feat = cur->feat;
    const int n_nodes = (int)group.size();
    const int nt_host = n_nodes >= 1024 ? c->host_threads : 1;   // OpenMP only where a group is worth a fork/join
    (void)nt_host;
    fo->group_sizes.push_back(n_nodes);

    // ---- tiles (tree-sharded multi-GPU: own trees only); sizes are upper bounds
    uint64_t total_entries = cur->total_ub;   // entry bounds of this rank's nodes, summed; the heaviest of them
    uint32_t max_weight = cur->max_w;
    if (row_sharded) {   // the bounds were replaced by exact sizes after the group was selected ...
      total_entries = 0;
      if (dev_tiles) {   // ... or not: the nodes of a tree are disjoint, so they hold at most the tree's entries together
        bound_ub.resize(n_nodes); bound_tree.resize(n_nodes);
        for (int i = 0; i < n_nodes; i++) { bound_ub[i] = group[i].ub; bound_tree[i] = group[i].tree; }
        total_entries = hl::group_entries_bound(bound_ub.data(), bound_tree.data(), n_nodes, tree_total.data(), tree_acc.data());
      } else {
        for (const GroupEntry& g : group) total_entries += g.ub;
      }
    }
    uint32_t tile_len = (uint32_t)align_up((size_t)((total_entries + (uint64_t)c->sm_count * 16 - 1) / ((uint64_t)c->sm_count * 16)), 256);
    // 16-bit launch with multi-tile nodes: a tile's counters cannot pass 65535 if its entries' weights cannot, i.e. if
    // tile_len * (largest bootstrap weight) <= 65535; its nodes' summed histograms live in HBM as 32-bit words.  Used
    // when the group's tiles would be about that short anyway (every group below the top levels); EMRF_HIST_U32=1: off
    const uint32_t tile16 = (65535u / (uint32_t)std::max(bag_maxw, 1)) & ~255u;
    const bool u16_mixed = allow_u16 && u16_mixed_ok && tile_len <= tile16 && tile16 >= 2048;
    tile_len = std::min<uint32_t>(std::max<uint32_t>(tile_len, 8192), 65536);
    if (u16_mixed) tile_len = std::min(tile_len, tile16);
    // TreePoint matrix larger than L2: many small tiles, launched in row-range order (below), so that the
    // CTAs resident at any moment gather rows from the same slice of the matrix and trees share it in L2
    if (l2_blocked) tile_len = 8192;

    // staging layout: [GroupNode n][node_feat n*f_sub][node_info n][tile_node tiles][tile_q tiles], written in place
    // (pinned); node_info = level | "this rank partitions the node" for decide_flags_kernel
    const size_t off_feat = align_up(sizeof(kern::GroupNode) * (size_t)n_nodes, 16);
    const size_t off_info = align_up(off_feat + (size_t)n_nodes * f_sub * 4, 16);
    const size_t off_tiles = align_up(off_info + (size_t)n_nodes * 4, 16);
    {
      const size_t tiles_bound = (size_t)n_nodes + (size_t)(total_entries / tile_len) + 1;
      const size_t bound = align_up(off_tiles + tiles_bound * 4, 16) + tiles_bound * 4 + 16 + (size_t)n_nodes * 4 + 32;
      CU(c->h_group.ensure(bound));
      CU(c->d_group.ensure(bound));
    }
    kern::GroupNode* gn = (kern::GroupNode*)c->h_group.p;
    nt_tmp.resize(n_nodes); tb_tmp.resize(n_nodes); slot_tmp.resize(n_nodes);
    // the partition grid's upper bound: every node that may still split, tiled by its entry bound
    uint64_t ptiles_ub64 = 0;
    {
      uint32_t* info = (uint32_t*)((char*)c->h_group.p + off_info);
      int32_t* feat_out = (int32_t*)((char*)c->h_group.p + off_feat);
#pragma omp parallel for num_threads(nt_host) schedule(static) reduction(+ : ptiles_ub64)
      for (int i = 0; i < n_nodes; i++) {
        const GroupEntry& hn = group[i];
        const bool m = mine(hn.tree);
        kern::GroupNode& d = gn[i];
        d.seg_start = 0;
        d.seg_len = 0;
        d.buf = 0;
        d.seg_slot = hn.slot;
        d.n_tiles = dev_tiles ? 1u : (m ? std::max<uint32_t>(1, (hn.ub + tile_len - 1) / tile_len) : 0);   // dev_tiles: set on the device
        nt_tmp[i] = d.n_tiles;
        const int level = hn.level;
        info[i] = (uint32_t)level | (m ? kern::kInfoMine : 0u);
        if (m && level < P.max_depth && !dev_tiles) ptiles_ub64 += std::max<uint32_t>(1, (hn.ub + kern::kPartTile - 1) / kern::kPartTile);
        memcpy(feat_out + (size_t)i * f_sub, group_feat.data() + group[i].feat_at, (size_t)f_sub * 4);
      }
    }
    // (the sequential part works on compact arrays; the staging records get the results in the parallel pass below)
    uint32_t n_tiles = 0;
    int n_slots = 0;
    // Rows sharded (NCCL path): a group is cut into pieces of consecutive nodes, each with its own block of histogram slots
    // [world][npr] — the reduce-scatter of piece p runs on the communication stream under the histogram kernel of piece
    // p + 1.  Piece q: nodes [pc_n0[q], pc_n0[q+1]), tiles [pc_t0[q], pc_t0[q+1]), slots from pc_s0[q], npr = pc_npr[q].
    constexpr int kMaxPieces = 8;
    bool pack_slots = false;
    int pack_share = 0;
    int n_pieces = 1;
    int pc_n0[kMaxPieces + 1], pc_npr[kMaxPieces], pc_s0[kMaxPieces + 1];
    uint32_t pc_t0[kMaxPieces + 1];
    {
      const bool all_slots = row_sharded && !push_mode;   // every histogram goes to HBM for the reduce-scatter
      if (all_slots) {
        // default 1: measured on 2 B200s, 2 / 4 / 8 pieces make the fit SLOWER (40.0 -> 41.5 / 43.9 / 49.8 ms): the collective
        // and the histogram kernel compete for the same HBM and SMs, and shorter launches fill the GPU worse
        n_pieces = std::max(1, std::min(rs_pieces_want, n_nodes / 128));      // small groups: one piece
      }
      // one piece: nodes whose weight over all ranks is <= 65535 travel as packed 16-bit counters, half the bytes;
      // each rank's share of the buffer is [its heavy nodes, whole slots][its light nodes, half slots]
      pack_slots = all_slots && n_pieces == 1 && rs_pack_ok;
      pc_s0[0] = 0;
      for (int q = 0; q <= n_pieces; q++) pc_n0[q] = (int)((int64_t)n_nodes * q / n_pieces);
      for (int q = 0; q < n_pieces; q++) {
        pc_npr[q] = all_slots ? (pc_n0[q + 1] - pc_n0[q] + c->world - 1) / c->world : 0;
        pc_s0[q + 1] = pc_s0[q] + pc_npr[q] * c->world;
      }
      const uint32_t* ntp = nt_tmp.data();
      uint32_t* tbp = tb_tmp.data();
      int32_t* slp = slot_tmp.data();
      // tile order = launch order: the tiles of multi-tile nodes (full tiles, the longest CTAs) first, then the larger
      // single-tile nodes, then the small ones, so that a launch does not end on a few long CTAs (a group arrives in
      // pop order: deepest, i.e. smallest, nodes first).  Row-range order (l2_blocked) has its own tile map.
      uint32_t cls_tiles[3 * kMaxPieces];
      for (int q = 0; q < 3 * n_pieces; q++) cls_tiles[q] = 0;
      for (int q = 0, i = 0; q < n_pieces; q++) {
        for (; i < pc_n0[q + 1]; i++) {
          slp[i] = all_slots ? pc_s0[q] + (i - pc_n0[q]) : (ntp[i] > 1 ? n_slots++ : -1);
          const int cls = 3 * q + (l2_blocked ? 0 : (ntp[i] > 1 ? 0 : (group[i].ub >= 2048 ? 1 : 2)));
          tbp[i] = (uint32_t)cls;              // class now, first tile below
          cls_tiles[cls] += ntp[i];
        }
      }
      if (all_slots) n_slots = pc_s0[n_pieces];
      if (pack_slots) {
        pack_w.resize(n_nodes); pack_owner.resize(n_nodes);
        for (int i = 0; i < n_nodes; i++) pack_w[i] = group[i].weight;
        pack_share = hl::packed_slot_layout(pack_w.data(), n_nodes, c->world, slp, pack_owner.data());   // half slots per rank
        own_nodes.clear();
        for (int i = 0; i < n_nodes; i++)
          if (pack_owner[i] == c->rank) own_nodes.push_back((uint32_t)i);
        n_slots = (pack_share * c->world + 1) / 2;
      }
      uint32_t cls_off[3 * kMaxPieces];
      for (int q = 0; q < 3 * n_pieces; q++) {
        cls_off[q] = n_tiles;
        if (q % 3 == 0) pc_t0[q / 3] = n_tiles;
        n_tiles += cls_tiles[q];
      }
      pc_t0[n_pieces] = n_tiles;
      for (int i = 0; i < n_nodes; i++) {
        const uint32_t cls = tbp[i];
        tbp[i] = cls_off[cls];
        cls_off[cls] += ntp[i];
      }
    }
    if (dev_tiles) {   // grids from the bounds: sum of ceil(len / tile) <= sum(len) / tile + nodes
      n_tiles = (uint32_t)hl::tiles_bound(total_entries, tile_len, n_nodes);
      pc_t0[n_pieces] = n_tiles;
      ptiles_ub64 = hl::tiles_bound(total_entries, kern::kPartTile, n_nodes);
    }
    const size_t off_tq = align_up(off_tiles + (size_t)n_tiles * 4, 16);
    const size_t off_own = align_up(l2_blocked ? off_tq + (size_t)n_tiles * 4 : off_tiles + (size_t)n_tiles * 4, 16);
    const size_t group_bytes = pack_slots ? off_own + own_nodes.size() * 4
                                          : (l2_blocked ? off_tq + (size_t)n_tiles * 4 : off_tiles + (size_t)n_tiles * 4);
    if (pack_slots && !own_nodes.empty()) memcpy((char*)c->h_group.p + off_own, own_nodes.data(), own_nodes.size() * 4);
    {
      uint32_t* tn = (uint32_t*)((char*)c->h_group.p + off_tiles);
      const uint32_t light_mark = pack_slots ? kern::kNodePacked : 0u;
      if (!l2_blocked) {
#pragma omp parallel for num_threads(nt_host) schedule(static)
        for (int i = 0; i < n_nodes; i++) {
          gn[i].tile_begin = tb_tmp[i];
          gn[i].slot = slot_tmp[i];
          if (group[i].weight <= 65535u) gn[i].buf = light_mark;
          for (uint32_t q = 0; q < gn[i].n_tiles; q++) tn[tb_tmp[i] + q] = (uint32_t)i;
        }
      } else {
        for (int i = 0; i < n_nodes; i++) {
          gn[i].tile_begin = tb_tmp[i]; gn[i].slot = slot_tmp[i];
          if (group[i].weight <= 65535u) gn[i].buf = light_mark;
        }
        // entries of a node are in ascending row order, so tile q of a node with n tiles covers about the
        // q/n-th part of the rows: emit range by range, every node's tiles of that range together
        uint32_t* tq = (uint32_t*)((char*)c->h_group.p + off_tq);
        uint32_t at = 0;
        for (int r = 0; r < l2_ranges; r++)
          for (int i = 0; i < n_nodes; i++) {
            const uint32_t n = gn[i].n_tiles;
            const uint32_t q0 = (uint32_t)((uint64_t)n * r / l2_ranges), q1 = (uint32_t)((uint64_t)n * (r + 1) / l2_ranges);
            for (uint32_t q = q0; q < q1; q++) { tn[at] = (uint32_t)i; tq[at] = q; at++; }
          }
      }
    }
    if (ptiles_ub64 > 0x7fffffffull) return fail(EMRF_ERR_LIMIT, "too many partition tiles in one group");
    const uint32_t ptiles_ub = (uint32_t)ptiles_ub64;
    // record block: [ScanResult r][left r*C][total r*C][decision word n]
    const size_t n_rec = (size_t)n_nodes * n_chunks;
    const size_t res_left = align_up(sizeof(kern::ScanResult) * n_rec, 16);
    const size_t res_total = res_left + n_rec * C * 4;
    const size_t res_dec = align_up(res_total + n_rec * C * 4, 16);
    const size_t res_bytes = align_up(res_dec + (size_t)n_nodes * 4, 256);
    CU(c->d_results.ensure(res_bytes));
    char* h_res = (char*)c->arena_alloc(res_bytes);   // kept until the forest is exported
    if (!h_res) return fail(EMRF_ERR_CUDA, "out of pinned host memory for the split-record arena");
    // single GPU: the HBM histogram slots and tickets a group dirtied are cleared right behind its partition
    // kernels (while the host schedules), not in front of the next histogram launch; zeroed in full once per
    // fit and whenever they grow
    const bool self_clean = (c->world == 1);
    {
      const void* pt = c->d_tickets.p;
      CU(c->d_tickets.ensure(n_rec * 4));
      if (self_clean && (c->d_tickets.p != pt || !tickets_zeroed)) { CU(cudaMemsetAsync(c->d_tickets.p, 0, c->d_tickets.bytes, st)); tickets_zeroed = true; }
    }
    const int nodes_per_rank = row_sharded ? (n_nodes + c->world - 1) / c->world : 0;   // the push path's single piece
    const int n_slots_alloc = n_slots;
    {
      const void* pg = c->d_ghist.p;
      CU(c->d_ghist.ensure((size_t)std::max(n_slots_alloc, 1) * slot_words * 4));
      if (self_clean && (c->d_ghist.p != pg || !ghist_zeroed)) { CU(cudaMemsetAsync(c->d_ghist.p, 0, c->d_ghist.bytes, st)); ghist_zeroed = true; }
    }
    // partition work of this group, written by decide_kernel: [SplitNode n][tile_node ptiles_ub] and
    // [tile_left ptiles_ub][n_left n][limits 2]
    const size_t off_ptn = align_up(sizeof(kern::SplitNode) * (size_t)n_nodes, 16);
    const size_t split_bytes = off_ptn + (size_t)std::max<uint32_t>(ptiles_ub, 4) * 4;
    const size_t off_nleft = align_up((size_t)std::max<uint32_t>(ptiles_ub, 4) * 4, 16);
    const size_t off_limits = off_nleft + (size_t)n_nodes * 8;
    const size_t off_sched = align_up(off_limits + 16, 16);
    const size_t off_ptiles = off_sched + (size_t)n_nodes * 16;
    CU(c->d_split.ensure(split_bytes));
    // one-pass partition: a status word per tile and the tile ticket behind them (cleared per group)
    const size_t off_status = align_up(off_ptiles + (size_t)n_nodes * 4, 16);
    const size_t status_bytes = (size_t)std::max<uint32_t>(ptiles_ub, 1) * 8 + 16;
    CU(c->d_part.ensure(off_status + status_bytes));
    if ((size_t)next_slot + 2 * (size_t)n_nodes > seg_cap) {   // grow the segment table
      CU(cudaStreamSynchronize(st));
      DevBuf bigger;
      CU(bigger.alloc(std::max(seg_cap * 2, (size_t)next_slot + 2 * (size_t)n_nodes) * sizeof(kern::Seg)));
      CU(cudaMemcpyAsync(bigger.p, c->p_segtab.p, (size_t)next_slot * sizeof(kern::Seg), cudaMemcpyDeviceToDevice, st));
      CU(cudaStreamSynchronize(st));
      std::swap(bigger.p, c->p_segtab.p);
      std::swap(bigger.bytes, c->p_segtab.bytes);
      std::swap(bigger.dev, c->p_segtab.dev);
      seg_cap = c->p_segtab.bytes / sizeof(kern::Seg);
    }
    { const double d_ = now_ms() - t_host0; host_ms += d_; host_sel += d_; }
    lap(0);

    // the previous group's records may still be on their way to the host (second stream): nothing of this group
    // may overwrite them before that copy is done
    if (pending.rec_done) CU(cudaStreamWaitEvent(st, pending.rec_done, 0));
    CU(cudaMemcpyAsync(c->d_group.p, c->h_group.p, group_bytes, cudaMemcpyHostToDevice, st));
    tm.h2d_bytes += (int64_t)group_bytes;
    tm.d2h_bytes += (int64_t)res_bytes;
    if (!self_clean) CU(cudaMemsetAsync(c->d_tickets.p, 0, n_rec * 4, st));
    if (c->world > 1) CU(cudaMemsetAsync(c->d_results.p, 0, res_bytes, st));
    if (n_slots_alloc && !self_clean) CU(cudaMemsetAsync(c->d_ghist.p, 0, (size_t)n_slots_alloc * slot_words * 4, st));
    kern::HistArgs A{};
    A.ent0 = d_ent[0]->as<uint64_t>();
    A.ent1 = d_ent[1]->as<uint64_t>();
    A.segtab = c->p_segtab.as<kern::Seg>();
    A.bins_rm = d_rm.as<uint8_t>();
    A.nodes = (const kern::GroupNode*)c->d_group.p;
    A.node_feat = (const int32_t*)((char*)c->d_group.p + off_feat);
    A.tile_node = (const uint32_t*)((char*)c->d_group.p + off_tiles);
    A.tile_q = l2_blocked ? (const uint32_t*)((char*)c->d_group.p + off_tq) : nullptr;
    A.num_bins_all = d_num_bins.as<int32_t>();
    A.ghist = c->d_ghist.as<uint32_t>();
    A.tickets = c->d_tickets.as<uint32_t>();
    A.Fp = Fp; A.f_sub = f_sub; A.B = B; A.C = C; A.k_chunk = k_chunk; A.n_chunks = n_chunks;
    A.tile_len = tile_len;
    if (dev_tiles) {
      uint32_t* d_tile_limit = (uint32_t*)((char*)c->d_group.p + align_up(group_bytes, 16));
      kern::tile_map_kernel<<<1, 1024, 0, st>>>((kern::GroupNode*)c->d_group.p, c->p_segtab.as<kern::Seg>(), n_nodes, tile_len,
                                               (uint32_t*)((char*)c->d_group.p + off_tiles), d_tile_limit, n_tiles);
      tm.kernel_launches++;
      CU(cudaGetLastError());
      A.tile_limit = d_tile_limit;
    }
    // every node in one tile and light enough: 16-bit counters, more CTAs per SM
    const bool u16 = u16_mixed || (allow_u16 && n_slots == 0 && max_weight <= 65535u);
    if (env_flag("EMRF_DEBUG_GROUPS"))
      fprintf(stderr, "[emrf] group %zu: nodes %d tiles %u slots %d max_weight %u tile_len %u u16 %d\n", fo->group_sizes.size(), n_nodes,
              n_tiles, n_slots, max_weight, tile_len, (int)u16);
    A.hist_smem_off = (uint32_t)hist_region;
    A.use_tma = hist_tma ? 1 : 0;
    A.tmap = rows_tmap;
    A.impurity = P.impurity; A.min_instances = P.min_instances_per_node; A.hist_only = row_sharded ? 1 : 0;
    A.push_world = push_mode ? c->world : 0; A.push_rank = c->rank; A.push_npr = nodes_per_rank;
    for (int r = 0; r < 8; r++) A.peer_stage[r] = (uint32_t*)c->peer_stage[r];
    A.slot_half = pack_slots ? 1 : 0;
    A.min_info_gain = P.min_info_gain;
    A.bins_cm = d_cm.as<uint8_t>(); A.Np = Np; A.cm_min_len = cm_min_len;
    A.results = (kern::ScanResult*)c->d_results.p;
    A.left_counts = (uint32_t*)((char*)c->d_results.p + res_left);
    A.total_counts = (uint32_t*)((char*)c->d_results.p + res_total);
    A.tile_base = 0;
    A.node_base = 0;
    // single GPU, one feature chunk: the scanning CTA also takes the decisions (decide_flags_kernel's work)
    const bool fuse_decide = c->world == 1 && n_chunks == 1 && !env_flag("EMRF_NO_FUSED_DECIDE");
    if (fuse_decide) {
      A.node_info = (const uint32_t*)((char*)c->d_group.p + off_info);
      A.out_dec = (uint32_t*)((char*)c->d_results.p + res_dec);
      A.out_sched = (uint4*)((char*)c->d_part.p + off_sched);
      // ... and the host's copy of the scheduling records is written by the kernel itself (zero-copy): the host then
      // waits for the kernel's own event instead of a copy on the second stream behind it
      CU(c->h_sched.ensure((size_t)n_nodes * 16));
      void* hs_dev = nullptr;
      CU(cudaHostGetDevicePointer(&hs_dev, c->h_sched.p, 0));
      A.out_sched_host = (uint4*)hs_dev;
      A.out_ptiles = (uint32_t*)((char*)c->d_part.p + off_ptiles);
      A.max_depth = P.max_depth;
    }
    const bool piped = row_sharded && !push_mode && c->world > 1;   // pieces of the group: histogram kernel || reduce-scatter
    cudaStream_t sc = c->stream_comm;
    // collectives are timed on their own (CUDA events right around each call, on the stream that carries it:
    // emrf_timings_t.comm_ms; with pieces in flight this is device time of the collectives, not exposed time)
    auto comm_begin = [&](cudaStream_t s_) -> cudaEvent_t { cudaEvent_t e = c->events.get(); cudaEventRecord(e, s_); return e; };
    auto comm_end = [&](cudaEvent_t e0, int64_t bytes, cudaStream_t s_) { cudaEvent_t e1 = c->events.get(); cudaEventRecord(e1, s_); comm_ev.emplace_back(e0, e1); tm.comm_bytes += bytes; };
    {
      cudaEvent_t e0 = c->events.get(), e1 = c->events.get();
      CU(cudaEventRecord(e0, st));
      for (int q = 0; q < n_pieces; q++) {
        const uint32_t t0 = pc_t0[q], nt = pc_t0[q + 1] - pc_t0[q];
        if (nt) {
          A.tile_base = t0;
          CU(hist_launch(row_ch, dim3(nt, n_chunks), u16 ? hist_smem16 : hist_smem, st, A, false, u16));
          tm.kernel_launches++;
          if (q == 0) tm.hist_launches++;
        }
        if (!piped) continue;
        // reduceByKey(_.merge(_)) of Spark's per-partition aggregators, keyed by node exactly as Spark does it: a
        // reduce-scatter leaves each rank with the summed histograms (exact integers) of ITS share of the piece's nodes, it
        // scans those; the small split records are combined at the end with an all-reduce over zero-filled slots.
        cudaEvent_t ev_piece = c->events.get();
        CU(cudaEventRecord(ev_piece, st));
        CU(cudaStreamWaitEvent(sc, ev_piece, 0));
        const int npr = pc_npr[q], size_q = pc_n0[q + 1] - pc_n0[q];
        if (size_q == 0) continue;
        uint32_t* piece_hist = c->d_ghist.as<uint32_t>() + (size_t)pc_s0[q] * slot_words;
        if (pack_slots) {
          const size_t share_words = (size_t)pack_share * (slot_words / 2);
          {
            const cudaEvent_t c0 = comm_begin(sc);
            TRY(nccl_reduce_scatter_u32(c, piece_hist, share_words, sc));
            comm_end(c0, (int64_t)c->world * (int64_t)share_words * 4, sc);
          }
          if (!own_nodes.empty()) {
            kern::scan_owned_kernel<<<dim3((unsigned)own_nodes.size(), n_chunks), kern::kHistThreads, hist_bytes, sc>>>(
                piece_hist, A.nodes, (const uint32_t*)((char*)c->d_group.p + off_own), A.node_feat, A.num_bins_all, f_sub, B, C,
                k_chunk, n_chunks, P.impurity, P.min_instances_per_node, P.min_info_gain, A.results, A.left_counts,
                A.total_counts);
            tm.kernel_launches++;
            CU(cudaGetLastError());
          }
          continue;
        }
        {
          const cudaEvent_t c0 = comm_begin(sc);
          TRY(nccl_reduce_scatter_u32(c, piece_hist, (size_t)npr * slot_words, sc));
          comm_end(c0, (int64_t)npr * c->world * (int64_t)slot_words * 4, sc);
        }
        const int first_local = c->rank * npr;
        const int mine_n = std::max(0, std::min(npr, size_q - first_local));
        if (mine_n > 0) {
          const size_t first = (size_t)pc_n0[q] + first_local;      // group index of this rank's first node of the piece
          kern::scan_kernel<<<dim3(mine_n, n_chunks), kern::kHistThreads, hist_bytes, sc>>>(
              piece_hist + (size_t)first_local * slot_words, A.node_feat + first * f_sub, A.num_bins_all, f_sub,
              B, C, k_chunk, n_chunks, P.impurity, P.min_instances_per_node, P.min_info_gain,
              A.results + first * n_chunks, A.left_counts + first * n_chunks * C, A.total_counts + first * n_chunks * C);
          tm.kernel_launches++;
          CU(cudaGetLastError());
        }
      }
      CU(cudaEventRecord(e1, st));
      hist_ev.emplace_back(e0, e1);
      CU(cudaGetLastError());
    }
    if (c->world > 1) {
      if (row_sharded && push_mode) {
        // every rank's partials must have landed in the owners' staging buffers: a 1-word all-reduce is the
        // cross-rank barrier (in stream order it follows this rank's histogram kernel)
        CU(cudaMemsetAsync(c->d_tickets.p, 0, 4, st));
        TRY(nccl_allreduce_u32(c, c->d_tickets.as<uint32_t>(), 1));
        const int first = c->rank * nodes_per_rank;
        const int mine_n = std::max(0, std::min(nodes_per_rank, n_nodes - first));
        if (mine_n > 0) {
          kern::scan_sum_kernel<<<dim3(mine_n, n_chunks), kern::kHistThreads, hist_bytes, st>>>(
              c->p_stage.as<uint32_t>(), c->world, nodes_per_rank, first, A.node_feat, A.num_bins_all, f_sub, B, C, k_chunk,
              n_chunks, P.impurity, P.min_instances_per_node, P.min_info_gain, A.results, A.left_counts, A.total_counts);
          tm.kernel_launches++;
          CU(cudaGetLastError());
        }
      }
      // the split records of all ranks (each rank wrote its nodes' records, zeros elsewhere)
      cudaStream_t sr = piped ? sc : st;
      {
        const cudaEvent_t c0 = comm_begin(sr);
        TRY(nccl_allreduce_u32(c, (uint32_t*)c->d_results.p, res_dec / 4, sr));
        comm_end(c0, (int64_t)res_dec, sr);
      }
      if (piped) {
        cudaEvent_t ev_comm = c->events.get();
        CU(cudaEventRecord(ev_comm, sc));
        CU(cudaStreamWaitEvent(st, ev_comm, 0));
      }
    }
    // ---- decisions and partition work on the device, the partition kernels right behind.  The host gets the
    // compact scheduling records first (zero-copy stores of the deciding kernel) and the full records after them
    // (second stream).
    kern::SplitNode* dsn = (kern::SplitNode*)c->d_split.p;
    uint32_t* dtn = (uint32_t*)((char*)c->d_split.p + off_ptn);
    uint32_t* d_tile_left = (uint32_t*)c->d_part.p;
    uint64_t* d_nleft = (uint64_t*)((char*)c->d_part.p + off_nleft);
    uint32_t* d_limits = (uint32_t*)((char*)c->d_part.p + off_limits);
    uint4* d_sched = (uint4*)((char*)c->d_part.p + off_sched);
    uint32_t* d_ptiles = (uint32_t*)((char*)c->d_part.p + off_ptiles);
    CU(c->h_sched.ensure((size_t)n_nodes * 16));
    kern::DecideArgs D;
    D.results = A.results; D.nodes = A.nodes; D.node_feat = A.node_feat;
    D.node_info = (const uint32_t*)((char*)c->d_group.p + off_info);
    D.segtab = c->p_segtab.as<kern::Seg>();
    D.out_splits = dsn; D.out_tile_node = dtn;
    D.out_dec = (uint32_t*)((char*)c->d_results.p + res_dec);
    D.out_sched = d_sched; D.out_tiles = d_ptiles;
    D.out_limits = d_limits;
    D.n_nodes = n_nodes; D.n_chunks = n_chunks; D.f_sub = f_sub; D.max_depth = P.max_depth;
    D.next_slot = next_slot;
    D.out_sched_host = nullptr;
    if (!fuse_decide) {   // zero-copy scheduling records here too: the host waits for this kernel's event
      void* hs_dev = nullptr;
      CU(cudaHostGetDevicePointer(&hs_dev, c->h_sched.p, 0));
      D.out_sched_host = (uint4*)hs_dev;
      kern::decide_flags_kernel<<<(n_nodes + 255) / 256, 256, 0, st>>>(D);
    }
    cudaEvent_t ev_dec = c->events.get(), rec_done = c->events.get();
    CU(cudaEventRecord(ev_dec, st));
    CU(cudaStreamWaitEvent(c->stream2, ev_dec, 0));
    CU(cudaMemcpyAsync(h_res, c->d_results.p, res_bytes, cudaMemcpyDeviceToHost, c->stream2));
    CU(cudaEventRecord(rec_done, c->stream2));
    tm.d2h_bytes += (int64_t)n_nodes * 16;
    kern::decide_scan_kernel<<<1, 1024, 0, st>>>(D);
    tm.kernel_launches += fuse_decide ? 1 : 2;
    CU(cudaGetLastError());
    if (ptiles_ub) {
      cudaEvent_t p0 = c->events.get(), p1 = c->events.get();
      CU(cudaEventRecord(p0, st));
      static const bool three_pass = env_flag("EMRF_PART_3PASS");   // count / scan / scatter (stable order), for A/B runs
      if (three_pass) {
        kern::part_count_kernel<<<ptiles_ub, kern::kPartThreads, 0, st>>>(d_ent[0]->as<uint64_t>(), d_ent[1]->as<uint64_t>(),
                                                                         d_cm.as<uint8_t>(), Np, dsn, dtn, d_tile_left,
                                                                         c->p_segtab.as<kern::Seg>(), d_limits);
        kern::part_scan_kernel<<<n_nodes, 256, 0, st>>>(dsn, d_tile_left, d_nleft, c->p_segtab.as<kern::Seg>(), d_limits);
        kern::part_scatter_kernel<<<ptiles_ub, kern::kPartThreads, 0, st>>>(d_ent[0]->as<uint64_t>(), d_ent[1]->as<uint64_t>(),
                                                                           d_ent[0]->as<uint64_t>(), d_ent[1]->as<uint64_t>(),
                                                                           d_cm.as<uint8_t>(), Np, dsn, dtn, d_tile_left, d_nleft,
                                                                           c->p_segtab.as<kern::Seg>(), d_limits);
        tm.kernel_launches += 3;
      } else {
        // one pass: every entry and bin byte is read once; tiles chain their left counts by look-back
        unsigned long long* d_status = (unsigned long long*)((char*)c->d_part.p + off_status);
        uint32_t* d_ticket = (uint32_t*)((char*)c->d_part.p + off_status + (size_t)std::max<uint32_t>(ptiles_ub, 1) * 8);
        CU(cudaMemsetAsync(d_status, 0, status_bytes, st));
        kern::part_onepass_kernel<<<ptiles_ub, kern::kPartThreads, 0, st>>>(d_ent[0]->as<uint64_t>(), d_ent[1]->as<uint64_t>(),
                                                                           d_ent[0]->as<uint64_t>(), d_ent[1]->as<uint64_t>(),
                                                                           d_cm.as<uint8_t>(), Np, dsn, dtn, d_status, d_ticket, d_nleft,
                                                                           c->p_segtab.as<kern::Seg>(), d_limits);
        tm.kernel_launches += 1;
      }
      CU(cudaEventRecord(p1, st));
      part_ev.emplace_back(p0, p1);
      CU(cudaGetLastError());
    }
    if (self_clean) {
      CU(cudaMemsetAsync(c->d_tickets.p, 0, n_rec * 4, st));
      if (n_slots_alloc) CU(cudaMemsetAsync(c->d_ghist.p, 0, (size_t)n_slots_alloc * slot_words * 4, st));
    }
    lap(1);
    // ---- bookkeeping nobody waits for, done while the GPU works: the previous group's node statistics, this
    // group's trace
    {
      t_host0 = now_ms();
      TRY(fill_stats(pending));
      const size_t at0 = fo->trace.size();
      fo->trace.resize(at0 + (size_t)n_nodes * (2 + f_sub));
      int32_t* tr_out = fo->trace.data() + at0;
#pragma omp parallel for num_threads(nt_host) schedule(static)
      for (int i = 0; i < n_nodes; i++) {
        const GroupEntry& g = group[i];
        int32_t* o = tr_out + (size_t)i * (2 + f_sub);
        o[0] = g.tree;
        o[1] = g.id;
        for (int k = 0; k < f_sub; k++) o[2 + k] = group_feat[g.feat_at + k];
      }
      int64_t built = 0;
      for (int i = 0; i < n_nodes; i++)
        if (gn[i].n_tiles) { hist_slots.push_back(gn[i].seg_slot); built++; }
      tm.hist_alg_bytes += built * (int64_t)slot_words * 4;
      tm.total_nodes_split += n_nodes;
      const double d_ = now_ms() - t_host0; host_ms += d_; host_sel += d_;
    }
    lap(2);

    // ---- host side of findBestSplits, the part the next group waits for: the scheduling records say what
    // became of every node; numbering repeats decide_scan_kernel's (group order); children get identity, segment
    // slot and size bounds.  Everything else about the nodes is filled in later (fill_stats).
    CU(cudaEventSynchronize(ev_dec));
    t_host0 = now_ms();
    lap(3);
    const uint4* sched = (const uint4*)c->h_sched.p;
    upd_vec.resize(n_nodes);
    Upd* upd = upd_vec.data();
    for (int t : cur->trees_seen) pool_next[t] = (int)trees[t].pool.size();
    int ns = 0;
    for (int gi = 0; gi < n_nodes; gi++) {
      Upd& u = upd[gi];
      u.dec = sched[gi].x;
      u.left_pool = -1; u.slot_l = 0; u.split_idx = 0;
      if (!(u.dec & kern::kDecNonLeaf)) continue;
      u.slot_l = next_slot; next_slot += 2;
      int& pn = pool_next[group[gi].tree];
      u.left_pool = pn; pn += 2;
      if (u.dec & kern::kDecSplit) u.split_idx = (uint32_t)ns++;
    }
    for (int t : cur->trees_seen) trees[t].pool.resize_uninit((size_t)pool_next[t]);
    lap(5);
    // rows sharded: the weight bound is global (world times too loose), so read the exact sizes back
    if (row_sharded && ns && !dev_tiles) {
      CU(c->h_part.ensure((size_t)ns * 8));
      CU(cudaMemcpyAsync(c->h_part.p, d_nleft, (size_t)ns * 8, cudaMemcpyDeviceToHost, st));
      tm.d2h_bytes += (int64_t)ns * 8;
      child_pos.assign((size_t)ns, ChildPos{-1, -1, 0});
    }
    lap(8);
    // ---- children go on the stack in Scala map iteration order (tree by tree, nodes in pop order), carrying their
    // identity, segment slot and size bounds: a segment cannot hold more entries than its parent's, nor more than the
    // node's total weight (every entry weighs >= 1) — good enough to tile the next launches without reading sizes back
    for (size_t q = 0; q < cur->trees_seen.size(); q++) tree_slot[cur->trees_seen[q]] = (int)q;
    const std::vector<int> order = hl::group_tree_visit_order(cur->trees_seen);
    // runs in visit order (rank of their tree, then pop order)
    for (size_t r = 0; r < order.size(); r++) tree_slot[order[r]] = (int)r;
    const std::vector<GroupSel::Run>& runs = cur->runs;
    run_order.resize(runs.size());
    for (size_t i = 0; i < runs.size(); i++) run_order[i] = (int)i;
    if (runs.size() > 1)
      std::stable_sort(run_order.begin(), run_order.end(), [&](int a, int b) { return tree_slot[runs[a].tree] < tree_slot[runs[b].tree]; });
    // pass 1 (sequential, light): where each node's children go
    push_off.resize(n_nodes);
    uint32_t n_push = 0;
    for (int ri : run_order) {
      const GroupSel::Run& rr = runs[ri];
      for (int gi = rr.start; gi < rr.start + rr.len; gi++) {
        const uint32_t d = upd[gi].dec;
        push_off[gi] = n_push;
        if (d & kern::kDecNonLeaf) n_push += ((d & kern::kDecLLeaf) ? 0u : 1u) + ((d & kern::kDecRLeaf) ? 0u : 1u);
      }
    }
    const size_t stack_base = stack.size();
    stack.resize(stack_base + n_push);
    // pass 2 (parallel): the entries themselves
#pragma omp parallel for num_threads(nt_host) schedule(static)
    for (int gi = 0; gi < n_nodes; gi++) {
      const Upd& u = upd[gi];
      if (!(u.dec & kern::kDecNonLeaf)) continue;
      const GroupEntry& par = group[gi];
      const uint32_t lw = sched[gi].y, tw = sched[gi].z;
      StackEnt ch;
      ch.tree = par.tree; ch.level = par.level + 1;
      size_t at = stack_base + push_off[gi];
      int pos_l = -1, pos_r = -1;
      if (!(u.dec & kern::kDecLLeaf)) {
        ch.pool = u.left_pool; ch.id = 2 * par.id; ch.slot = u.slot_l; ch.ub = std::min(par.ub, lw); ch.weight = lw;
        pos_l = (int)at;
        stack[at++] = ch;
      }
      if (!(u.dec & kern::kDecRLeaf)) {
        ch.pool = u.left_pool + 1; ch.id = 2 * par.id + 1; ch.slot = u.slot_l + 1; ch.ub = std::min(par.ub, tw - lw); ch.weight = tw - lw;
        pos_r = (int)at;
        stack[at++] = ch;
      }
      if (row_sharded && !dev_tiles && (u.dec & kern::kDecSplit)) child_pos[u.split_idx] = ChildPos{pos_l, pos_r, par.ub};
    }
    for (int t : cur->trees_seen) tree_slot[t] = -1;
    const size_t stack_before_select = stack.size();
    // the tree nodes themselves (children included) are built from the full records one group later (fill_stats)
    pending.sel = cur; pending.h_res = h_res; pending.res_left = res_left; pending.res_total = res_total;
    pending.n_nodes = n_nodes; pending.rec_done = rec_done; pending.nt = nt_host;
    std::swap(pending.upd, upd_vec);
    { const double d_ = now_ms() - t_host0; host_ms += d_; host_upd += d_; }
    lap(9);
    // ---- the host selects the next group (needs only the stack) while the GPU partitions
    {
      const double t0 = now_ms();
      select_group(*nxt);
      const double d_ = now_ms() - t0; host_ms += d_; host_sel += d_;
    }
    TRY(select_rc);
    lap(10);
    if (ns && row_sharded && !dev_tiles) {
      CU(cudaStreamSynchronize(st));
      t_host0 = now_ms();
      // exact sizes of the children just created: they are either still on the stack or already in the next group
      // (popped strictly from the top, so group index i holds what was at stack position stack_before_select - 1 - i)
      const uint64_t* nl = (const uint64_t*)c->h_part.p;
      auto patch = [&](int pos, uint32_t v) {
        if (pos < 0) return;
        if ((size_t)pos < stack.size()) stack[pos].ub = v;
        else nxt->group[stack_before_select - 1 - (size_t)pos].ub = v;
      };
      for (int i = 0; i < ns; i++) {
        patch(child_pos[i].l, (uint32_t)nl[i]);
        patch(child_pos[i].r, child_pos[i].parent_ub - (uint32_t)nl[i]);
      }
      { const double d_ = now_ms() - t_host0; host_ms += d_; host_upd += d_; }
    }
    std::swap(cur, nxt);
  }

  TRY(fill_stats(pending));   // the last group's
  c->last_nodes_split = hist_slots.size();
  c->last_trace_len = fo->trace.size();
  if (env_flag("EMRF_DEBUG_TIMERS"))
    fprintf(stderr, "[emrf] host phases ms: prep %.2f | launch %.2f | bookkeeping %.2f | wait %.2f | numbering %.2f | nodes %.2f | "
            "stack %.2f | select %.2f | other %.2f\n", ph[0], ph[1], ph[2], ph[3], ph[5], ph[8], ph[9], ph[10], ph[11]);
  // ---- exact entry counts of every histogram built (the roofline's algorithmic bytes), from the table
  // (the table comes back into page-locked memory while the forest is exported)
  CU(c->h_seg.ensure((size_t)next_slot * sizeof(kern::Seg)));
  CU(cudaMemcpyAsync(c->h_seg.p, c->p_segtab.p, (size_t)next_slot * sizeof(kern::Seg), cudaMemcpyDeviceToHost, st));
  // ---- topNodes.map(_.toNode(prune = true))
  double t_host0 = now_ms();
  // many ranks on one host: each exports its share of the trees (EMRF_EXPORT_SHARD=0/1 overrides the world > 4 rule;
  // measured on B200s with 16 host cores: 2 ranks 1.7 -> 3.2 ms, 4 ranks 3.1 -> 2.9 ms including the copies' tail)
  static const int export_shard_env = [] { const char* e = getenv("EMRF_EXPORT_SHARD"); return e && *e ? atoi(e) : -1; }();
  const bool export_shard = c->world > 1 && (export_shard_env >= 0 ? export_shard_env != 0 : c->world > 4);
  if (export_shard) TRY(export_forest_sharded(c, trees, C, *fo, st));
  else if (!export_forest(trees, C, *fo, c->host_threads)) return fail(EMRF_ERR_INTERNAL, "out of host memory for the model");
  { const double d_ = now_ms() - t_host0; host_ms += d_; tm.host_export_ms = d_; }
  CU(cudaEventRecord(ev_end, st));
  CU(cudaStreamSynchronize(st));
  {
    const kern::Seg* segs = (const kern::Seg*)c->h_seg.p;
    const size_t n_hs = hist_slots.size();
    int64_t units = 0;
#pragma omp parallel for num_threads(c->host_threads) schedule(static) reduction(+ : units)
    for (size_t i = 0; i < n_hs; i++) units += segs[hist_slots[i]].len;
    tm.hist_units = units;
    tm.hist_alg_bytes += units * (int64_t)(f_sub + 6);
  }

  float ms = 0;
  CU(cudaEventElapsedTime(&ms, ev_start, ev_end)); tm.total_ms = ms;
  tm.h2d_ms = 0.0;   // host-resident features: until the last chunk has landed (findSplits and binning run underneath)
  if (copied) { CU(cudaEventElapsedTime(&ms, ev_start, x_chunks.back().ev)); tm.h2d_ms = ms; }
  CU(cudaEventElapsedTime(&ms, ev_in, ev_fs)); tm.find_splits_ms = ms;
  CU(cudaEventElapsedTime(&ms, ev_fs, ev_bin)); tm.bin_ms = ms;
  CU(cudaEventElapsedTime(&ms, ev_bin, ev_bag)); tm.bag_ms = ms;   // wait for the bag kernel (if still running) + list building
  CU(cudaEventElapsedTime(&ms, ev_b0, ev_b1)); tm.bag_kernel_ms = ms;
  for (auto& pr : hist_ev) { CU(cudaEventElapsedTime(&ms, pr.first, pr.second)); tm.hist_kernel_ms += ms; }
  for (auto& pr : part_ev) { CU(cudaEventElapsedTime(&ms, pr.first, pr.second)); tm.partition_ms += ms; }
  for (auto& pr : comm_ev) { CU(cudaEventElapsedTime(&ms, pr.first, pr.second)); tm.comm_ms += ms; }
  if (env_flag("EMRF_DEBUG_TIMERS") && copied) {
    fprintf(stderr, "[emrf] feature chunks landed at ms:");
    for (const XChunk& xc : x_chunks) { CU(cudaEventElapsedTime(&ms, ev_start, xc.ev)); fprintf(stderr, " %.2f", ms); }
    fprintf(stderr, "\n");
  }
  if (env_flag("EMRF_DEBUG_TIMERS")) {
    // GPU timeline of the group loop: hist end -> (decide) -> partition start, partition end -> next hist start
    double gap_dec = 0, gap_next = 0, first_to_last = 0;
    const bool paired = hist_ev.size() == part_ev.size();
    for (size_t i = 0; paired && i < hist_ev.size(); i++) {
      CU(cudaEventElapsedTime(&ms, hist_ev[i].second, part_ev[i].first)); gap_dec += ms;
      if (i + 1 < hist_ev.size()) { CU(cudaEventElapsedTime(&ms, part_ev[i].second, hist_ev[i + 1].first)); gap_next += ms; }
    }
    if (!hist_ev.empty()) { CU(cudaEventElapsedTime(&ms, hist_ev.front().first, hist_ev.back().second)); first_to_last = ms; }
    CU(cudaEventElapsedTime(&ms, ev_bag, hist_ev.front().first));
    fprintf(stderr, "[emrf] GPU timeline ms: lists->first hist %.2f | loop %.2f (hist %.2f, part %.2f, hist->part gaps %.2f, part->next hist gaps %.2f)%s\n",
            (double)ms, first_to_last, tm.hist_kernel_ms, tm.partition_ms, gap_dec, gap_next, paired ? "" : " [some groups had no partition]");
    CU(cudaEventElapsedTime(&ms, hist_ev.back().second, ev_end));
    fprintf(stderr, "[emrf] last hist -> end %.2f ms\n", (double)ms);
  }
  tm.find_best_splits_ms = tm.hist_kernel_ms;
  tm.host_scheduler_ms = host_ms;
  tm.host_select_ms = host_sel;
  tm.host_update_ms = host_upd;
  tm.num_groups = (int64_t)fo->group_sizes.size();
  *out = fo_guard.release();
  return EMRF_OK;
}

// Nothing may be thrown across the C boundary, and a call that fails half-way must not leave work of its own in
// flight on the context's streams (the next call would reuse the buffers under it).
template <class Fn>
static int guarded(emrf_ctx* c, Fn&& fn) {
  int rc;
  try {
    rc = fn();
  } catch (const std::bad_alloc&) {
    rc = fail(EMRF_ERR_INTERNAL, "out of host memory");
  } catch (const std::exception& e) {
    rc = fail(EMRF_ERR_INTERNAL, "internal error: %s", e.what());
  } catch (...) {
    rc = fail(EMRF_ERR_INTERNAL, "internal error");
  }
  if (rc != EMRF_OK && c) {
    const std::string msg = emrf_last_error();      // the synchronisations below must not replace the message
    if (c->stream) cudaStreamSynchronize(c->stream);
    if (c->stream2) cudaStreamSynchronize(c->stream2);
    if (c->stream3) cudaStreamSynchronize(c->stream3);
    if (c->stream_comm) cudaStreamSynchronize(c->stream_comm);
    cudaGetLastError();
    fail(rc, "%s", msg.c_str());
  }
  return rc;
}

extern "C" int emrf_fit(emrf_ctx_t* c, const emrf_params_t* params, const double* features_hd, const double* labels_hd,
                        int64_t n_rows, int32_t n_features, const int64_t* partition_offsets, int32_t n_partitions,
                        emrf_forest_t** out) {
  if (!c || !out || !features_hd || !labels_hd) return fail(EMRF_ERR_ARG, "NULL argument");
  TRY(check_params(params));
  if (n_rows < 1 || n_features < 1) return fail(EMRF_ERR_ARG, "empty dataset");
  TRY(check_partitions(n_rows, partition_offsets, n_partitions));
  std::lock_guard<std::mutex> lk(c->fit_mu);
  return guarded(c, [&] { return fit_impl(c, *params, features_hd, labels_hd, n_rows, n_features, partition_offsets, n_partitions, out); });
}

extern "C" int emrf_fit_sharded(emrf_ctx_t* c, const emrf_params_t* params, const double* features_hd, const double* labels_hd,
                                int64_t n_rows_local, int32_t n_features, const int64_t* partition_offsets_global,
                                int32_t n_partitions_global, int32_t first_partition, int32_t n_partitions_local,
                                emrf_forest_t** out) {
  if (!c || !out || !features_hd || !labels_hd || !partition_offsets_global) return fail(EMRF_ERR_ARG, "NULL argument");
  TRY(check_params(params));
  if (n_partitions_global < 1 || first_partition < 0 || n_partitions_local < 1 ||
      first_partition + n_partitions_local > n_partitions_global)
    return fail(EMRF_ERR_ARG, "bad partition range [%d, %d) of %d", first_partition, first_partition + n_partitions_local,
                n_partitions_global);
  const int64_t n_global = partition_offsets_global[n_partitions_global];
  TRY(check_partitions(n_global, partition_offsets_global, n_partitions_global));
  const int64_t row0 = partition_offsets_global[first_partition];
  if (partition_offsets_global[first_partition + n_partitions_local] - row0 != n_rows_local)
    return fail(EMRF_ERR_ARG, "n_rows_local does not match the partition range");
  if (n_rows_local < 1 || n_features < 1) return fail(EMRF_ERR_ARG, "every rank needs at least one row");
  if (params->num_classes < 1 && c->world > 1) {}  // inferred classes are all-reduced below
  std::vector<int64_t> local_off(n_partitions_local + 1);
  for (int i = 0; i <= n_partitions_local; i++) local_off[i] = partition_offsets_global[first_partition + i] - row0;
  ShardSpec sh;
  sh.first_part = first_partition;
  sh.n_rows_global = n_global;
  sh.goff = partition_offsets_global;
  sh.n_parts_global = n_partitions_global;
  sh.row0 = row0;
  std::lock_guard<std::mutex> lk(c->fit_mu);
  return guarded(c, [&] {
    return fit_impl(c, *params, features_hd, labels_hd, n_rows_local, n_features, local_off.data(), n_partitions_local, out, &sh);
  });
}

// ---------------------------------------------------------------------------------------------
// forest accessors
// ---------------------------------------------------------------------------------------------
extern "C" void emrf_forest_free(emrf_forest_t* f) { delete f; }
extern "C" int emrf_forest_info(const emrf_forest_t* f, int32_t* T, int32_t* C, int32_t* F, int64_t* total_nodes) {
  if (!f) return fail(EMRF_ERR_ARG, "forest is NULL");
  if (T) *T = f->num_trees;
  if (C) *C = f->num_classes;
  if (F) *F = f->num_features;
  if (total_nodes) *total_nodes = (int64_t)f->nodes.size();
  return EMRF_OK;
}
extern "C" int emrf_forest_export(const emrf_forest_t* f, int64_t* tree_off, int32_t* node_id, int32_t* left, int32_t* right,
                                  int32_t* feature, double* threshold, double* gain, double* impurity, double* prediction,
                                  double* stats) {
  if (!f) return fail(EMRF_ERR_ARG, "forest is NULL");
  ensure_double_stats(const_cast<emrf_forest*>(f));
  for (size_t t = 0; t < f->tree_off.size(); t++) tree_off[t] = f->tree_off[t];
  for (size_t i = 0; i < f->nodes.size(); i++) {
    const FlatNode& n = f->nodes[i];
    node_id[i] = n.id; left[i] = n.left; right[i] = n.right; feature[i] = n.feature;
    threshold[i] = n.threshold; gain[i] = n.gain; impurity[i] = n.impurity; prediction[i] = n.prediction;
  }
  memcpy(stats, f->stats.data(), f->stats.size() * 8);
  return EMRF_OK;
}
extern "C" int emrf_forest_import(emrf_ctx_t* c, int32_t T, int32_t C, int32_t F, const int64_t* tree_off, const int32_t* node_id,
                                  const int32_t* left, const int32_t* right, const int32_t* feature, const double* threshold,
                                  const double* gain, const double* impurity, const double* prediction, const double* stats,
                                  emrf_forest_t** out) {
  (void)c;
  if (!out || !tree_off || T < 1 || C < 1 || C > 256 || F < 1) return fail(EMRF_ERR_ARG, "bad forest description");
  if (!left || !right || !feature || !threshold || !stats) return fail(EMRF_ERR_ARG, "forest description: a required array is NULL");
  // The description usually comes from a file (RandomForestClassificationModel.load / fromJson), and predict walks the
  // children on the host and on the device without further checks: reject anything that is not a set of proper trees.
  // Exports are pre-order, so a child index above its parent's rules out cycles; every node but the root must have
  // exactly one parent.
  if (tree_off[0] != 0) return fail(EMRF_ERR_ARG, "forest description: tree_off[0] must be 0");
  for (int32_t t = 0; t < T; t++)
    if (tree_off[t + 1] <= tree_off[t]) return fail(EMRF_ERR_ARG, "forest description: tree %d is empty or tree_off decreases", t);
  if (tree_off[T] >= (1ll << 31)) return fail(EMRF_ERR_LIMIT, "forest too large");
  {
    std::vector<uint8_t> parents;
    for (int32_t t = 0; t < T; t++) {
      const int64_t base = tree_off[t], sz = tree_off[t + 1] - base;
      parents.assign((size_t)sz, 0);
      for (int64_t i = 0; i < sz; i++) {
        const int32_t l = left[base + i], r = right[base + i];
        if (l < 0 && r < 0) continue;
        if (l <= i || r <= i || l >= sz || r >= sz || l == r)
          return fail(EMRF_ERR_ARG, "forest description: tree %d node %lld has bad children (%d, %d)", t, (long long)i, l, r);
        const int32_t ft = feature[base + i];
        if (ft < 0 || ft >= F) return fail(EMRF_ERR_ARG, "forest description: tree %d node %lld: feature %d out of range", t, (long long)i, ft);
        if (parents[l]++ || parents[r]++)
          return fail(EMRF_ERR_ARG, "forest description: tree %d node %lld shares a child with another node", t, (long long)i);
      }
      for (int64_t i = 1; i < sz; i++)
        if (!parents[i]) return fail(EMRF_ERR_ARG, "forest description: tree %d node %lld is unreachable", t, (long long)i);
    }
  }
  emrf_forest* f = new emrf_forest();
  f->num_trees = T; f->num_classes = C; f->num_features = F;
  f->tree_off.assign(tree_off, tree_off + T + 1);
  const int64_t n = tree_off[T];
  if (!f->nodes.resize_uninit((size_t)n)) { delete f; return fail(EMRF_ERR_INTERNAL, "out of host memory for the model"); }
  for (int64_t i = 0; i < n; i++) {
    FlatNode& q = f->nodes[i];
    q.id = node_id ? node_id[i] : 0; q.left = left[i]; q.right = right[i]; q.feature = feature[i];
    q.threshold = threshold[i]; q.gain = gain ? gain[i] : 0.0; q.impurity = impurity ? impurity[i] : 0.0;
    q.prediction = prediction ? prediction[i] : 0.0;
    if (q.left >= 0 && (q.feature < 0 || q.feature >= F)) { delete f; return fail(EMRF_ERR_ARG, "node %lld: bad feature", (long long)i); }
  }
  f->stats.assign(stats, stats + n * C);
  *out = f;
  return EMRF_OK;
}
extern "C" int emrf_forest_trace_info(const emrf_forest_t* f, int32_t* num_groups, int64_t* trace_len, int32_t* f_sub) {
  if (!f) return fail(EMRF_ERR_ARG, "forest is NULL");
  if (num_groups) *num_groups = (int32_t)f->group_sizes.size();
  if (trace_len) *trace_len = (int64_t)f->trace.size();
  if (f_sub) *f_sub = f->f_sub;
  return EMRF_OK;
}
extern "C" int emrf_forest_trace(const emrf_forest_t* f, int32_t* group_sizes, int32_t* trace) {
  if (!f) return fail(EMRF_ERR_ARG, "forest is NULL");
  if (group_sizes) memcpy(group_sizes, f->group_sizes.data(), f->group_sizes.size() * 4);
  if (trace) memcpy(trace, f->trace.data(), f->trace.size() * 4);
  return EMRF_OK;
}
// TreeEnsembleModel.featureImportances (ml/tree/treeModels.scala): per tree, sum over internal
// nodes of gain * count; normalise per tree; sum over trees; normalise.
extern "C" int emrf_forest_feature_importances(const emrf_forest_t* f, double* out) {
  if (!f || !out) return fail(EMRF_ERR_ARG, "NULL argument");
  ensure_double_stats(const_cast<emrf_forest*>(f));
  const int F = f->num_features, C = f->num_classes;
  std::vector<double> total(F, 0.0), one(F);
  for (int t = 0; t < f->num_trees; t++) {
    std::fill(one.begin(), one.end(), 0.0);
    for (int64_t i = f->tree_off[t]; i < f->tree_off[t + 1]; i++) {
      const FlatNode& n = f->nodes[i];
      if (n.left < 0) continue;
      double cnt = 0.0;
      for (int c = 0; c < C; c++) cnt += f->stats[i * C + c];
      one[n.feature] += n.gain * cnt;
    }
    double s = 0.0;
    for (int j = 0; j < F; j++) s += one[j];
    if (s != 0.0) for (int j = 0; j < F; j++) total[j] += one[j] / s;
  }
  double s = 0.0;
  for (int j = 0; j < F; j++) s += total[j];
  for (int j = 0; j < F; j++) out[j] = (s != 0.0) ? total[j] / s : total[j];
  return EMRF_OK;
}

// ---------------------------------------------------------------------------------------------
// emrf_predict
// ---------------------------------------------------------------------------------------------
static int forest_to_device(emrf_ctx* c, emrf_forest* f) {
  std::lock_guard<std::mutex> lk(f->dev_mu);
  if (f->dev_device == c->device) return EMRF_OK;
  if (f->dev_device >= 0) {   // the forest was last used on another GPU: that copy goes back to its own device's cache
    for (DevBuf* b : {&f->d_nodes, &f->d_roots, &f->d_leaf_prob, &f->d_wnodes, &f->d_wroots, &f->d_pthr, &f->d_pns, &f->d_wblocks,
                      &f->d_broots, &f->d_leaf_rec})
      b->release();
    f->dev_device = -1;
  }
  ensure_double_stats(f);
  const int C = f->num_classes;
  const int64_t n = (int64_t)f->nodes.size();
  std::vector<kern::DevNode> dn(n);
  std::vector<int64_t> roots(f->num_trees);
  std::vector<double> leaf_prob;
  int64_t n_leaves = 0;
  for (int t = 0; t < f->num_trees; t++) {
    const int64_t base = f->tree_off[t];
    roots[t] = base;
    for (int64_t i = base; i < f->tree_off[t + 1]; i++) {
      const FlatNode& q = f->nodes[i];
      kern::DevNode& d = dn[i];
      d.pad = 0;
      if (q.left >= 0) {
        d.feature = q.feature; d.threshold = q.threshold;
        d.left = (int32_t)(base + q.left); d.right = (int32_t)(base + q.right);
      } else {
        d.feature = -1; d.threshold = 0.0; d.left = (int32_t)n_leaves;
        double total = 0.0;
        for (int cc = 0; cc < C; cc++) total += f->stats[i * C + cc];
        d.right = (total != 0.0) ? 1 : 0;
        leaf_prob.resize((size_t)(n_leaves + 1) * C);
        for (int cc = 0; cc < C; cc++)
          leaf_prob[(size_t)n_leaves * C + cc] = (total != 0.0) ? f->stats[i * C + cc] / total : 0.0;
        n_leaves++;
      }
    }
  }
  if (n >= (1ll << 31)) return fail(EMRF_ERR_LIMIT, "forest too large");
  CU(f->d_nodes.alloc((size_t)n * sizeof(kern::DevNode)));
  CU(f->d_roots.alloc((size_t)f->num_trees * 8));
  CU(f->d_leaf_prob.alloc(leaf_prob.size() * 8));
  CU(cudaMemcpy(f->d_nodes.p, dn.data(), (size_t)n * sizeof(kern::DevNode), cudaMemcpyHostToDevice));
  CU(cudaMemcpy(f->d_roots.p, roots.data(), (size_t)f->num_trees * 8, cudaMemcpyHostToDevice));
  CU(cudaMemcpy(f->d_leaf_prob.p, leaf_prob.data(), leaf_prob.size() * 8, cudaMemcpyHostToDevice));
  // ---- binned-walk layout: per-feature sorted tables of every threshold the forest compares with,
  // nodes re-laid breadth-first so that siblings are adjacent (right = left + 1)
  {
    const int F = f->num_features;
    std::vector<std::vector<double>> tab(F);
    for (int64_t i = 0; i < n; i++)
      if (f->nodes[i].left >= 0) tab[f->nodes[i].feature].push_back(f->nodes[i].threshold);
    size_t ts = 1;
    bool ok = (F <= 65535);
    for (int j = 0; j < F; j++) {
      std::sort(tab[j].begin(), tab[j].end());
      tab[j].erase(std::unique(tab[j].begin(), tab[j].end()), tab[j].end());
      for (double v : tab[j]) if (v != v) ok = false;
      ts = std::max(ts, tab[j].size());
    }
    if (ts > 255) ok = false;   // bins must fit uint8; such forests use the f64 walk
    f->walk_ok = ok;
    if (ok) {
      f->walk_ts = (int)ts;
      std::vector<double> pthr((size_t)F * ts, 0.0);
      std::vector<int32_t> pns(F);
      for (int j = 0; j < F; j++) {
        pns[j] = (int32_t)tab[j].size();
        std::copy(tab[j].begin(), tab[j].end(), pthr.begin() + (size_t)j * ts);
      }
      std::vector<kern::WalkNode> wn(n);
      std::vector<int32_t> wroots(f->num_trees);
      std::vector<int64_t> queue;
      int64_t next_free = 0;
      // leaf rows must match the legacy numbering (leaf_prob was filled in flat pre-order)
      std::vector<int32_t> leaf_of(n, -1);
      for (int64_t i = 0; i < n; i++)
        if (f->nodes[i].left < 0) leaf_of[i] = dn[i].right ? dn[i].left : -1;
      for (int t = 0; t < f->num_trees; t++) {
        const int64_t base = f->tree_off[t];
        wroots[t] = (int32_t)next_free;
        queue.clear();
        queue.push_back(base);
        std::vector<int64_t> slot_of;   // BFS slot for queue entry q = wroots[t] + q
        next_free += 1;
        for (size_t q = 0; q < queue.size(); q++) {
          const int64_t i = queue[q];
          const FlatNode& nd = f->nodes[i];
          kern::WalkNode& w = wn[wroots[t] + (int64_t)q];
          if (nd.left >= 0) {
            const auto& tb = tab[nd.feature];
            const int sb = (int)(std::lower_bound(tb.begin(), tb.end(), nd.threshold) - tb.begin());
            w.left = (int32_t)next_free; w.feature = (uint16_t)nd.feature; w.split_bin = (uint8_t)sb; w.is_leaf = 0;
            queue.push_back(base + nd.left);
            queue.push_back(base + nd.right);
            next_free += 2;
          } else {
            w.left = leaf_of[i]; w.feature = 0; w.split_bin = 0; w.is_leaf = 1;
          }
        }
      }
      CU(f->d_wnodes.alloc((size_t)n * sizeof(kern::WalkNode)));
      CU(f->d_wroots.alloc((size_t)f->num_trees * 4));
      CU(f->d_pthr.alloc(pthr.size() * 8));
      CU(f->d_pns.alloc((size_t)F * 4));
      CU(cudaMemcpy(f->d_wnodes.p, wn.data(), (size_t)n * sizeof(kern::WalkNode), cudaMemcpyHostToDevice));
      CU(cudaMemcpy(f->d_wroots.p, wroots.data(), (size_t)f->num_trees * 4, cudaMemcpyHostToDevice));
      CU(cudaMemcpy(f->d_pthr.p, pthr.data(), pthr.size() * 8, cudaMemcpyHostToDevice));
      CU(cudaMemcpy(f->d_pns.p, pns.data(), (size_t)F * 4, cudaMemcpyHostToDevice));
      // compact leaf records: integer class counts (true of every forest trained here or by Spark) instead of the
      // per-class probabilities; the walk recomputes count / total exactly
      bool integral = C <= 256;
      double max_count = 0.0;
      for (int64_t i = 0; i < n && integral; i++) {
        if (f->nodes[i].left >= 0) continue;
        double total = 0.0;
        for (int cc = 0; cc < C; cc++) {
          const double v = f->stats[i * C + cc];
          if (!(v >= 0.0) || v != std::floor(v) || v >= 4294967296.0) { integral = false; break; }
          total += v;
          max_count = std::max(max_count, v);
        }
        if (total >= 4294967296.0) integral = false;
      }
      f->leaf_count_bytes = 0;
      if (integral) {
        const int cb = max_count <= 65535.0 ? 2 : 4;
        const size_t rb = kern::leaf_rec_bytes(C, cb);
        std::vector<uint8_t> recs((size_t)std::max<int64_t>(n_leaves, 1) * rb, 0);
        for (int64_t i = 0; i < n; i++) {
          if (f->nodes[i].left >= 0 || !dn[i].right) continue;     // internal node, or an empty leaf (never read)
          uint8_t* r = recs.data() + (size_t)dn[i].left * rb;
          double total = 0.0;
          for (int cc = 0; cc < C; cc++) total += f->stats[i * C + cc];
          const double ty[2] = {total, 1.0 / total};
          memcpy(r, ty, 16);
          for (int cc = 0; cc < C; cc++) {
            const size_t slot = (size_t)(cc >> 6) * 64 + (size_t)(cc & 31) * 2 + ((cc >> 5) & 1);   // lane's pair: (c, c + 32)
            const uint32_t v = (uint32_t)f->stats[i * C + cc];
            if (cb == 2) reinterpret_cast<uint16_t*>(r + 16)[slot] = (uint16_t)v;
            else reinterpret_cast<uint32_t*>(r + 16)[slot] = v;
          }
        }
        CU(f->d_leaf_rec.alloc(recs.size()));
        CU(cudaMemcpy(f->d_leaf_rec.p, recs.data(), recs.size(), cudaMemcpyHostToDevice));
        // two-level blocks (kern::WalkBlock): breadth-first over block roots, a split node's grandchildren blocks allocated
        // as they are met; a leaf's record index (-1: empty leaf) sits where the pointer to it would be
        std::vector<kern::WalkBlock> blocks;
        std::vector<int32_t> broots(f->num_trees);
        blocks.reserve((size_t)n / 2 + f->num_trees);
        std::vector<std::pair<int64_t, int32_t>> todo;     // (flat node index of a block's root, block index)
        auto leaf_idx = [&](int64_t i) -> int32_t { return dn[i].right ? dn[i].left : -1; };
        auto split_bin_of = [&](const FlatNode& nd) -> uint8_t {
          const auto& tb = tab[nd.feature];
          return (uint8_t)(std::lower_bound(tb.begin(), tb.end(), nd.threshold) - tb.begin());
        };
        for (int t = 0; t < f->num_trees; t++) {
          const int64_t base = f->tree_off[t];
          broots[t] = (int32_t)blocks.size();
          blocks.emplace_back();
          todo.clear();
          todo.emplace_back(base, broots[t]);
          for (size_t q = 0; q < todo.size(); q++) {
            const int64_t i = todo[q].first;
            const int32_t bi = todo[q].second;
            kern::WalkBlock b;
            memset(&b, 0, sizeof(b));
            const FlatNode& nd = f->nodes[i];
            if (nd.left < 0) { b.fl0 = 1; b.leaf0 = leaf_idx(i); blocks[bi] = b; continue; }
            b.f0 = (uint16_t)nd.feature; b.sb0 = split_bin_of(nd);
            for (int side = 0; side < 2; side++) {
              const int64_t ci = base + (side ? nd.right : nd.left);
              const FlatNode& ch = f->nodes[ci];
              uint8_t fl = 0;
              int32_t pl = 0, pr = 0;
              uint16_t cf = 0;
              uint8_t csb = 0;
              if (ch.left < 0) { fl = 1; pl = leaf_idx(ci); }
              else {
                cf = (uint16_t)ch.feature; csb = split_bin_of(ch);
                const int64_t gl = base + ch.left, gr = base + ch.right;
                if (f->nodes[gl].left < 0) { fl |= 2; pl = leaf_idx(gl); }
                else { pl = (int32_t)blocks.size(); blocks.emplace_back(); todo.emplace_back(gl, pl); }
                if (f->nodes[gr].left < 0) { fl |= 4; pr = leaf_idx(gr); }
                else { pr = (int32_t)blocks.size(); blocks.emplace_back(); todo.emplace_back(gr, pr); }
              }
              if (side == 0) { b.f1 = cf; b.sb1 = csb; b.fl1 = fl; b.p1l = pl; b.p1r = pr; }
              else { b.f2 = cf; b.sb2 = csb; b.fl2 = fl; b.p2l = pl; b.p2r = pr; }
            }
            blocks[bi] = b;
          }
        }
        CU(f->d_wblocks.alloc(blocks.size() * sizeof(kern::WalkBlock)));
        CU(f->d_broots.alloc((size_t)f->num_trees * 4));
        CU(cudaMemcpy(f->d_wblocks.p, blocks.data(), blocks.size() * sizeof(kern::WalkBlock), cudaMemcpyHostToDevice));
        CU(cudaMemcpy(f->d_broots.p, broots.data(), (size_t)f->num_trees * 4, cudaMemcpyHostToDevice));
        f->leaf_count_bytes = cb;
      }
    }
  }
  f->dev_device = c->device;
  return EMRF_OK;
}

template <class CT, int CB>
static cudaError_t walk_compact_t(const emrf_forest* f, const uint8_t* bins, int Fp, int64_t n, int C, double* raw, double* prob,
                                  double* pred, const double* thr, cudaStream_t st) {
  const size_t smem = (size_t)kern::kWalk2Warps * 32 * Fp;
  if (smem > 48 * 1024) {
    const cudaError_t e = cudaFuncSetAttribute(kern::predict_walk_compact_kernel<CT, CB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  const unsigned grid = (unsigned)((n + kern::kWalk2Warps * 32 - 1) / (kern::kWalk2Warps * 32));
  kern::predict_walk_compact_kernel<CT, CB><<<grid, kern::kWalk2Warps * 32, smem, st>>>(
      bins, Fp, n, f->d_wblocks.as<kern::WalkBlock>(), f->d_broots.as<int32_t>(), f->num_trees, f->d_leaf_rec.as<uint8_t>(), C, raw, prob,
      pred, thr);
  return cudaGetLastError();
}
static int launch_walk_compact(const emrf_forest* f, const uint8_t* bins, int Fp, int64_t n, int C, double* raw, double* prob, double* pred,
                               const double* thr, cudaStream_t st) {
  const int cb = (C + 63) / 64;
  const bool u16 = f->leaf_count_bytes == 2;
  cudaError_t e;
#define EMRF_WALK_CASE(CB_)                                                                     \
  e = u16 ? walk_compact_t<uint16_t, CB_>(f, bins, Fp, n, C, raw, prob, pred, thr, st)        \
          : walk_compact_t<uint32_t, CB_>(f, bins, Fp, n, C, raw, prob, pred, thr, st)
  if (cb <= 1) { EMRF_WALK_CASE(1); }
  else if (cb == 2) { EMRF_WALK_CASE(2); }
  else { EMRF_WALK_CASE(4); }
  CU(e);
#undef EMRF_WALK_CASE
  return EMRF_OK;
}
static int predict_impl(emrf_ctx_t* c, const emrf_forest_t* f_in, const double* features_hd, int64_t N, double* raw_hd,
                        double* prob_hd, double* pred_hd, const double* thresholds_h) {
  if (!c || !f_in || !features_hd) return fail(EMRF_ERR_ARG, "NULL argument");
  if (N < 0) return fail(EMRF_ERR_ARG, "negative row count");
  emrf_forest* f = const_cast<emrf_forest*>(f_in);
  const int C = f->num_classes, F = f->num_features;
  if (thresholds_h) {   // ProbabilisticClassifierParams.setThresholds: all >= 0, at most one 0
    int zeros = 0;
    for (int q = 0; q < C; q++) {
      if (!(thresholds_h[q] >= 0.0)) return fail(EMRF_ERR_ARG, "thresholds must all be >= 0 (class %d: %g)", q, thresholds_h[q]);
      zeros += thresholds_h[q] == 0.0;
    }
    if (zeros > 1) return fail(EMRF_ERR_ARG, "at most one threshold may be 0 (%d are)", zeros);
  }
  if (N == 0) return EMRF_OK;
  CU(cudaSetDevice(c->device));
  TRY(forest_to_device(c, f));
  // two streams and two sets of staging buffers: chunk i+1's host->device copy and binning run under chunk i's walk and
  // its device->host copies (a stream reuses its own buffer set in stream order, so no host wait inside the loop)
  cudaStream_t sts[2] = {nullptr, nullptr};
  struct StreamGuard { cudaStream_t* s; ~StreamGuard() { for (int i = 0; i < 2; i++) if (s[i]) cudaStreamDestroy(s[i]); } } sg{sts};
  CU(cudaStreamCreateWithFlags(&sts[0], cudaStreamNonBlocking));
  CU(cudaStreamCreateWithFlags(&sts[1], cudaStreamNonBlocking));
  cudaStream_t st = sts[0];
  DevBuf dThr;
  const double* d_thresholds = nullptr;
  if (thresholds_h) {
    CU(dThr.alloc((size_t)C * 8));
    CU(cudaMemcpyAsync(dThr.p, thresholds_h, (size_t)C * 8, cudaMemcpyHostToDevice, st));
    d_thresholds = dThr.as<double>();
  }
  const bool x_dev = is_device_ptr(features_hd);
  const bool raw_dev = raw_hd && is_device_ptr(raw_hd), prob_dev = prob_hd && is_device_ptr(prob_hd),
             pred_dev = pred_hd && is_device_ptr(pred_hd);
  // chunk so that staging buffers stay modest
  const int64_t chunk = std::max<int64_t>(1, std::min<int64_t>(N, (int64_t)(256ll << 20) / ((int64_t)F * 8)));
  const int n_sets = N > chunk ? 2 : 1;
  DevBuf dXs[2], dRaws[2], dProbs[2], dPreds[2], dBinss[2];
  for (int i = 0; i < n_sets; i++) {
    CU(dBinss[i].alloc((size_t)align_up((size_t)chunk, 128) * align_up(F, 16)));
    if (!x_dev) CU(dXs[i].alloc((size_t)chunk * F * 8));
    if (raw_hd && !raw_dev) CU(dRaws[i].alloc((size_t)chunk * C * 8));
    if (prob_hd && !prob_dev) CU(dProbs[i].alloc((size_t)chunk * C * 8));
    if (pred_hd && !pred_dev) CU(dPreds[i].alloc((size_t)chunk * 8));
  }
  if (d_thresholds) CU(cudaStreamSynchronize(sts[0]));   // the thresholds were copied on stream 0; stream 1 reads them too
  int64_t chunk_idx = 0;
  for (int64_t r0 = 0; r0 < N; r0 += chunk, chunk_idx++) {
    const int set = (int)(chunk_idx & 1) % n_sets;
    st = sts[set];
    DevBuf &dX = dXs[set], &dRaw = dRaws[set], &dProb = dProbs[set], &dPred = dPreds[set], &dBins = dBinss[set];
    const int64_t n = std::min(chunk, N - r0);
    const double* x = features_hd + r0 * F;
    if (!x_dev) {
      CU(cudaMemcpyAsync(dX.p, x, (size_t)n * F * 8, cudaMemcpyHostToDevice, st));
      x = dX.as<double>();
    }
    double* raw = raw_hd ? (raw_dev ? raw_hd + r0 * C : dRaw.as<double>()) : nullptr;
    double* prob = prob_hd ? (prob_dev ? prob_hd + r0 * C : dProb.as<double>()) : nullptr;
    double* pred = pred_hd ? (pred_dev ? pred_hd + r0 : dPred.as<double>()) : nullptr;
    const int Fp = (int)align_up(F, 16);
    const size_t walk_smem = kern::predict_smem_bytes(Fp, C);
    static const bool force_f64_walk = env_flag("EMRF_PREDICT_F64_WALK");
    if (f->walk_ok && !force_f64_walk && walk_smem <= 200 * 1024) {
      // bin the chunk once (numeric compare against the forest's own threshold tables), then walk bins
      const size_t tile_bytes = kern::bin_tile_bytes(Fp);
      const size_t thr_bytes = kern::bin_thr_smem_bytes(F, f->walk_ts);
      const unsigned bgrid = (unsigned)std::min<int64_t>((n + kern::kBinRows - 1) / kern::kBinRows, (int64_t)c->sm_count * 8);
      if (tile_bytes + thr_bytes <= 160 * 1024) {
        const size_t smem = tile_bytes + thr_bytes;
        TRY((bin_launch_smem<true>(kern::bin_key_row(F), bgrid, smem, st, x, n, F, Fp, 0, f->d_pthr.as<double>(), f->d_pns.as<int>(),
                                   f->walk_ts, f->walk_ts, dBins.as<uint8_t>(), nullptr)));
      } else {
        if (tile_bytes > 48 * 1024) CU(cudaFuncSetAttribute(kern::bin_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tile_bytes));
        kern::bin_kernel<false, true><<<bgrid, 256, tile_bytes, st>>>(x, n, F, Fp, 0, f->d_pthr.as<double>(), f->d_pns.as<int>(),
                                                                      f->walk_ts, f->walk_ts, dBins.as<uint8_t>(), nullptr);
      }
      static const bool dense_leaves = env_flag("EMRF_PREDICT_DENSE_LEAVES");   // the f64 probability records (A/B runs, parity test)
      if (f->leaf_count_bytes && !dense_leaves) {
        TRY(launch_walk_compact(f, dBins.as<uint8_t>(), Fp, n, C, raw, prob, pred, d_thresholds, st));
      } else {
        if (walk_smem > 48 * 1024) CU(cudaFuncSetAttribute(kern::predict_walk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)walk_smem));
        const unsigned wgrid = (unsigned)((n + kern::kPredWarps * 32 - 1) / (kern::kPredWarps * 32));
        kern::predict_walk_kernel<<<wgrid, kern::kPredWarps * 32, walk_smem, st>>>(
            dBins.as<uint8_t>(), Fp, n, f->d_wnodes.as<kern::WalkNode>(), f->d_wroots.as<int32_t>(), f->num_trees,
            f->d_leaf_prob.as<double>(), C, raw, prob, pred, d_thresholds);
      }
    } else {
      const int64_t threads = n * 32;
      kern::predict_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, st>>>(x, n, F, f->d_nodes.as<kern::DevNode>(),
                                                                           f->d_roots.as<int64_t>(), f->num_trees,
                                                                           f->d_leaf_prob.as<double>(), C, raw, prob, pred, d_thresholds);
    }
    CU(cudaGetLastError());
    if (raw_hd && !raw_dev) CU(cudaMemcpyAsync(raw_hd + r0 * C, raw, (size_t)n * C * 8, cudaMemcpyDeviceToHost, st));
    if (prob_hd && !prob_dev) CU(cudaMemcpyAsync(prob_hd + r0 * C, prob, (size_t)n * C * 8, cudaMemcpyDeviceToHost, st));
    if (pred_hd && !pred_dev) CU(cudaMemcpyAsync(pred_hd + r0, pred, (size_t)n * 8, cudaMemcpyDeviceToHost, st));
  }
  CU(cudaStreamSynchronize(sts[0]));
  CU(cudaStreamSynchronize(sts[1]));
  return EMRF_OK;
}
extern "C" int emrf_predict(emrf_ctx_t* c, const emrf_forest_t* f, const double* features_hd, int64_t N, double* raw_hd,
                            double* prob_hd, double* pred_hd) {
  return guarded(nullptr, [&] { return predict_impl(c, f, features_hd, N, raw_hd, prob_hd, pred_hd, nullptr); });
}
extern "C" int emrf_predict_thresholds(emrf_ctx_t* c, const emrf_forest_t* f, const double* features_hd, int64_t N,
                                       const double* thresholds, double* raw_hd, double* prob_hd, double* pred_hd) {
  return guarded(nullptr, [&] { return predict_impl(c, f, features_hd, N, raw_hd, prob_hd, pred_hd, thresholds); });
}

// ---------------------------------------------------------------------------------------------
// stage-level entry points
// ---------------------------------------------------------------------------------------------
extern "C" int emrf_find_splits(emrf_ctx_t* c, const emrf_params_t* params, const double* features_hd, int64_t N, int32_t F,
                                const int64_t* part_off, int32_t n_parts, double* thresholds_out, int32_t* n_splits_out) {
  if (!c || !features_hd || !thresholds_out || !n_splits_out) return fail(EMRF_ERR_ARG, "NULL argument");
  TRY(check_params(params));
  TRY(check_partitions(N, part_off, n_parts));
  std::lock_guard<std::mutex> lk(c->fit_mu);
  CU(cudaSetDevice(c->device));
  DevBuf own_X, d_thr, d_ns;
  const void* dX = nullptr;
  TRY(stage_input(c, features_hd, (size_t)N * F * 8, own_X, &dX, nullptr));
  std::vector<double> h_thr;
  std::vector<int> h_ns;
  TRY(run_find_splits(c, *params, (const double*)dX, N, F, part_off, n_parts, d_thr, d_ns, h_thr, h_ns));
  memcpy(thresholds_out, h_thr.data(), h_thr.size() * 8);
  memcpy(n_splits_out, h_ns.data(), h_ns.size() * 4);
  return EMRF_OK;
}
extern "C" int64_t emrf_find_splits_sample_rows(const emrf_params_t* params, int64_t N, const int64_t* part_off, int32_t n_parts,
                                                int64_t* rows_out, int64_t cap) {
  if (check_params(params) != EMRF_OK || check_partitions(N, part_off, n_parts) != EMRF_OK) return -1;
  std::vector<int64_t> rows;
  if (!hl::find_splits_rows(*params, N, part_off, n_parts, &rows)) {
    for (int64_t i = 0; i < N && i < cap; i++) rows_out[i] = i;
    return N;
  }
  for (size_t i = 0; i < rows.size() && (int64_t)i < cap; i++) rows_out[i] = rows[i];
  return (int64_t)rows.size();
}
extern "C" int emrf_bin(emrf_ctx_t* c, const double* features_hd, int64_t N, int32_t F, const double* thresholds,
                        const int32_t* n_splits, int32_t thr_stride, uint8_t* bins_cm_out, uint8_t* bins_rm_out) {
  if (!c || !features_hd || !thresholds || !n_splits) return fail(EMRF_ERR_ARG, "NULL argument");
  if (N < 1 || F < 1 || thr_stride < 1) return fail(EMRF_ERR_ARG, "bad shape");
  std::lock_guard<std::mutex> lk(c->fit_mu);
  CU(cudaSetDevice(c->device));
  DevBuf own_X, d_thr, d_ns, d_rm, d_cm;
  const void* dX = nullptr;
  TRY(stage_input(c, features_hd, (size_t)N * F * 8, own_X, &dX, nullptr));
  CU(d_thr.alloc((size_t)F * thr_stride * 8));
  CU(d_ns.alloc((size_t)F * 4));
  CU(cudaMemcpyAsync(d_thr.p, thresholds, (size_t)F * thr_stride * 8, cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(d_ns.p, n_splits, (size_t)F * 4, cudaMemcpyHostToDevice, c->stream));
  const int Fp = (int)align_up(F, 16);
  const int64_t Np = (int64_t)align_up((size_t)N, 128);
  CU(d_rm.alloc((size_t)Np * Fp));
  CU(d_cm.alloc((size_t)F * Np));
  TRY(run_bin(c, (const double*)dX, N, F, d_thr.as<double>(), d_ns.as<int>(), thr_stride, thr_stride, Fp, Np, d_rm.as<uint8_t>(),
              d_cm.as<uint8_t>()));
  if (bins_rm_out)
    CU(cudaMemcpy2DAsync(bins_rm_out, (size_t)F, d_rm.p, (size_t)Fp, (size_t)F, (size_t)N, cudaMemcpyDeviceToHost, c->stream));
  if (bins_cm_out)
    CU(cudaMemcpy2DAsync(bins_cm_out, (size_t)N, d_cm.p, (size_t)Np, (size_t)N, (size_t)F, cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return EMRF_OK;
}
extern "C" int emrf_bag_weights(emrf_ctx_t* c, const emrf_params_t* params, int64_t N, const int64_t* part_off, int32_t n_parts,
                                uint8_t* weights_out, int32_t* max_weight_out) {
  if (!c || !weights_out) return fail(EMRF_ERR_ARG, "NULL argument");
  TRY(check_params(params));
  TRY(check_partitions(N, part_off, n_parts));
  std::lock_guard<std::mutex> lk(c->fit_mu);
  CU(cudaSetDevice(c->device));
  DevBuf d_w;
  CU(d_w.alloc((size_t)N * params->num_trees));
  int mw = 0;
  TRY(run_bag(c, *params, N, part_off, n_parts, d_w.as<uint8_t>(), &mw));
  if (max_weight_out) *max_weight_out = mw;
  CU(cudaMemcpyAsync(weights_out, d_w.p, (size_t)N * params->num_trees, cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return EMRF_OK;
}

extern "C" int emrf_histogram(emrf_ctx_t* c, const uint8_t* bins_rm, int64_t N, int32_t F, const uint64_t* entries,
                              const int64_t* node_seg_off, int32_t n_nodes, const int32_t* subsets, int32_t f_sub,
                              const int32_t* num_bins_all, int32_t C, uint32_t* hist_out) {
  if (!c || !bins_rm || !entries || !node_seg_off || !subsets || !num_bins_all || !hist_out) return fail(EMRF_ERR_ARG, "NULL argument");
  if (n_nodes < 1 || f_sub < 1 || C < 1 || C > 256 || F < 1) return fail(EMRF_ERR_ARG, "bad shape");
  std::lock_guard<std::mutex> lk(c->fit_mu);
  CU(cudaSetDevice(c->device));
  cudaStream_t st = c->stream;
  int B = 1;
  for (int f = 0; f < F; f++) B = std::max(B, num_bins_all[f]);
  const int Fp = row_pitch(F);
  int row_ch, k_chunk, n_chunks;
  size_t region, smem;
  bool tma = false;
  TRY(hist_geometry(c, Fp, f_sub, B, C, &row_ch, &k_chunk, &n_chunks, &region, &smem, &tma));
  CU(hist_launch(row_ch, dim3(1), smem, st, kern::HistArgs(), true));
  const int64_t n_ent = node_seg_off[n_nodes];
  DevBuf d_rm, d_ent, d_nodes, d_feat, d_tiles, d_nb, d_hist;
  CU(d_rm.alloc((size_t)N * Fp));
  CU(cudaMemsetAsync(d_rm.p, 0, (size_t)N * Fp, st));
  CU(cudaMemcpy2DAsync(d_rm.p, (size_t)Fp, bins_rm, (size_t)F, (size_t)F, (size_t)N, cudaMemcpyHostToDevice, st));
  CU(d_ent.alloc((size_t)n_ent * 8));
  CU(cudaMemcpyAsync(d_ent.p, entries, (size_t)n_ent * 8, cudaMemcpyHostToDevice, st));
  const uint32_t tile_len = 8192;
  std::vector<kern::GroupNode> gn(n_nodes);
  std::vector<uint32_t> tn;
  for (int i = 0; i < n_nodes; i++) {
    kern::GroupNode& d = gn[i];
    d.seg_start = (uint64_t)node_seg_off[i];
    d.seg_len = (uint32_t)(node_seg_off[i + 1] - node_seg_off[i]);
    d.buf = 0; d.seg_slot = 0; d.slot = i;
    d.tile_begin = (uint32_t)tn.size();
    d.n_tiles = std::max<uint32_t>(1, (d.seg_len + tile_len - 1) / tile_len);
    for (uint32_t q = 0; q < d.n_tiles; q++) tn.push_back((uint32_t)i);
  }
  const size_t slot_words = (size_t)f_sub * B * C;
  CU(d_nodes.alloc(sizeof(kern::GroupNode) * (size_t)n_nodes));
  CU(d_feat.alloc((size_t)n_nodes * f_sub * 4));
  CU(d_tiles.alloc(tn.size() * 4));
  CU(d_nb.alloc((size_t)F * 4));
  CU(d_hist.alloc((size_t)n_nodes * slot_words * 4));
  CU(cudaMemcpyAsync(d_nodes.p, gn.data(), sizeof(kern::GroupNode) * (size_t)n_nodes, cudaMemcpyHostToDevice, st));
  CU(cudaMemcpyAsync(d_feat.p, subsets, (size_t)n_nodes * f_sub * 4, cudaMemcpyHostToDevice, st));
  CU(cudaMemcpyAsync(d_tiles.p, tn.data(), tn.size() * 4, cudaMemcpyHostToDevice, st));
  CU(cudaMemcpyAsync(d_nb.p, num_bins_all, (size_t)F * 4, cudaMemcpyHostToDevice, st));
  CU(cudaMemsetAsync(d_hist.p, 0, (size_t)n_nodes * slot_words * 4, st));
  kern::HistArgs A{};
  memset(&A, 0, sizeof(A));
  A.ent0 = d_ent.as<uint64_t>(); A.ent1 = d_ent.as<uint64_t>();
  A.bins_rm = d_rm.as<uint8_t>();
  A.nodes = d_nodes.as<kern::GroupNode>();
  A.tile_node = d_tiles.as<uint32_t>();
  A.node_feat = d_feat.as<int32_t>();
  A.num_bins_all = d_nb.as<int32_t>();
  A.ghist = d_hist.as<uint32_t>();
  A.Fp = Fp; A.f_sub = f_sub; A.B = B; A.C = C; A.k_chunk = k_chunk; A.n_chunks = n_chunks;
  A.tile_len = tile_len; A.hist_only = 1;
  A.hist_smem_off = (uint32_t)region;
  A.use_tma = tma ? 1 : 0;
  if (tma) TRY(make_rows_tmap(&A.tmap, d_rm.as<uint8_t>(), N, Fp));
  CU(hist_launch(row_ch, dim3((unsigned)tn.size(), n_chunks), smem, st, A, false));
  CU(cudaMemcpyAsync(hist_out, d_hist.p, (size_t)n_nodes * slot_words * 4, cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  return EMRF_OK;
}

extern "C" int emrf_best_splits(emrf_ctx_t* c, const uint32_t* hist, int32_t n_nodes, const int32_t* subsets, int32_t f_sub,
                                const int32_t* num_bins_all, int32_t F, int32_t C, int32_t impurity, int32_t min_instances,
                                double min_info_gain, int32_t* k_out, int32_t* split_out, double* gain_out, int32_t* valid_out,
                                uint32_t* left_out, uint32_t* total_out) {
  if (!c || !hist || !subsets || !num_bins_all) return fail(EMRF_ERR_ARG, "NULL argument");
  if (n_nodes < 1 || f_sub < 1 || C < 1 || C > 256 || F < 1) return fail(EMRF_ERR_ARG, "bad shape");
  std::lock_guard<std::mutex> lk(c->fit_mu);
  CU(cudaSetDevice(c->device));
  cudaStream_t st = c->stream;
  int B = 1;
  for (int f = 0; f < F; f++) B = std::max(B, num_bins_all[f]);
  int row_ch, k_chunk, n_chunks;
  size_t region, smem;
  TRY(hist_geometry(c, 16, f_sub, B, C, &row_ch, &k_chunk, &n_chunks, &region, &smem));
  smem = kern::hist_smem_bytes(k_chunk, B, C);
  CU(cudaFuncSetAttribute(kern::scan_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const size_t slot_words = (size_t)f_sub * B * C;
  const size_t n_rec = (size_t)n_nodes * n_chunks;
  DevBuf d_hist, d_feat, d_nb, d_res, d_left, d_tot;
  CU(d_hist.alloc((size_t)n_nodes * slot_words * 4));
  CU(d_feat.alloc((size_t)n_nodes * f_sub * 4));
  CU(d_nb.alloc((size_t)F * 4));
  CU(d_res.alloc(n_rec * sizeof(kern::ScanResult)));
  CU(d_left.alloc(n_rec * C * 4));
  CU(d_tot.alloc(n_rec * C * 4));
  CU(cudaMemcpyAsync(d_hist.p, hist, (size_t)n_nodes * slot_words * 4, cudaMemcpyHostToDevice, st));
  CU(cudaMemcpyAsync(d_feat.p, subsets, (size_t)n_nodes * f_sub * 4, cudaMemcpyHostToDevice, st));
  CU(cudaMemcpyAsync(d_nb.p, num_bins_all, (size_t)F * 4, cudaMemcpyHostToDevice, st));
  kern::scan_kernel<<<dim3(n_nodes, n_chunks), kern::kHistThreads, smem, st>>>(
      d_hist.as<uint32_t>(), d_feat.as<int32_t>(), d_nb.as<int32_t>(), f_sub, B, C, k_chunk, n_chunks, impurity, min_instances,
      min_info_gain, d_res.as<kern::ScanResult>(), d_left.as<uint32_t>(), d_tot.as<uint32_t>());
  CU(cudaGetLastError());
  std::vector<kern::ScanResult> res(n_rec);
  std::vector<uint32_t> l(n_rec * C), t(n_rec * C);
  CU(cudaMemcpyAsync(res.data(), d_res.p, n_rec * sizeof(kern::ScanResult), cudaMemcpyDeviceToHost, st));
  CU(cudaMemcpyAsync(l.data(), d_left.p, n_rec * C * 4, cudaMemcpyDeviceToHost, st));
  CU(cudaMemcpyAsync(t.data(), d_tot.p, n_rec * C * 4, cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  for (int g = 0; g < n_nodes; g++) {
    int best = -1;
    for (int ch = 0; ch < n_chunks; ch++) {
      const kern::ScanResult& r = res[(size_t)g * n_chunks + ch];
      if (r.k < 0) continue;
      if (best < 0 || r.gain > res[(size_t)g * n_chunks + best].gain) best = ch;
    }
    const size_t rec = (size_t)g * n_chunks + (best < 0 ? 0 : best);
    if (k_out) k_out[g] = best < 0 ? -1 : res[rec].k;
    if (split_out) split_out[g] = best < 0 ? 0 : res[rec].split;
    if (gain_out) gain_out[g] = best < 0 ? kern::kInvalidGain : res[rec].gain;
    if (valid_out) valid_out[g] = best < 0 ? 0 : res[rec].valid;
    if (left_out) memcpy(left_out + (size_t)g * C, &l[rec * C], (size_t)C * 4);
    if (total_out) memcpy(total_out + (size_t)g * C, &t[rec * C], (size_t)C * 4);
  }
  return EMRF_OK;
}

extern "C" int emrf_update_partition(emrf_ctx_t* c, const uint8_t* bins_cm, int64_t N, int32_t F, const uint64_t* entries,
                                     const int64_t* node_seg_off, int32_t n_nodes, const int32_t* split_feature,
                                     const int32_t* split_bin, uint64_t* entries_out, int64_t* n_left_out) {
  if (!c || !bins_cm || !entries || !node_seg_off || !split_feature || !split_bin || !entries_out || !n_left_out)
    return fail(EMRF_ERR_ARG, "NULL argument");
  if (n_nodes < 1 || F < 1 || N < 1) return fail(EMRF_ERR_ARG, "bad shape");
  std::lock_guard<std::mutex> lk(c->fit_mu);
  CU(cudaSetDevice(c->device));
  cudaStream_t st = c->stream;
  const int64_t Np = (int64_t)align_up((size_t)N, 128);
  const int64_t n_ent = node_seg_off[n_nodes];
  DevBuf d_cm, d_in, d_out, d_nodes, d_tiles, d_tl, d_nl;
  CU(d_cm.alloc((size_t)F * Np));
  CU(cudaMemcpy2DAsync(d_cm.p, (size_t)Np, bins_cm, (size_t)N, (size_t)N, (size_t)F, cudaMemcpyHostToDevice, st));
  CU(d_in.alloc((size_t)n_ent * 8));
  CU(d_out.alloc((size_t)n_ent * 8));
  CU(cudaMemcpyAsync(d_in.p, entries, (size_t)n_ent * 8, cudaMemcpyHostToDevice, st));
  CU(cudaMemsetAsync(d_out.p, 0, (size_t)n_ent * 8, st));
  std::vector<kern::SplitNode> sn(n_nodes);
  std::vector<uint32_t> tn;
  for (int i = 0; i < n_nodes; i++) {
    kern::SplitNode& s = sn[i];
    s.seg_start = (uint64_t)node_seg_off[i];
    s.seg_len = (uint32_t)(node_seg_off[i + 1] - node_seg_off[i]);
    s.buf = 0; s.feature = split_feature[i]; s.split_bin = split_bin[i];
    s.seg_slot = 0; s.left_slot = 0; s.right_slot = 0; s.pad = 0;
    if (s.feature < 0 || s.feature >= F) return fail(EMRF_ERR_ARG, "split feature out of range");
    s.tile_begin = (uint32_t)tn.size();
    s.n_tiles = (s.seg_len + kern::kPartTile - 1) / kern::kPartTile;
    for (uint32_t q = 0; q < s.n_tiles; q++) tn.push_back((uint32_t)i);
  }
  CU(d_nodes.alloc(sizeof(kern::SplitNode) * (size_t)n_nodes));
  CU(d_tiles.alloc(tn.size() * 4));
  CU(d_tl.alloc(tn.size() * 4));
  CU(d_nl.alloc((size_t)n_nodes * 8));
  CU(cudaMemcpyAsync(d_nodes.p, sn.data(), sizeof(kern::SplitNode) * (size_t)n_nodes, cudaMemcpyHostToDevice, st));
  CU(cudaMemcpyAsync(d_tiles.p, tn.data(), tn.size() * 4, cudaMemcpyHostToDevice, st));
  CU(cudaMemsetAsync(d_nl.p, 0, (size_t)n_nodes * 8, st));
  if (!tn.empty()) {
    kern::part_count_kernel<<<(unsigned)tn.size(), kern::kPartThreads, 0, st>>>(d_in.as<uint64_t>(), d_in.as<uint64_t>(),
                                                                              d_cm.as<uint8_t>(), Np, d_nodes.as<kern::SplitNode>(),
                                                                              d_tiles.as<uint32_t>(), d_tl.as<uint32_t>(), nullptr);
    kern::part_scan_kernel<<<n_nodes, 256, 0, st>>>(d_nodes.as<kern::SplitNode>(), d_tl.as<uint32_t>(), d_nl.as<uint64_t>(), nullptr);
    kern::part_scatter_kernel<<<(unsigned)tn.size(), kern::kPartThreads, 0, st>>>(
        d_in.as<uint64_t>(), d_in.as<uint64_t>(), d_out.as<uint64_t>(), d_out.as<uint64_t>(), d_cm.as<uint8_t>(), Np,
        d_nodes.as<kern::SplitNode>(), d_tiles.as<uint32_t>(), d_tl.as<uint32_t>(), d_nl.as<uint64_t>(), nullptr);
  }
  CU(cudaGetLastError());
  std::vector<uint64_t> nl(n_nodes);
  CU(cudaMemcpyAsync(entries_out, d_out.p, (size_t)n_ent * 8, cudaMemcpyDeviceToHost, st));
  CU(cudaMemcpyAsync(nl.data(), d_nl.p, (size_t)n_nodes * 8, cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  for (int i = 0; i < n_nodes; i++) n_left_out[i] = (int64_t)nl[i];
  return EMRF_OK;
}

// ---------------------------------------------------------------------------------------------
// host-logic probes (no GPU)
// ---------------------------------------------------------------------------------------------
extern "C" int64_t emrf_host_java_random_next_long(int64_t seed, int32_t nth) {
  hl::JavaUtilRandom r(seed);
  int64_t v = 0;
  for (int i = 0; i <= nth; i++) v = r.next_long();
  return v;
}
extern "C" double emrf_host_xorshift_next_double(int64_t seed, int32_t nth) {
  hl::SparkXorShift r(seed);
  double v = 0;
  for (int i = 0; i <= nth; i++) v = r.next_double();
  return v;
}
extern "C" int32_t emrf_host_reservoir_sample(int32_t n_items, int32_t k, int64_t seed, int32_t* out) {
  int len = 0;
  hl::reservoir_features(n_items, k, seed, out, &len);
  return len;
}
extern "C" int32_t emrf_host_group_tree_order(const int32_t* trees_first_seen, int32_t n, int32_t* out) {
  std::vector<int> in(trees_first_seen, trees_first_seen + n);
  std::vector<int> o = hl::group_tree_visit_order(in);
  for (size_t i = 0; i < o.size(); i++) out[i] = o[i];
  return (int32_t)o.size();
}
extern "C" int32_t emrf_host_packed_slot_layout(const uint32_t* weight, int32_t n, int32_t world, int32_t* slot_out,
                                                int32_t* owner_out) {
  if (!weight || !slot_out || !owner_out || n < 0 || world < 1) return -1;
  return hl::packed_slot_layout(weight, n, world, slot_out, owner_out);
}
extern "C" int64_t emrf_host_group_tiles_bound(const uint32_t* ub, const int32_t* tree, int32_t n, const uint64_t* tree_total,
                                               int32_t n_trees, uint32_t tile_len) {
  if (!ub || !tree || !tree_total || n < 0 || n_trees < 1 || tile_len < 1) return -1;
  for (int i = 0; i < n; i++)
    if (tree[i] < 0 || tree[i] >= n_trees) return -1;
  std::vector<uint64_t> acc((size_t)n_trees, 0);
  return (int64_t)hl::tiles_bound(hl::group_entries_bound(ub, tree, n, tree_total, acc.data()), tile_len, n);
}
extern "C" int32_t emrf_host_num_features_per_node(const emrf_params_t* params, int32_t n_features) {
  return hl::features_per_node(*params, n_features);
}

// ---------------------------------------------------------------------------------------------
// emrf_update_partition_noalloc — the partition update for callers that must not allocate.
//
// Some hosts of this library cannot let it call cudaMalloc: a Spark executor that has handed the
// whole device to a pooling allocator at start-up (nothing is left for a plain cudaMalloc to
// find), or a serving loop that cannot afford the device-wide synchronisation cudaMalloc/cudaFree
// imply.  For them this variant keeps everything the three partition kernels need besides the
// caller's own inputs and outputs — node descriptors, the tile→node map and the per-tile left
// counts — in ONE file-scope device array of fixed size.  Every GPU that loads libemrf.so holds
// those kNoAllocScratchBytes until the process exits.
//
// The array's name below is only a host-side handle; the device address the kernels need is
// looked up with cudaGetSymbolAddress the first time the entry point runs on a device (with that
// device current).  Make that first call outside any stream capture.  All calls share the one
// array and it keeps whatever the previous call left there: this call overwrites every word it
// later reads, and clears the caller's n_left output itself when there is nothing to partition.
// Two calls that overlap on the GPU (different streams, contexts or threads) would race on the
// array — serialise them.
// ---------------------------------------------------------------------------------------------
constexpr size_t kNoAllocMaxNodes = 4096;
constexpr size_t kNoAllocMaxTiles = 262144;   // 262144 tiles x 2048 entries = 536 870 912 entries
constexpr size_t kNoAllocNodesOff = 0;
constexpr size_t kNoAllocTileNodeOff = kNoAllocMaxNodes * sizeof(kern::SplitNode);
constexpr size_t kNoAllocTileLeftOff = kNoAllocTileNodeOff + kNoAllocMaxTiles * 4;
constexpr size_t kNoAllocScratchBytes = kNoAllocTileLeftOff + kNoAllocMaxTiles * 4;  // 2 293 760 B
static_assert(kNoAllocScratchBytes == 2293760, "documented in include/emrf.h, DESIGN.md and INTEGRATION.md");
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocScratchBytes];

static std::mutex g_noalloc_mu;
static void* g_noalloc_ptr[64] = {nullptr};   // per device ordinal

extern "C" int emrf_update_partition_noalloc(emrf_ctx_t* c, const uint8_t* bins_cm_dev, int64_t col_stride,
                                             int32_t n_features, const uint64_t* entries_dev,
                                             const int64_t* node_seg_off, int32_t n_nodes, const int32_t* split_feature,
                                             const int32_t* split_bin, uint64_t* entries_out_dev, uint64_t* n_left_dev) {
  if (!c || !bins_cm_dev || !entries_dev || !node_seg_off || !split_feature || !split_bin || !entries_out_dev || !n_left_dev)
    return fail(EMRF_ERR_ARG, "NULL argument");
  if (n_nodes < 1 || n_features < 1 || col_stride < 1) return fail(EMRF_ERR_ARG, "bad shape");
  if ((size_t)n_nodes > kNoAllocMaxNodes)
    return fail(EMRF_ERR_LIMIT, "emrf_update_partition_noalloc handles at most %zu nodes per call (got %d)", kNoAllocMaxNodes, n_nodes);
  if (c->device < 0 || c->device >= 64) return fail(EMRF_ERR_LIMIT, "device ordinal %d not supported", c->device);
  if (node_seg_off[0] < 0) return fail(EMRF_ERR_ARG, "node_seg_off[0] is negative");
  return guarded(c, [&]() -> int {       // no exception (std::bad_alloc of the descriptor vectors) may cross the C boundary
  // host-side descriptors first, so that the size check happens before anything is enqueued
  std::vector<kern::SplitNode> sn(n_nodes);
  size_t n_tiles = 0;
  for (int i = 0; i < n_nodes; i++) {
    const int64_t len = node_seg_off[i + 1] - node_seg_off[i];
    if (len < 0 || len >= (1ll << 32)) return fail(EMRF_ERR_ARG, "node %d: bad segment", i);
    if (split_feature[i] < 0 || split_feature[i] >= n_features) return fail(EMRF_ERR_ARG, "split feature out of range");
    kern::SplitNode& s = sn[i];
    s.seg_start = (uint64_t)node_seg_off[i];
    s.seg_len = (uint32_t)len;
    s.buf = 0;
    s.feature = split_feature[i];
    s.split_bin = split_bin[i];
    s.seg_slot = 0; s.left_slot = 0; s.right_slot = 0; s.pad = 0;
    s.tile_begin = (uint32_t)n_tiles;
    s.n_tiles = (uint32_t)((len + kern::kPartTile - 1) / kern::kPartTile);
    n_tiles += s.n_tiles;
  }
  if (n_tiles > kNoAllocMaxTiles)
    return fail(EMRF_ERR_LIMIT, "emrf_update_partition_noalloc handles at most %zu entries per call (%zu tiles of %d; got %zu tiles)",
                kNoAllocMaxTiles * (size_t)kern::kPartTile, kNoAllocMaxTiles, kern::kPartTile, n_tiles);
  // the kernels read entries_dev while they write entries_out_dev: the two must be distinct device buffers (checked with
  // cudaPointerGetAttributes, which allocates nothing)
  if (!is_device_ptr(bins_cm_dev) || !is_device_ptr(entries_dev) || !is_device_ptr(entries_out_dev) || !is_device_ptr(n_left_dev))
    return fail(EMRF_ERR_ARG, "emrf_update_partition_noalloc takes device pointers for bins_cm, entries, entries_out and n_left");
  {
    const uint64_t n_ent = (uint64_t)node_seg_off[n_nodes];
    const uintptr_t a0 = (uintptr_t)entries_dev, b0 = (uintptr_t)entries_out_dev;
    if (a0 == b0 || (a0 < b0 + n_ent * 8 && b0 < a0 + n_ent * 8))
      return fail(EMRF_ERR_ARG, "entries_out_dev must not alias entries_dev (the partition is not done in place)");
  }
  std::vector<uint32_t> tn(n_tiles);
  for (int i = 0; i < n_nodes; i++)
    for (uint32_t q = 0; q < sn[i].n_tiles; q++) tn[sn[i].tile_begin + q] = (uint32_t)i;

  std::lock_guard<std::mutex> lk(c->fit_mu);
  CU(cudaSetDevice(c->device));
  unsigned char* scratch = nullptr;
  {
    std::lock_guard<std::mutex> lk2(g_noalloc_mu);
    if (!g_noalloc_ptr[c->device]) {  // once per device, with that device current
      void* p = nullptr;
      CU(cudaGetSymbolAddress(&p, g_noalloc_scratch));
      g_noalloc_ptr[c->device] = p;
    }
    scratch = static_cast<unsigned char*>(g_noalloc_ptr[c->device]);
  }
  cudaStream_t st = c->stream;
  kern::SplitNode* d_nodes = reinterpret_cast<kern::SplitNode*>(scratch + kNoAllocNodesOff);
  uint32_t* d_tile_node = reinterpret_cast<uint32_t*>(scratch + kNoAllocTileNodeOff);
  uint32_t* d_tile_left = reinterpret_cast<uint32_t*>(scratch + kNoAllocTileLeftOff);
  CU(cudaMemcpyAsync(d_nodes, sn.data(), sizeof(kern::SplitNode) * (size_t)n_nodes, cudaMemcpyHostToDevice, st));
  if (n_tiles) {
    CU(cudaMemcpyAsync(d_tile_node, tn.data(), n_tiles * 4, cudaMemcpyHostToDevice, st));
    // part_count writes tile_left[0..n_tiles), part_scan turns it into offsets and writes n_left for
    // every node, part_scatter reads both: no word of the scratch is read before this call wrote it.
    kern::part_count_kernel<<<(unsigned)n_tiles, kern::kPartThreads, 0, st>>>(entries_dev, entries_dev, bins_cm_dev, col_stride,
                                                                          d_nodes, d_tile_node, d_tile_left, nullptr);
    kern::part_scan_kernel<<<n_nodes, 256, 0, st>>>(d_nodes, d_tile_left, n_left_dev, nullptr);
    kern::part_scatter_kernel<<<(unsigned)n_tiles, kern::kPartThreads, 0, st>>>(entries_dev, entries_dev, entries_out_dev,
                                                                            entries_out_dev, bins_cm_dev, col_stride, d_nodes,
                                                                            d_tile_node, d_tile_left, n_left_dev, nullptr);
    CU(cudaGetLastError());
  } else {
    CU(cudaMemsetAsync(n_left_dev, 0, (size_t)n_nodes * 8, st));  // every segment is empty
  }
  CU(cudaStreamSynchronize(st));  // sn / tn are pageable host memory read by the async copies
  return EMRF_OK;
  });
}

// ---------------------------------------------------------------------------------------------
// emrf_build_lag_features — draw table -> lag-window feature matrix + label column, on the device
// ---------------------------------------------------------------------------------------------
extern "C" int emrf_build_lag_features(emrf_ctx_t* c, const int32_t* draws, int32_t n_draws, int32_t lag, int32_t ball,
                                       double* features_hd, double* labels_hd) {
  if (!c || !draws || !features_hd) return fail(EMRF_ERR_ARG, "NULL argument");
  if (lag < 1 || n_draws <= lag || ball < 0 || ball > 6) return fail(EMRF_ERR_ARG, "need n_draws > lag >= 1 and ball in 0..6");
  std::lock_guard<std::mutex> lk(c->fit_mu);
  CU(cudaSetDevice(c->device));
  cudaStream_t st = c->stream;
  const int64_t n_rows = n_draws - lag, F = 7 * (int64_t)lag;
  DevBuf d_draws, d_X, d_y;
  CU(d_draws.alloc((size_t)n_draws * 7 * 4));
  CU(cudaMemcpyAsync(d_draws.p, draws, (size_t)n_draws * 7 * 4, cudaMemcpyHostToDevice, st));
  const bool x_dev = is_device_ptr(features_hd), y_dev = labels_hd && is_device_ptr(labels_hd);
  double* dX = features_hd;
  double* dy = labels_hd;
  if (!x_dev) { CU(d_X.alloc((size_t)n_rows * F * 8)); dX = d_X.as<double>(); }
  if (labels_hd && !y_dev) { CU(d_y.alloc((size_t)n_rows * 8)); dy = d_y.as<double>(); }
  kern::lag_features_kernel<<<(unsigned)((n_rows * F + 255) / 256), 256, 0, st>>>(d_draws.as<int32_t>(), n_draws, lag, ball, dX, dy);
  CU(cudaGetLastError());
  if (!x_dev) CU(cudaMemcpyAsync(features_hd, dX, (size_t)n_rows * F * 8, cudaMemcpyDeviceToHost, st));
  if (labels_hd && !y_dev) CU(cudaMemcpyAsync(labels_hd, dy, (size_t)n_rows * 8, cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  return EMRF_OK;
}

extern "C" int emrf_synth_flat(emrf_ctx_t* c, int64_t n_rows, int32_t n_features, int32_t n_classes, int64_t seed, int64_t row_start,
                               double* features_dev, double* labels_dev) {
  if (!c || !features_dev) return fail(EMRF_ERR_ARG, "NULL argument");
  if (n_rows < 0 || n_features < 1 || n_classes < 1 || row_start < 0) return fail(EMRF_ERR_ARG, "bad shape");
  if (!is_device_ptr(features_dev) || (labels_dev && !is_device_ptr(labels_dev))) return fail(EMRF_ERR_ARG, "outputs must be device buffers");
  return guarded(c, [&]() -> int {
    CU(cudaSetDevice(c->device));
    if (n_rows == 0) return EMRF_OK;
    const int64_t total = n_rows * n_features;
    const unsigned grid = (unsigned)std::min<int64_t>((total + 255) / 256, (int64_t)c->sm_count * 32);
    kern::synth_flat_kernel<<<grid, 256, 0, c->stream>>>(n_rows, n_features, n_classes, (uint64_t)seed, row_start, features_dev, labels_dev);
    CU(cudaGetLastError());
    CU(cudaStreamSynchronize(c->stream));
    return EMRF_OK;
  });
}

#ifdef EMRF_HIST_PROF
// debug builds only (make prof): read and clear the histogram kernel's phase counters
extern "C" int emrf_debug_hist_prof(unsigned long long* out16) {
  CU(cudaDeviceSynchronize());
  CU(cudaMemcpyFromSymbol(out16, kern::g_hist_prof, sizeof(unsigned long long) * 16));
  unsigned long long z[16] = {0};
  CU(cudaMemcpyToSymbol(kern::g_hist_prof, z, sizeof(z)));
  return EMRF_OK;
}
extern "C" int emrf_debug_bag_prof(unsigned long long* out8) {
  CU(cudaDeviceSynchronize());
  CU(cudaMemcpyFromSymbol(out8, kern::g_bag_prof, sizeof(unsigned long long) * 8));
  unsigned long long z[8] = {0};
  CU(cudaMemcpyToSymbol(kern::g_bag_prof, z, sizeof(z)));
  return EMRF_OK;
}
#endif

