// C-ABI plumbing: thread-local error message, device queries.
#include <cstdarg>
#include <atomic>
#include <cstdio>
#include <mutex>
#include <vector>

#include "common.cuh"

static thread_local char g_err[512] = "";

void xchg_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int xchg_sm_count() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

extern "C" int xchg_abi_version(void) { return XCHG_ABI_VERSION; }
extern "C" const char* xchg_last_error(void) { return g_err; }
extern "C" int xchg_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}

// ---- measurement hooks ---------------------------------------------------------
namespace {
std::atomic<uint64_t> g_launches{0};
std::atomic<int> g_prof_on{0};
std::mutex g_prof_mu;
struct Pending {
  int id;
  cudaEvent_t a, b;
};
std::vector<Pending> g_pending;
std::vector<cudaEvent_t> g_free_events;
double g_ms[XK_COUNT] = {0};
uint64_t g_cnt[XK_COUNT] = {0};
thread_local cudaEvent_t t_open = nullptr;
constexpr size_t MAX_PENDING = 1 << 16;

cudaEvent_t get_event() {
  if (!g_free_events.empty()) {
    cudaEvent_t e = g_free_events.back();
    g_free_events.pop_back();
    return e;
  }
  cudaEvent_t e = nullptr;
  if (cudaEventCreate(&e) != cudaSuccess) return nullptr;
  return e;
}

void drain_locked() {
  for (auto& p : g_pending) {
    float ms = 0.f;
    if (cudaEventSynchronize(p.b) == cudaSuccess && cudaEventElapsedTime(&ms, p.a, p.b) == cudaSuccess) {
      g_ms[p.id] += ms;
      g_cnt[p.id] += 1;
    }
    g_free_events.push_back(p.a);
    g_free_events.push_back(p.b);
  }
  g_pending.clear();
}
}  // namespace

void xchg_prof_begin(int kernel_id, cudaStream_t s) {
  g_launches.fetch_add(1, std::memory_order_relaxed);
  t_open = nullptr;
  if (!g_prof_on.load(std::memory_order_relaxed)) return;
  std::lock_guard<std::mutex> lk(g_prof_mu);
  if (g_pending.size() >= MAX_PENDING) return;
  cudaEvent_t e = get_event();
  if (!e) return;
  cudaEventRecord(e, s);
  t_open = e;
}

void xchg_prof_end(int kernel_id, cudaStream_t s) {
  if (!t_open) return;
  std::lock_guard<std::mutex> lk(g_prof_mu);
  cudaEvent_t e = get_event();
  if (!e) {
    g_free_events.push_back(t_open);
    t_open = nullptr;
    return;
  }
  cudaEventRecord(e, s);
  g_pending.push_back({kernel_id, t_open, e});
  t_open = nullptr;
}

extern "C" uint64_t xchg_launch_count(void) { return g_launches.load(); }

extern "C" int xchg_profile_enable(int on) {
  std::lock_guard<std::mutex> lk(g_prof_mu);
  if (on) {
    drain_locked();
    for (int i = 0; i < XK_COUNT; ++i) {
      g_ms[i] = 0;
      g_cnt[i] = 0;
    }
  }
  g_prof_on.store(on ? 1 : 0);
  return XCHG_OK;
}

extern "C" int xchg_profile_read(int kernel_id, double* total_ms, uint64_t* launches) {
  XCHG_REQUIRE(kernel_id >= 0 && kernel_id < XK_COUNT, "bad kernel id %d", kernel_id);
  XCHG_REQUIRE(total_ms && launches, "NULL out pointer");
  std::lock_guard<std::mutex> lk(g_prof_mu);
  drain_locked();
  *total_ms = g_ms[kernel_id];
  *launches = g_cnt[kernel_id];
  return XCHG_OK;
}

extern "C" int xchg_set_l2_fetch_granularity(int bytes) {
  XCHG_REQUIRE(bytes == 32 || bytes == 64 || bytes == 128, "granularity must be 32, 64 or 128");
  XCHG_CUDA_TRY(cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, (size_t)bytes));
  return XCHG_OK;
}

// ---- gather grid cap -------------------------------------------------------------
namespace {
std::atomic<int> g_gather_ctas_per_sm{0};
}
int xchg_gather_ctas_per_sm() { return g_gather_ctas_per_sm.load(std::memory_order_relaxed); }

namespace {
std::atomic<int> g_gather_stage_kb{0};
}
int xchg_gather_stage_kb() { return g_gather_stage_kb.load(std::memory_order_relaxed); }

extern "C" int xchg_set_gather_stage_kb(int kb) {
  XCHG_REQUIRE(kb == 0 || (kb >= 4 && kb <= 56), "kb must be 0 or in [4, 56]");
  g_gather_stage_kb.store(kb);
  return XCHG_OK;
}

extern "C" int xchg_set_gather_ctas_per_sm(int ctas_per_sm) {
  XCHG_REQUIRE(ctas_per_sm >= 0 && ctas_per_sm <= 32, "ctas_per_sm must be in [0, 32]");
  g_gather_ctas_per_sm.store(ctas_per_sm);
  return XCHG_OK;
}

// ---- caller-owned host memory ---------------------------------------------------------
extern "C" int xchg_host_register(const void* ptr, uint64_t bytes, int* already_out) {
  XCHG_REQUIRE(ptr != nullptr && bytes > 0, "NULL / empty host range");
  if (already_out) *already_out = 0;
  cudaError_t e = cudaHostRegister(const_cast<void*>(ptr), (size_t)bytes, cudaHostRegisterDefault);
  if (e == cudaErrorHostMemoryAlreadyRegistered) {
    cudaGetLastError();  // not an error for the caller: clear it
    if (already_out) *already_out = 1;
    return XCHG_OK;
  }
  if (e != cudaSuccess) {
    cudaGetLastError();
    // read-only mappings (a sealed shared-memory object, an mmap'ed file) need the flag
    e = cudaHostRegister(const_cast<void*>(ptr), (size_t)bytes, cudaHostRegisterReadOnly);
  }
  if (e != cudaSuccess) {
    cudaGetLastError();
    xchg_set_error("cudaHostRegister(%p, %llu): %s", ptr, (unsigned long long)bytes, cudaGetErrorString(e));
    return XCHG_ERR_CUDA;
  }
  return XCHG_OK;
}

extern "C" int xchg_host_unregister(const void* ptr) {
  XCHG_REQUIRE(ptr != nullptr, "NULL host pointer");
  XCHG_CUDA_TRY(cudaHostUnregister(const_cast<void*>(ptr)));
  return XCHG_OK;
}
