// Shared helpers for the exchange kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include "../../include/xchg_b200.h"

void xchg_set_error(const char* fmt, ...);

#define XCHG_CUDA_TRY(expr)                                                        \
  do {                                                                             \
    cudaError_t e__ = (expr);                                                      \
    if (e__ != cudaSuccess) {                                                      \
      xchg_set_error("%s: %s (%s:%d)", #expr, cudaGetErrorString(e__), __FILE__,   \
                     __LINE__);                                                    \
      return XCHG_ERR_CUDA;                                                        \
    }                                                                              \
  } while (0)

#define XCHG_LAUNCH_CHECK() XCHG_CUDA_TRY(cudaGetLastError())

#define XCHG_REQUIRE(cond, ...)            \
  do {                                     \
    if (!(cond)) {                         \
      xchg_set_error(__VA_ARGS__);         \
      return XCHG_ERR_INVALID;             \
    }                                      \
  } while (0)

// Runs the statements the first time the enclosing call site is reached with each CUDA device
// current (function attributes such as the dynamic shared-memory limit are per device /
// context, not per process).  Thread-safe: two host threads may both run the body, which
// is idempotent; a body that fails (XCHG_CUDA_TRY returns) leaves the device unmarked.
#define XCHG_ONCE_PER_DEVICE(...)                                             \
  do {                                                                        \
    static std::atomic<uint64_t> done_mask__{0};                              \
    int dev__ = 0;                                                            \
    XCHG_CUDA_TRY(cudaGetDevice(&dev__));                                     \
    const uint64_t bit__ = 1ull << (dev__ & 63);                              \
    if (!(done_mask__.load(std::memory_order_acquire) & bit__)) {             \
      __VA_ARGS__;                                                            \
      done_mask__.fetch_or(bit__, std::memory_order_release);                 \
    }                                                                         \
  } while (0)

static inline uint64_t xchg_align_up(uint64_t x, uint64_t a) { return (x + a - 1) / a * a; }

// Number of SMs of the current device (cached per process; B200 = 148).
int xchg_sm_count();
// CTAs per SM the gather kernels may use (0 = each kernel's own default); see
// xchg_set_gather_ctas_per_sm.
int xchg_gather_ctas_per_sm();
// Staging bytes per CTA (KB) of the uniform-row gather for the launches that follow (0 = the
// default); see xchg_set_gather_stage_kb.
int xchg_gather_stage_kb();

__device__ __forceinline__ bool bit_is_set(const uint8_t* bm, int64_t i) {
  return (bm[i >> 3] >> (i & 7)) & 1;
}

__device__ __forceinline__ unsigned lanemask_lt() {
  unsigned m;
  asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
  return m;
}

// Streaming (read-once) loads / stores: keep them out of L1.
__device__ __forceinline__ uint4 ld_stream_u4(const uint4* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ void st_stream_u4(uint4* p, const uint4& v) {
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x),
               "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}

// Random-access loads: ask the L2 for a 64-byte fill instead of the default whole
// 128-byte line (measured on B200: a random 32-byte read costs 127 B of HBM traffic by
// default, 64 B with .L2::64B -- profiles/r01_fetch_granularity.md).
template <typename T>
__device__ __forceinline__ T ld_random(const T* p) {
  static_assert(sizeof(T) == 1 || sizeof(T) == 2 || sizeof(T) == 4 || sizeof(T) == 8 || sizeof(T) == 16,
                "unsupported width");
  if constexpr (sizeof(T) == 1) {
    uint32_t v;
    asm volatile("ld.global.L2::64B.u8 %0, [%1];" : "=r"(v) : "l"(p));
    uint8_t b = (uint8_t)v;
    return *reinterpret_cast<T*>(&b);
  } else if constexpr (sizeof(T) == 2) {
    uint16_t v;
    asm volatile("ld.global.L2::64B.u16 %0, [%1];" : "=h"(v) : "l"(p));
    return *reinterpret_cast<T*>(&v);
  } else if constexpr (sizeof(T) == 4) {
    uint32_t v;
    asm volatile("ld.global.L2::64B.u32 %0, [%1];" : "=r"(v) : "l"(p));
    return *reinterpret_cast<T*>(&v);
  } else if constexpr (sizeof(T) == 8) {
    uint64_t v;
    asm volatile("ld.global.L2::64B.u64 %0, [%1];" : "=l"(v) : "l"(p));
    return *reinterpret_cast<T*>(&v);
  } else {
    uint4 v;
    asm volatile("ld.global.L2::64B.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "l"(p));
    return *reinterpret_cast<T*>(&v);
  }
}

// ---- measurement hooks (api.cu) ----------------------------------------------
// Every kernel launch is counted; when profiling is enabled a launch is bracketed
// by CUDA events on its own stream so bench.py can report per-kernel device time.
enum XchgKernelId {
  XK_NORMALIZE = 0,
  XK_SCAN_HIST = 1,
  XK_ONESWEEP = 2,
  XK_GATHER_ROWS = 3,
  XK_GATHER_FIXED = 4,
  XK_GATHER_BITMAP = 5,
  XK_GATHER_OFFSETS = 6,
  XK_PARTITION = 7,
  XK_OFFSETS_UTIL = 8,
  XK_MISC = 9,
  XK_MERGE = 10,
  XK_SHUFFLE = 11,
  XK_AGGREGATE = 12,
  XK_LOCAL_SORT = 13,
  XK_MSD_AUX = 14,
  XK_COUNT = 15
};

void xchg_prof_begin(int kernel_id, cudaStream_t s);
void xchg_prof_end(int kernel_id, cudaStream_t s);

struct XchgProfScope {
  int id;
  cudaStream_t s;
  XchgProfScope(int id_, cudaStream_t s_) : id(id_), s(s_) { xchg_prof_begin(id, s); }
  ~XchgProfScope() { xchg_prof_end(id, s); }
};

#define XCHG_LAUNCH(id, stream, ...)      \
  do {                                    \
    XchgProfScope prof__((id), (stream)); \
    __VA_ARGS__;                          \
  } while (0)
