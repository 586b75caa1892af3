// Shared helpers for the exchange kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
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

static inline uint64_t xchg_align_up(uint64_t x, uint64_t a) { return (x + a - 1) / a * a; }

// Number of SMs of the current device (cached per process; B200 = 148).
int xchg_sm_count();

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
  XK_COUNT = 13
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
