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
