// K2 as device functions: raw key value -> order-preserving unsigned 64-bit radix key
// (shared by csrc/radix_sort.cu and csrc/exchange.cu), and the destination rule of the
// multi-GPU exchange's map side.
#pragma once
#include <stdint.h>

#include <type_traits>

#include "common.cuh"

template <typename T>
__device__ __forceinline__ uint64_t norm_value(T x, bool desc) {
  if constexpr (std::is_same<T, float>::value) {
    if (x != x) return 0xFFFFFFFEull;  // NaN after every number, both directions
    uint32_t b = __float_as_uint(x);
    if (x == 0.0f) b = 0;  // -0.0 ties with +0.0 (Arrow compares numerically)
    uint32_t t = (b & 0x80000000u) ? ~b : (b | 0x80000000u);
    if (desc) t = ~t;
    return (uint64_t)t;
  } else if constexpr (std::is_same<T, double>::value) {
    if (x != x) return 0xFFFFFFFFFFFFFFFEull;
    uint64_t b = (uint64_t)__double_as_longlong(x);
    if (x == 0.0) b = 0;
    uint64_t t = (b >> 63) ? ~b : (b | 0x8000000000000000ull);
    if (desc) t = ~t;
    return t;
  } else {
    using U = typename std::make_unsigned<T>::type;
    U u = (U)x;
    if constexpr (std::is_signed<T>::value) u ^= (U)((U)1 << (sizeof(T) * 8 - 1));
    if (desc) u = (U)~u;
    return (uint64_t)u;
  }
}

// Key of a null row: one past the largest key a value can take (nulls last in both
// directions).  64-bit integers have no spare value: they use UINT64_MAX and the
// sort appends a 9th, validity-driven pass.
template <typename T>
__device__ __forceinline__ uint64_t null_key() {
  if constexpr (std::is_same<T, float>::value) return 0xFFFFFFFFull;
  else if constexpr (sizeof(T) == 8) return 0xFFFFFFFFFFFFFFFFull;
  else return 1ull << (sizeof(T) * 8);
}

// Calls f(XchgTypeTag<T>{}) with T = the C type of key_type; false if it is unknown.
template <typename T>
struct XchgTypeTag {
  using type = T;
};
template <typename F>
inline bool xchg_dispatch_key_type(int key_type, F&& f) {
  switch (key_type) {
    case XCHG_I8: f(XchgTypeTag<int8_t>{}); return true;
    case XCHG_I16: f(XchgTypeTag<int16_t>{}); return true;
    case XCHG_I32: f(XchgTypeTag<int32_t>{}); return true;
    case XCHG_I64: f(XchgTypeTag<int64_t>{}); return true;
    case XCHG_U8: f(XchgTypeTag<uint8_t>{}); return true;
    case XCHG_U16: f(XchgTypeTag<uint16_t>{}); return true;
    case XCHG_U32: f(XchgTypeTag<uint32_t>{}); return true;
    case XCHG_U64: f(XchgTypeTag<uint64_t>{}); return true;
    case XCHG_F32: f(XchgTypeTag<float>{}); return true;
    case XCHG_F64: f(XchgTypeTag<double>{}); return true;
    default: return false;
  }
}

// Destination of a row under <= 255 thresholds in sort order: the number of thresholds the
// row is NOT before, i.e. the range partition find_partitions
// (python/ray/data/_internal/util.py:806-828) slices it into.  A threshold t counts for
// radix key k when t < k, or t == k and the threshold is not strict.  Ascending keys: a
// boundary value goes right (never strict).  Descending keys: a boundary the column can
// hold goes left -- with the larger values, util.py:803 returns the right edge -- so it is
// strict; a boundary between two representable values is stated by its floor and is not.
constexpr int XCHG_MAX_THRESHOLDS = 255;
struct DestRule {
  uint32_t pick[XCHG_MAX_THRESHOLDS];   // threshold i = bounds[pick[i]]
  uint8_t strict[XCHG_MAX_THRESHOLDS];
  int nthr;
};

__device__ __forceinline__ uint32_t dest_of(uint64_t k, const uint64_t* __restrict__ thr,
                                            const uint8_t* __restrict__ strict, int nthr) {
  int lo = 0, hi = nthr;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    const uint64_t t = thr[mid];
    if (t < k || (t == k && !strict[mid])) lo = mid + 1;
    else hi = mid;
  }
  return (uint32_t)lo;
}
