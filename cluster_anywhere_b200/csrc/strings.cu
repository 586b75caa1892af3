// String (binary) sort keys for sm_100a: order-preserving radix keys of a variable-width
// column, and the boundary rules on a sorted / unsorted string column.
//
// Replaces, for string / binary key columns, pyarrow.compute.sort_indices
// (python/ray/data/_internal/arrow_ops/transform_pyarrow.py:41) -- Arrow orders strings
// bytewise (unsigned), a proper prefix before the longer string, nulls last in both
// directions -- and find_partition_index (python/ray/data/_internal/util.py:728-803) on an
// object column, where nulls are replaced by NULL_SENTINEL (:765-772).
//
// K2 for strings: a string is cut into 8-byte chunks, chunk c of row i = bytes
// [8c, 8c+8) as a big-endian u64, zero padded; (chunk 0, chunk 1, ..., length) compared
// lexicographically is exactly the bytewise order with "prefix first".  The sort is LSD
// over those columns: length first, then the chunks from the last to the first, every
// one a stable K3 sort (xchg_sort_column on u64 / u32 keys, row ids chained through
// perm_in); digit places that are constant over the block are skipped as usual, so the
// cost follows the entropy of the keys, not their maximum length.
// Algorithmic bytes per chunk: the string bytes once + 8 B key written per row.
#include <algorithm>

#include "common.cuh"

namespace {

constexpr int ST_THREADS = 256;

int st_grid(uint64_t n) {
  return (int)std::max<uint64_t>(1, std::min<uint64_t>((n + ST_THREADS - 1) / ST_THREADS,
                                                       (uint64_t)xchg_sm_count() * 8));
}

__global__ void __launch_bounds__(ST_THREADS)
k_string_lengths(const int64_t* __restrict__ off, uint64_t n, uint32_t* __restrict__ len_out,
                 unsigned long long* max_len) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  unsigned long long mx = 0;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const uint64_t l = (uint64_t)(off[i + 1] - off[i]);
    len_out[i] = l > 0xFFFFFFFFull ? 0xFFFFFFFFu : (uint32_t)l;
    mx = l > mx ? l : mx;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const unsigned long long y = __shfl_xor_sync(0xffffffffu, mx, o);
    mx = y > mx ? y : mx;
  }
  if ((threadIdx.x & 31) == 0 && mx) atomicMax(max_len, mx);
}

__global__ void __launch_bounds__(ST_THREADS)
k_string_chunk_keys(const int64_t* __restrict__ off, const uint8_t* __restrict__ data, uint64_t n,
                    uint32_t chunk, uint64_t* __restrict__ out) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const int64_t b = off[i], e = off[i + 1];
    const int64_t p = b + (int64_t)chunk * 8;
    uint64_t k = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const uint64_t byte = (p + j < e) ? (uint64_t)data[p + j] : 0ull;
      k = (k << 8) | byte;
    }
    out[i] = k;
  }
}

// bytewise compare of row r with boundary b: -1 row < b, 0 equal, +1 row > b
__device__ __forceinline__ int cmp_bytes(const uint8_t* __restrict__ a, int64_t la,
                                         const uint8_t* __restrict__ b, int64_t lb) {
  const int64_t m = la < lb ? la : lb;
  for (int64_t i = 0; i < m; ++i) {
    const int x = a[i], y = b[i];
    if (x != y) return x < y ? -1 : 1;
  }
  return la < lb ? -1 : (la > lb ? 1 : 0);
}

struct StrCol {
  const int64_t* off;
  const uint8_t* data;
  const uint8_t* validity;  // may be NULL
};

__device__ __forceinline__ bool row_is_null(const StrCol& c, uint64_t i) {
  return c.validity && !bit_is_set(c.validity, (int64_t)i);
}

// cut index of every boundary on a SORTED string column (nulls last)
__global__ void k_partition_bounds_bytes(StrCol col, uint64_t n, const int64_t* __restrict__ boff,
                                         const uint8_t* __restrict__ bdata,
                                         const uint8_t* __restrict__ bflags, int nb, int desc,
                                         uint64_t* __restrict__ cuts) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= nb) return;
  const bool key_none = bflags && (bflags[b] & 4);
  const uint8_t* kb = bdata + boff[b];
  const int64_t kl = boff[b + 1] - boff[b];
  if (!desc) {
    // util.py:782-787: left = searchsorted(vals, key, 'left'), nulls are NULL_SENTINEL
    // (greater than every string, equal to a None key); the first key is ascending, so
    // the cut is `left`: the first row that is not before the boundary
    uint64_t lo = 0, hi = n;
    while (lo < hi) {
      const uint64_t mid = lo + ((hi - lo) >> 1);
      bool before;  // row < key
      if (row_is_null(col, mid)) before = false;
      else if (key_none) before = true;
      else before = cmp_bytes(col.data + col.off[mid], col.off[mid + 1] - col.off[mid], kb, kl) < 0;
      if (before) lo = mid + 1;
      else hi = mid;
    }
    cuts[b] = lo;
    return;
  }
  // descending: numpy.searchsorted over the reversed view, probe by probe (the view is
  // not sorted for numpy when null rows -- last in storage, first in the view -- exist;
  // util.py:775-799).  Element j of the view = row n-1-j.
  const uint64_t m = n;
  uint64_t mn = 0, mx = m;
  while (mn < mx) {  // side='left': advance while arr < key
    const uint64_t mid = mn + ((mx - mn) >> 1);
    const uint64_t r = n - 1 - mid;
    const bool el = row_is_null(col, r);
    bool lt;
    if (key_none) lt = !el;  // everything but the sentinel itself is below a None key
    else lt = !el && cmp_bytes(col.data + col.off[r], col.off[r + 1] - col.off[r], kb, kl) < 0;
    if (lt) mn = mid + 1;
    else mx = mid;
  }
  const uint64_t left_idx = mn;
  // first key descending: the cut is `right` of util.py:799-803 = n - left_idx
  cuts[b] = m - left_idx;
}

// destination partition of every row of an UNSORTED string column (per-row form of the
// rule above, as xchg_assign_partitions states it for numeric keys)
__global__ void __launch_bounds__(ST_THREADS)
k_assign_partitions_bytes(StrCol col, uint64_t n, const int64_t* __restrict__ boff,
                          const uint8_t* __restrict__ bdata, const uint8_t* __restrict__ bflags,
                          int nb, int desc, uint32_t* __restrict__ dest, unsigned long long* counts) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const bool null_row = row_is_null(col, i);
    const uint8_t* s = col.data + col.off[i];
    const int64_t sl = col.off[i + 1] - col.off[i];
    int lo = 0, hi = nb;
    while (lo < hi) {  // boundaries are in sort order: "row is right of b" is monotone in b
      const int mid = (lo + hi) >> 1;
      const bool none = bflags && (bflags[mid] & 4);
      bool right;
      if (!desc) {
        // right of b  <=>  not (row < b); a None boundary equals the null rows only
        if (null_row) right = true;
        else if (none) right = false;
        else right = cmp_bytes(s, sl, bdata + boff[mid], boff[mid + 1] - boff[mid]) >= 0;
      } else {
        // descending order: strictly after b  <=>  row < b (value order), nulls after all
        // values; a None boundary sorts before every row
        if (none) right = true;
        else if (null_row) right = true;
        else right = cmp_bytes(s, sl, bdata + boff[mid], boff[mid + 1] - boff[mid]) < 0;
      }
      if (right) lo = mid + 1;
      else hi = mid;
    }
    dest[i] = (uint32_t)lo;
    if (counts) atomicAdd(&counts[lo], 1ull);
  }
}

}  // namespace

extern "C" int xchg_string_lengths(const int64_t* offsets, uint64_t n, uint32_t* len_out,
                                   uint64_t* max_len_out, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  XCHG_REQUIRE(max_len_out != nullptr, "NULL max_len_out");
  XCHG_CUDA_TRY(cudaMemsetAsync(max_len_out, 0, 8, s));
  if (n == 0) return XCHG_OK;
  XCHG_REQUIRE(offsets && len_out, "NULL offsets / len_out");
  XCHG_LAUNCH(XK_NORMALIZE, s, k_string_lengths<<<st_grid(n), ST_THREADS, 0, s>>>(
      offsets, n, len_out, (unsigned long long*)max_len_out));
  XCHG_LAUNCH_CHECK();
  return XCHG_OK;
}

extern "C" int xchg_string_chunk_keys(const int64_t* offsets, const uint8_t* data, uint64_t n,
                                      uint32_t chunk, uint64_t* keys_out, void* stream) {
  if (n == 0) return XCHG_OK;
  XCHG_REQUIRE(offsets && keys_out, "NULL offsets / keys_out");
  XCHG_LAUNCH(XK_NORMALIZE, (cudaStream_t)stream, k_string_chunk_keys<<<st_grid(n), ST_THREADS, 0, (cudaStream_t)stream>>>(
      offsets, data, n, chunk, keys_out));
  XCHG_LAUNCH_CHECK();
  return XCHG_OK;
}

extern "C" int xchg_partition_bounds_bytes(const int64_t* sorted_offsets, const uint8_t* sorted_data,
                                           const uint8_t* sorted_validity, uint64_t n,
                                           const int64_t* bound_offsets, const uint8_t* bound_data,
                                           const uint8_t* bound_flags, int nb, int descending,
                                           uint64_t* cuts_out, void* stream) {
  if (nb == 0) return XCHG_OK;
  XCHG_REQUIRE(bound_offsets && cuts_out, "NULL argument");
  XCHG_REQUIRE(n == 0 || sorted_offsets, "NULL offsets");
  StrCol c{sorted_offsets, sorted_data, sorted_validity};
  XCHG_LAUNCH(XK_PARTITION, (cudaStream_t)stream, k_partition_bounds_bytes<<<(nb + 127) / 128, 128, 0, (cudaStream_t)stream>>>(
      c, n, bound_offsets, bound_data, bound_flags, nb, descending, cuts_out));
  XCHG_LAUNCH_CHECK();
  return XCHG_OK;
}

extern "C" int xchg_assign_partitions_bytes(const int64_t* offsets, const uint8_t* data,
                                            const uint8_t* validity, uint64_t n,
                                            const int64_t* bound_offsets, const uint8_t* bound_data,
                                            const uint8_t* bound_flags, int nb, int descending,
                                            uint32_t* dest_out, uint64_t* counts_out, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  if (counts_out) XCHG_CUDA_TRY(cudaMemsetAsync(counts_out, 0, (size_t)(nb + 1) * 8, s));
  if (n == 0) return XCHG_OK;
  XCHG_REQUIRE(offsets && dest_out, "NULL argument");
  XCHG_REQUIRE(nb == 0 || bound_offsets, "NULL boundaries");
  StrCol c{offsets, data, validity};
  XCHG_LAUNCH(XK_PARTITION, s, k_assign_partitions_bytes<<<st_grid(n), ST_THREADS, 0, s>>>(
      c, n, bound_offsets, bound_data, bound_flags, nb, descending, dest_out,
      (unsigned long long*)counts_out));
  XCHG_LAUNCH_CHECK();
  return XCHG_OK;
}
