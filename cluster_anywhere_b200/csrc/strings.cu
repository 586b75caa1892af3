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

// ---- MSD refinement of a string sort (levels after the first 8 bytes) ----
//
// After the rows have been sorted by chunk 0 only the rows that TIE there need another look,
// and of those only the ones whose group holds at least two different strings: a Zipf head
// of identical keys is final after one comparison with its neighbour.  The kernels below
// work on a LIST of m rows in their current order (level 0: all rows in sorted order; later:
// the rows still unresolved).

// gflag[j] = row j starts a new tie group (different previous-level group, different key at
// this level, or null against non-null); fflag[j] = gflag[j], or the two strings differ
// somewhere from byte `from_byte` on (lengths included).
__global__ void __launch_bounds__(ST_THREADS)
k_string_tie_flags(StrCol col, const uint32_t* __restrict__ rows, const uint64_t* __restrict__ keys,
                   const uint32_t* __restrict__ gid_prev, uint64_t m, int64_t from_byte,
                   uint8_t* __restrict__ gflag, uint8_t* __restrict__ fflag) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; j < m; j += stride) {
    bool g = j == 0, f = j == 0;
    if (!g) {
      const uint32_t a = rows[j - 1], b = rows[j];
      const bool na = row_is_null(col, a), nb = row_is_null(col, b);
      g = (gid_prev && gid_prev[j] != gid_prev[j - 1]) || keys[j] != keys[j - 1] || na != nb;
      f = g;
      if (!g && !nb) {
        const int64_t a0 = col.off[a], a1 = col.off[a + 1], b0 = col.off[b], b1 = col.off[b + 1];
        if (a1 - a0 != b1 - b0) {
          f = true;
        } else {
          for (int64_t q = from_byte; q < a1 - a0; ++q)
            if (col.data[a0 + q] != col.data[b0 + q]) {
              f = true;
              break;
            }
        }
      }
    }
    gflag[j] = g ? 1 : 0;
    fflag[j] = f ? 1 : 0;
  }
}

// a group is unresolved when two neighbours inside it differ
__global__ void __launch_bounds__(ST_THREADS)
k_mark_unresolved(const uint8_t* __restrict__ gflag, const uint8_t* __restrict__ fflag,
                  const uint32_t* __restrict__ gid, uint64_t m, uint8_t* __restrict__ unresolved) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; j < m; j += stride)
    if (fflag[j] && !gflag[j]) unresolved[gid[j]] = 1;
}

__global__ void __launch_bounds__(ST_THREADS)
k_flag_active(const uint8_t* __restrict__ unresolved, const uint32_t* __restrict__ gid, uint64_t m,
              uint32_t* __restrict__ active) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; j < m; j += stride)
    active[j] = unresolved[gid[j]] ? 1u : 0u;
}

// head[j] = row j (in level-0 order) does NOT repeat the string of row j-1 in the final order:
// it differs from it, or its group is still open (its rows move again, so nothing is known)
// A run is also cut every `period` rows: each stretch of the output then reads its own copy
// of the value, instead of every SM reading ONE source row of a heavy hitter (a single hot
// line serialises in its L2 slice).
__global__ void __launch_bounds__(ST_THREADS)
k_string_run_heads(const uint8_t* __restrict__ fflag, const uint32_t* __restrict__ active, uint64_t m,
                   uint32_t period, uint8_t* __restrict__ head) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; j < m; j += stride)
    head[j] = (fflag[j] || (active && active[j]) || (period && j % period == 0)) ? 1 : 0;
}

// list[excl[j]] = j for every flagged j (excl = exclusive scan of the flags)
__global__ void __launch_bounds__(ST_THREADS)
k_scatter_flagged(const uint32_t* __restrict__ active, const int64_t* __restrict__ excl, uint64_t m,
                  uint32_t* __restrict__ list) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; j < m; j += stride)
    if (active[j]) list[excl[j]] = (uint32_t)j;
}

// chunk keys (or, chunk == UINT32_MAX, lengths) of the listed rows
__global__ void __launch_bounds__(ST_THREADS)
k_string_chunk_keys_rows(const int64_t* __restrict__ off, const uint8_t* __restrict__ data,
                         const uint32_t* __restrict__ rows, uint64_t m, uint32_t chunk,
                         uint64_t* __restrict__ out) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; j < m; j += stride) {
    const uint32_t r = rows[j];
    const int64_t b = off[r], e = off[r + 1];
    uint64_t k = 0;
    if (chunk == 0xFFFFFFFFu) {
      k = (uint64_t)(e - b);
    } else {
      const int64_t p = b + (int64_t)chunk * 8;
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        const uint64_t byte = (p + q < e) ? (uint64_t)data[p + q] : 0ull;
        k = (k << 8) | byte;
      }
    }
    out[j] = k;
  }
}

__global__ void __launch_bounds__(ST_THREADS)
k_scatter_u32(const uint32_t* __restrict__ idx, const uint32_t* __restrict__ vals, uint64_t m,
              uint32_t* __restrict__ out) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; j < m; j += stride) out[idx[j]] = vals[j];
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

extern "C" int xchg_string_tie_flags(const int64_t* offsets, const uint8_t* data, const uint8_t* validity,
                                     const uint32_t* rows, const uint64_t* keys, const uint32_t* gid_prev,
                                     uint64_t m, uint64_t from_byte, uint8_t* gflag_out, uint8_t* fflag_out,
                                     void* stream) {
  if (m == 0) return XCHG_OK;
  XCHG_REQUIRE(offsets && rows && keys && gflag_out && fflag_out, "NULL argument");
  StrCol c{offsets, data, validity};
  XCHG_LAUNCH(XK_PARTITION, (cudaStream_t)stream,
              k_string_tie_flags<<<st_grid(m), ST_THREADS, 0, (cudaStream_t)stream>>>(
                  c, rows, keys, gid_prev, m, (int64_t)from_byte, gflag_out, fflag_out));
  XCHG_LAUNCH_CHECK();
  return XCHG_OK;
}

extern "C" int xchg_string_active_rows(const uint8_t* gflag, const uint8_t* fflag, const uint32_t* gid,
                                       uint64_t m, uint64_t num_groups, uint8_t* unresolved_scratch,
                                       uint32_t* active_out, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  if (m == 0) return XCHG_OK;
  XCHG_REQUIRE(gflag && fflag && gid && unresolved_scratch && active_out, "NULL argument");
  XCHG_CUDA_TRY(cudaMemsetAsync(unresolved_scratch, 0, num_groups, s));
  XCHG_LAUNCH(XK_PARTITION, s, k_mark_unresolved<<<st_grid(m), ST_THREADS, 0, s>>>(gflag, fflag, gid, m,
                                                                                unresolved_scratch));
  XCHG_LAUNCH(XK_PARTITION, s, k_flag_active<<<st_grid(m), ST_THREADS, 0, s>>>(unresolved_scratch, gid, m,
                                                                            active_out));
  XCHG_LAUNCH_CHECK();
  return XCHG_OK;
}

extern "C" int xchg_string_run_heads(const uint8_t* fflag, const uint32_t* active, uint64_t m, uint32_t period,
                                     uint8_t* head_out, void* stream) {
  if (m == 0) return XCHG_OK;
  XCHG_REQUIRE(fflag && head_out, "NULL argument");
  XCHG_LAUNCH(XK_MISC, (cudaStream_t)stream,
              k_string_run_heads<<<st_grid(m), ST_THREADS, 0, (cudaStream_t)stream>>>(fflag, active, m, period,
                                                                                    head_out));
  XCHG_LAUNCH_CHECK();
  return XCHG_OK;
}

extern "C" int xchg_scatter_flagged(const uint32_t* flags, const int64_t* excl_scan, uint64_t m,
                                    uint32_t* list_out, void* stream) {
  if (m == 0) return XCHG_OK;
  XCHG_REQUIRE(flags && excl_scan && list_out, "NULL argument");
  XCHG_LAUNCH(XK_MISC, (cudaStream_t)stream,
              k_scatter_flagged<<<st_grid(m), ST_THREADS, 0, (cudaStream_t)stream>>>(flags, excl_scan, m, list_out));
  XCHG_LAUNCH_CHECK();
  return XCHG_OK;
}

extern "C" int xchg_string_chunk_keys_rows(const int64_t* offsets, const uint8_t* data, const uint32_t* rows,
                                           uint64_t m, uint32_t chunk, uint64_t* keys_out, void* stream) {
  if (m == 0) return XCHG_OK;
  XCHG_REQUIRE(offsets && rows && keys_out, "NULL argument");
  XCHG_LAUNCH(XK_NORMALIZE, (cudaStream_t)stream,
              k_string_chunk_keys_rows<<<st_grid(m), ST_THREADS, 0, (cudaStream_t)stream>>>(offsets, data, rows, m,
                                                                                          chunk, keys_out));
  XCHG_LAUNCH_CHECK();
  return XCHG_OK;
}

extern "C" int xchg_scatter_u32(const uint32_t* idx, const uint32_t* vals, uint64_t m, uint32_t* out,
                                void* stream) {
  if (m == 0) return XCHG_OK;
  XCHG_REQUIRE(idx && vals && out, "NULL argument");
  XCHG_LAUNCH(XK_MISC, (cudaStream_t)stream,
              k_scatter_u32<<<st_grid(m), ST_THREADS, 0, (cudaStream_t)stream>>>(idx, vals, m, out));
  XCHG_LAUNCH_CHECK();
  return XCHG_OK;
}
