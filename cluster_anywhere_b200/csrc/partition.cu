// K4: boundary search on sorted, normalised key columns -> partition cut indices.
//
// Replaces find_partition_index / find_partitions
// (python/ray/data/_internal/util.py:728-828).  The reference runs, per boundary
// and per key column, two numpy.searchsorted calls (left/right; descending columns
// through a reversing sorter) after an O(N) `col == None` scan; here all key
// columns are already order-preserving unsigned radix keys sorted ascending, so
// every column -- whatever its direction -- is one lower_bound and one upper_bound
// inside the run left by the previous column.  One thread per boundary: the work
// is O(nb * ncols * log n) dependent loads, nowhere near any roofline.
#include <algorithm>

#include "common.cuh"

namespace {

constexpr int MAX_KEY_COLS = 8;
struct ColPtrs {
  const uint64_t* c[MAX_KEY_COLS];
  uint64_t last_thr[MAX_KEY_COLS];  // keys >= this are NaN / null ("last" class)
  uint8_t desc[MAX_KEY_COLS];       // column direction
  uint8_t has_last[MAX_KEY_COLS];   // column may hold NaN / null rows
};

__device__ __forceinline__ uint64_t lower_bound_u64(const uint64_t* __restrict__ a, uint64_t lo,
                                                    uint64_t hi, uint64_t v) {
  while (lo < hi) {
    const uint64_t mid = lo + ((hi - lo) >> 1);
    if (a[mid] < v) lo = mid + 1;
    else hi = mid;
  }
  return lo;
}
__device__ __forceinline__ uint64_t upper_bound_u64(const uint64_t* __restrict__ a, uint64_t lo,
                                                    uint64_t hi, uint64_t v) {
  while (lo < hi) {
    const uint64_t mid = lo + ((hi - lo) >> 1);
    if (a[mid] <= v) lo = mid + 1;
    else hi = mid;
  }
  return lo;
}

__global__ void k_partition_bounds(ColPtrs cols, int ncols, uint64_t n,
                                   const uint64_t* __restrict__ lower,
                                   const uint64_t* __restrict__ upper,
                                   const uint8_t* __restrict__ flags, int nb, int first_desc,
                                   uint64_t* __restrict__ cuts) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= nb) return;
  uint64_t lo = 0, hi = n;
  bool early = false;
  for (int c = 0; c < ncols; ++c) {
    if (lo == hi) {  // util.py:757-758: `if left == right: return right`
      early = true;
      break;
    }
    const int e = b * ncols + c;
    const uint8_t f = flags ? flags[e] : 0;
    if (f & 2) break;  // boundary tuple shorter than the key list
    if (f & 4) {
      // None boundary component: NULL_SENTINEL compares greater than every value
      // (util.py:67-89, :765-766).  Ascending column: it sorts after every row of the
      // current run (lo = hi = run end).  Descending column: the reversed search puts
      // it BEFORE every row (util.py:775-799 -> lo = hi = run start).
      if (cols.desc[c]) hi = lo;
      else lo = hi;
      continue;
    }
    const uint64_t L = lower[e], U = upper[e];
    if (!cols.desc[c]) {
      // ascending: numpy's order (numbers, then NaN) agrees with the stored order
      const uint64_t l2 = lower_bound_u64(cols.c[c], lo, hi, L);
      const uint64_t h2 = (f & 1) ? l2 : upper_bound_u64(cols.c[c], lo, hi, U);
      lo = l2;
      hi = h2;
    } else {
      // descending: the reference searches a REVERSED view with numpy.searchsorted
      // (util.py:775-799).  NaN / null rows sit at the end of the stored order, i.e.
      // at the FRONT of that view, where numpy's NaN-last comparator does not
      // expect them; the outcome then depends on the probe sequence.  Reproduce
      // numpy's binary search probe by probe so the cut points stay bit-exact.
      const uint64_t* a = cols.c[c];
      const uint64_t m = hi > lo ? hi - lo : 0;
      const uint64_t thr = cols.last_thr[c];
      const bool has_last = cols.has_last[c] != 0;
      const bool key_last = (f & 8) != 0;  // NaN boundary
      uint64_t mn = 0, mx = m;
      while (mn < mx) {  // side='left': advance while arr < key
        const uint64_t mid = mn + ((mx - mn) >> 1);
        const uint64_t k = a[hi - 1 - mid];
        const bool el = has_last && k >= thr;
        const bool before_eq = (f & 1) ? (k < L) : (k <= U);  // arr >= key (value space)
        const bool lt = key_last ? !el : (!el && !before_eq);
        if (lt) mn = mid + 1;
        else mx = mid;
      }
      const uint64_t left_idx = mn;
      mn = 0;
      mx = m;
      while (mn < mx) {  // side='right': advance while arr <= key
        const uint64_t mid = mn + ((mx - mn) >> 1);
        const uint64_t k = a[hi - 1 - mid];
        const bool el = has_last && k >= thr;
        const bool before = k < L;  // arr > key (value space)
        const bool key_lt_arr = !key_last && (el || before);
        if (!key_lt_arr) mn = mid + 1;
        else mx = mid;
      }
      const uint64_t right_idx = mn;
      const uint64_t base = lo;
      lo = base + (m - right_idx);
      hi = base + (m - left_idx);
    }
  }
  cuts[b] = early ? hi : (first_desc ? hi : lo);
}

}  // namespace

extern "C" int xchg_partition_bounds(const uint64_t* const* norm_cols_host,
                                     const uint8_t* col_desc_host,
                                     const uint64_t* last_threshold_host,
                                     const uint8_t* has_last_host, int ncols, uint64_t n,
                                     const uint64_t* lower, const uint64_t* upper,
                                     const uint8_t* flags, int nb, int first_descending,
                                     uint64_t* cuts_out, void* stream) {
  if (nb == 0) return XCHG_OK;
  XCHG_REQUIRE(norm_cols_host && col_desc_host && last_threshold_host && has_last_host && lower &&
                   upper && cuts_out,
               "NULL argument");
  XCHG_REQUIRE(ncols >= 1 && ncols <= MAX_KEY_COLS, "ncols must be in [1, %d]", MAX_KEY_COLS);
  ColPtrs cp;
  for (int c = 0; c < MAX_KEY_COLS; ++c) {
    cp.c[c] = c < ncols ? norm_cols_host[c] : nullptr;
    cp.last_thr[c] = c < ncols ? last_threshold_host[c] : ~0ull;
    cp.desc[c] = c < ncols ? col_desc_host[c] : 0;
    cp.has_last[c] = c < ncols ? has_last_host[c] : 0;
  }
  for (int c = 0; c < ncols; ++c) XCHG_REQUIRE(n == 0 || cp.c[c], "NULL key column %d", c);
  XCHG_LAUNCH(XK_PARTITION, (cudaStream_t)stream, k_partition_bounds<<<(nb + 127) / 128, 128, 0, (cudaStream_t)stream>>>(
      cp, ncols, n, lower, upper, flags, nb, first_descending, cuts_out));
  XCHG_LAUNCH_CHECK();
  return XCHG_OK;
}

// ------------------------------------------- key tuples with string columns ---
//
// The same search for key tuples that mix numeric and string / binary columns: the
// reference narrows [left, right) column by column whatever the column types
// (find_partition_index, util.py:757-802).  Numeric columns go through the radix-key search
// above; a string column is searched on the SORTED column itself (int64 offsets + bytes):
// object arrays have their nulls replaced by NULL_SENTINEL (greater than every value, equal
// to a None boundary, util.py:765-772); descending columns replay numpy.searchsorted over the
// reversed view probe by probe, both sides, as the single-column string kernel does for one.
namespace {

struct MixedCols {
  xchg_mixed_col c[MAX_KEY_COLS];
};

// three-way compare of a column element with a boundary value, NULL_SENTINEL semantics:
// -1 element < key, 0 equal, +1 element > key
__device__ __forceinline__ int cmp_elem_key(const xchg_mixed_col& col, uint64_t row, bool key_none,
                                            const uint8_t* kb, int64_t kl) {
  const bool el_null = col.validity && !bit_is_set(col.validity, (int64_t)row);
  if (el_null) return key_none ? 0 : 1;
  if (key_none) return -1;
  const int64_t a0 = col.offsets[row], a1 = col.offsets[row + 1];
  const uint8_t* a = col.data + a0;
  const int64_t la = a1 - a0;
  const int64_t m = la < kl ? la : kl;
  for (int64_t i = 0; i < m; ++i) {
    const int x = a[i], y = kb[i];
    if (x != y) return x < y ? -1 : 1;
  }
  return la < kl ? -1 : (la > kl ? 1 : 0);
}

__global__ void k_partition_bounds_mixed(MixedCols cols, int ncols, uint64_t n,
                                         const uint64_t* __restrict__ lower,
                                         const uint64_t* __restrict__ upper,
                                         const uint8_t* __restrict__ flags, int nb, int first_desc,
                                         uint64_t* __restrict__ cuts) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= nb) return;
  uint64_t lo = 0, hi = n;
  bool early = false;
  for (int c = 0; c < ncols; ++c) {
    if (lo == hi) {
      early = true;
      break;
    }
    const xchg_mixed_col& col = cols.c[c];
    const int e = b * ncols + c;
    const uint8_t f = flags ? flags[e] : 0;
    if (f & 2) break;
    if (col.kind == 1) {
      const bool key_none = (f & 4) != 0;
      const uint8_t* kb = col.bound_data + col.bound_offsets[b];
      const int64_t kl = col.bound_offsets[b + 1] - col.bound_offsets[b];
      const uint64_t m = hi - lo;
      if (!col.descending) {
        uint64_t l = lo, h = hi;
        while (l < h) {  // side='left': first element that is not < key
          const uint64_t mid = l + ((h - l) >> 1);
          if (cmp_elem_key(col, mid, key_none, kb, kl) < 0) l = mid + 1;
          else h = mid;
        }
        const uint64_t left = l;
        h = hi;
        while (l < h) {  // side='right': first element that is > key
          const uint64_t mid = l + ((h - l) >> 1);
          if (cmp_elem_key(col, mid, key_none, kb, kl) <= 0) l = mid + 1;
          else h = mid;
        }
        lo = left;
        hi = l;
      } else {
        // numpy.searchsorted(col_vals, key, side, sorter=reversed): view element j = row hi-1-j
        uint64_t mn = 0, mx = m;
        while (mn < mx) {  // side='left': advance while view[mid] < key
          const uint64_t mid = mn + ((mx - mn) >> 1);
          if (cmp_elem_key(col, hi - 1 - mid, key_none, kb, kl) < 0) mn = mid + 1;
          else mx = mid;
        }
        const uint64_t left_idx = mn;
        mn = 0;
        mx = m;
        while (mn < mx) {  // side='right': advance while not (key < view[mid])
          const uint64_t mid = mn + ((mx - mn) >> 1);
          if (cmp_elem_key(col, hi - 1 - mid, key_none, kb, kl) <= 0) mn = mid + 1;
          else mx = mid;
        }
        const uint64_t right_idx = mn;
        const uint64_t base = lo;
        lo = base + (m - right_idx);
        hi = base + (m - left_idx);
      }
      continue;
    }
    // numeric column: as k_partition_bounds
    if (f & 4) {
      if (col.descending) hi = lo;
      else lo = hi;
      continue;
    }
    const uint64_t L = lower[e], U = upper[e];
    const uint64_t* a = col.norm;
    if (!col.descending) {
      const uint64_t l2 = lower_bound_u64(a, lo, hi, L);
      const uint64_t h2 = (f & 1) ? l2 : upper_bound_u64(a, lo, hi, U);
      lo = l2;
      hi = h2;
    } else {
      const uint64_t m = hi > lo ? hi - lo : 0;
      const uint64_t thr = col.last_threshold;
      const bool has_last = col.has_last != 0;
      const bool key_last = (f & 8) != 0;
      uint64_t mn = 0, mx = m;
      while (mn < mx) {
        const uint64_t mid = mn + ((mx - mn) >> 1);
        const uint64_t k = a[hi - 1 - mid];
        const bool el = has_last && k >= thr;
        const bool before_eq = (f & 1) ? (k < L) : (k <= U);
        const bool lt = key_last ? !el : (!el && !before_eq);
        if (lt) mn = mid + 1;
        else mx = mid;
      }
      const uint64_t left_idx = mn;
      mn = 0;
      mx = m;
      while (mn < mx) {
        const uint64_t mid = mn + ((mx - mn) >> 1);
        const uint64_t k = a[hi - 1 - mid];
        const bool el = has_last && k >= thr;
        const bool key_lt_arr = !key_last && (el || k < L);
        if (!key_lt_arr) mn = mid + 1;
        else mx = mid;
      }
      const uint64_t right_idx = mn;
      const uint64_t base = lo;
      lo = base + (m - right_idx);
      hi = base + (m - left_idx);
    }
  }
  cuts[b] = early ? hi : (first_desc ? hi : lo);
}

}  // namespace

extern "C" int xchg_partition_bounds_mixed(const xchg_mixed_col* cols_host, int ncols, uint64_t n,
                                           const uint64_t* lower, const uint64_t* upper,
                                           const uint8_t* flags, int nb, int first_descending,
                                           uint64_t* cuts_out, void* stream) {
  if (nb == 0) return XCHG_OK;
  XCHG_REQUIRE(cols_host && cuts_out, "NULL argument");
  XCHG_REQUIRE(ncols >= 1 && ncols <= MAX_KEY_COLS, "ncols must be in [1, %d]", MAX_KEY_COLS);
  MixedCols mc;
  for (int c = 0; c < ncols; ++c) {
    mc.c[c] = cols_host[c];
    if (mc.c[c].kind == 1) {
      XCHG_REQUIRE(n == 0 || mc.c[c].offsets, "NULL offsets of string column %d", c);
      XCHG_REQUIRE(mc.c[c].bound_offsets, "NULL boundary offsets of string column %d", c);
    } else {
      XCHG_REQUIRE(mc.c[c].kind == 0, "column %d: kind must be 0 or 1", c);
      XCHG_REQUIRE((n == 0 || mc.c[c].norm) && lower && upper, "NULL radix keys / bounds of column %d", c);
    }
  }
  XCHG_LAUNCH(XK_PARTITION, (cudaStream_t)stream,
              k_partition_bounds_mixed<<<(nb + 127) / 128, 128, 0, (cudaStream_t)stream>>>(
                  mc, ncols, n, lower, upper, flags, nb, first_descending, cuts_out));
  XCHG_LAUNCH_CHECK();
  return XCHG_OK;
}

// ------------------------------------------------------- per-row assignment ---
//
// xchg_assign_partitions: the same boundary rule as xchg_partition_bounds, stated
// per ROW instead of per boundary, for rows that are NOT sorted yet.  The
// reference only ever partitions a sorted block (find_partitions,
// python/ray/data/_internal/util.py:806-828: partition i = table[cut[i-1]:cut[i]]);
// a row therefore lands in partition  #(boundaries whose cut index is <= the row's
// sorted position).  With cut = lo (first key ascending) that is the number of
// boundaries the row is NOT strictly before; with cut = hi (first key descending)
// the number of boundaries the row is strictly AFTER (util.py:803).  Evaluating
// that predicate row by row lets a rank split its rows by destination GPU before
// any sort (the exchange's "map" side), and is what the reduce side re-derives
// with xchg_partition_bounds on the merged, sorted rows.
//
// Column-wise comparison of row key t against boundary component (L, U, flags):
//   bit1 (tuple ended)      -> equal on all remaining columns
//   bit2 (None component)   -> the boundary sorts after the row (ascending column) /
//                              before the row (descending column)
//   t <  L                  -> row before boundary
//   bit0 (no equal value)   -> t >= L: row after boundary
//   t >  U                  -> row after boundary;  else equal, next column
namespace {

constexpr int AP_THREADS = 256;
constexpr int AP_MAX_SMEM_BINS = 8192;

// -1 row before boundary b, 0 equal, +1 row after
__device__ __forceinline__ int cmp_row_boundary(const ColPtrs& cols, int ncols, uint64_t row, uint64_t t0,
                                                const uint64_t* __restrict__ lower,
                                                const uint64_t* __restrict__ upper,
                                                const uint8_t* __restrict__ flags, int b) {
  for (int c = 0; c < ncols; ++c) {
    const int e = b * ncols + c;
    const uint8_t f = flags ? flags[e] : 0;
    if (f & 2) return 0;
    if (f & 4) return cols.desc[c] ? 1 : -1;
    const uint64_t t = c == 0 ? t0 : cols.c[c][row];
    const uint64_t L = lower[e];
    if (t < L) return -1;
    if (f & 1) return 1;
    if (t > upper[e]) return 1;
  }
  return 0;
}

__global__ void __launch_bounds__(AP_THREADS)
k_assign_partitions(ColPtrs cols, int ncols, uint64_t n, const uint64_t* __restrict__ lower,
                    const uint64_t* __restrict__ upper, const uint8_t* __restrict__ flags, int nb,
                    int first_desc, uint32_t* __restrict__ dest, unsigned long long* counts,
                    int smem_bins) {
  extern __shared__ uint32_t s_cnt[];
  for (int j = threadIdx.x; j < smem_bins; j += blockDim.x) s_cnt[j] = 0;
  __syncthreads();
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const uint64_t t0 = cols.c[0][i];
    // boundaries are in sort order, so "row is right of boundary b" is monotone in b:
    // binary search for the first boundary the row is left of
    int lo = 0, hi = nb;
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      const int c = cmp_row_boundary(cols, ncols, i, t0, lower, upper, flags, mid);
      const bool right = first_desc ? (c > 0) : (c >= 0);
      if (right) lo = mid + 1;
      else hi = mid;
    }
    dest[i] = (uint32_t)lo;
    if (counts) {
      if (smem_bins) atomicAdd(&s_cnt[lo], 1u);
      else atomicAdd(&counts[lo], 1ull);
    }
  }
  if (counts && smem_bins) {
    __syncthreads();
    for (int j = threadIdx.x; j < smem_bins; j += blockDim.x)
      if (s_cnt[j]) atomicAdd(&counts[j], (unsigned long long)s_cnt[j]);
  }
}

}  // namespace

extern "C" int xchg_assign_partitions(const uint64_t* const* norm_cols_host,
                                      const uint8_t* col_desc_host, int ncols, uint64_t n,
                                      const uint64_t* lower, const uint64_t* upper,
                                      const uint8_t* flags, int nb, int first_descending,
                                      uint32_t* dest_out, uint64_t* counts_out, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  XCHG_REQUIRE(nb >= 0, "nb must be >= 0");
  if (counts_out) XCHG_CUDA_TRY(cudaMemsetAsync(counts_out, 0, (size_t)(nb + 1) * 8, s));
  if (n == 0) return XCHG_OK;
  XCHG_REQUIRE(norm_cols_host && col_desc_host && dest_out, "NULL argument");
  XCHG_REQUIRE(nb == 0 || (lower && upper), "NULL boundaries");
  XCHG_REQUIRE(ncols >= 1 && ncols <= MAX_KEY_COLS, "ncols must be in [1, %d]", MAX_KEY_COLS);
  ColPtrs cp;
  for (int c = 0; c < MAX_KEY_COLS; ++c) {
    cp.c[c] = c < ncols ? norm_cols_host[c] : nullptr;
    cp.last_thr[c] = ~0ull;
    cp.desc[c] = c < ncols ? col_desc_host[c] : 0;
    cp.has_last[c] = 0;
  }
  for (int c = 0; c < ncols; ++c) XCHG_REQUIRE(cp.c[c], "NULL key column %d", c);
  const int bins = (counts_out && nb + 1 <= AP_MAX_SMEM_BINS) ? nb + 1 : 0;
  const int grid = (int)std::max<uint64_t>(
      1, std::min<uint64_t>((n + AP_THREADS - 1) / AP_THREADS, (uint64_t)xchg_sm_count() * 8));
  XCHG_LAUNCH(XK_PARTITION, s,
              k_assign_partitions<<<grid, AP_THREADS, (size_t)bins * 4, s>>>(
                  cp, ncols, n, lower, upper, flags, nb, first_descending, dest_out,
                  (unsigned long long*)counts_out, bins));
  XCHG_LAUNCH_CHECK();
  return XCHG_OK;
}

// ---- per-row form for key tuples with string columns (see xchg_partition_bounds_mixed) ----
namespace {

// three-way compare, in SORT order, of row `row` with boundary b: -1 row before, 0 equal, +1 after.
// Numeric columns as cmp_row_boundary; string columns as k_assign_partitions_bytes states the rule.
__device__ __forceinline__ int cmp_row_boundary_mixed(const MixedCols& cols, int ncols, uint64_t row,
                                                      const uint64_t* __restrict__ lower,
                                                      const uint64_t* __restrict__ upper,
                                                      const uint8_t* __restrict__ flags, int b) {
  for (int c = 0; c < ncols; ++c) {
    const xchg_mixed_col& col = cols.c[c];
    const int e = b * ncols + c;
    const uint8_t f = flags ? flags[e] : 0;
    if (f & 2) return 0;
    if (col.kind == 1) {
      const bool none = (f & 4) != 0;
      const bool null_row = col.validity && !bit_is_set(col.validity, (int64_t)row);
      if (!col.descending) {
        if (null_row) {
          if (none) continue;  // NULL_SENTINEL == NULL_SENTINEL: next column decides
          return 1;
        }
        if (none) return -1;
      } else {
        if (none || null_row) return 1;  // a None boundary sorts before every row; nulls after all values
      }
      const int64_t a0 = col.offsets[row], a1 = col.offsets[row + 1];
      const uint8_t* kb = col.bound_data + col.bound_offsets[b];
      const int64_t kl = col.bound_offsets[b + 1] - col.bound_offsets[b];
      const uint8_t* a = col.data + a0;
      const int64_t la = a1 - a0, m = la < kl ? la : kl;
      int r = 0;
      for (int64_t i = 0; i < m && r == 0; ++i) {
        const int x = a[i], y = kb[i];
        if (x != y) r = x < y ? -1 : 1;
      }
      if (r == 0) r = la < kl ? -1 : (la > kl ? 1 : 0);
      if (r != 0) return col.descending ? -r : r;
      continue;
    }
    if (f & 4) return col.descending ? 1 : -1;
    const uint64_t t = col.norm[row];
    if (t < lower[e]) return -1;
    if (f & 1) return 1;
    if (t > upper[e]) return 1;
  }
  return 0;
}

__global__ void __launch_bounds__(AP_THREADS)
k_assign_partitions_mixed(MixedCols cols, int ncols, uint64_t n, const uint64_t* __restrict__ lower,
                          const uint64_t* __restrict__ upper, const uint8_t* __restrict__ flags, int nb,
                          int first_desc, uint32_t* __restrict__ dest, unsigned long long* counts) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    int lo = 0, hi = nb;
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      const int c = cmp_row_boundary_mixed(cols, ncols, i, lower, upper, flags, mid);
      const bool right = first_desc ? (c > 0) : (c >= 0);
      if (right) lo = mid + 1;
      else hi = mid;
    }
    dest[i] = (uint32_t)lo;
    if (counts) atomicAdd(&counts[lo], 1ull);
  }
}

}  // namespace

extern "C" int xchg_assign_partitions_mixed(const xchg_mixed_col* cols_host, int ncols, uint64_t n,
                                            const uint64_t* lower, const uint64_t* upper,
                                            const uint8_t* flags, int nb, int first_descending,
                                            uint32_t* dest_out, uint64_t* counts_out, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  XCHG_REQUIRE(nb >= 0, "nb must be >= 0");
  if (counts_out) XCHG_CUDA_TRY(cudaMemsetAsync(counts_out, 0, (size_t)(nb + 1) * 8, s));
  if (n == 0) return XCHG_OK;
  XCHG_REQUIRE(cols_host && dest_out, "NULL argument");
  XCHG_REQUIRE(ncols >= 1 && ncols <= MAX_KEY_COLS, "ncols must be in [1, %d]", MAX_KEY_COLS);
  MixedCols mc;
  for (int c = 0; c < ncols; ++c) {
    mc.c[c] = cols_host[c];
    if (mc.c[c].kind == 1) {
      XCHG_REQUIRE(mc.c[c].offsets, "NULL offsets of string column %d", c);
      XCHG_REQUIRE(nb == 0 || mc.c[c].bound_offsets, "NULL boundary offsets of string column %d", c);
    } else {
      XCHG_REQUIRE(mc.c[c].kind == 0, "column %d: kind must be 0 or 1", c);
      XCHG_REQUIRE(mc.c[c].norm && (nb == 0 || (lower && upper)), "NULL radix keys / bounds of column %d", c);
    }
  }
  const int grid = (int)std::max<uint64_t>(
      1, std::min<uint64_t>((n + AP_THREADS - 1) / AP_THREADS, (uint64_t)xchg_sm_count() * 8));
  XCHG_LAUNCH(XK_PARTITION, s,
              k_assign_partitions_mixed<<<grid, AP_THREADS, 0, s>>>(mc, ncols, n, lower, upper, flags, nb,
                                                                    first_descending, dest_out,
                                                                    (unsigned long long*)counts_out));
  XCHG_LAUNCH_CHECK();
  return XCHG_OK;
}
