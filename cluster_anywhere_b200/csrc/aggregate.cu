// K8 / K9: hash group-by aggregate for sm_100a.
//
// Replaces the reference's sort-based aggregate arithmetic: the per-group Python loop
// of ArrowBlockAccessor.combine (python/ray/data/_internal/arrow_block.py:452-526,
// one pyarrow.compute call per group and aggregate: :346-423) and the per-row
// heapq.merge of aggregate_combined_blocks (:544-663), for the built-in accumulators
// Count / Sum / Min / Max / Mean / Std (python/ray/data/aggregate.py:97-326) under the
// null policy of _internal/null_aggregate.py:11-25.
//
// Plan (one key column; keys are the order-preserving u64 radix keys of K2, so that
// "sorted by key" later is a radix sort of the few group keys):
//   k_hash_build      open-addressing table (linear probing, atomicCAS on the key
//                     word): every row finds / claims the slot of its key and
//                     records it.  Null rows and the one key value that collides
//                     with the EMPTY sentinel use two slots outside the table.
//   k_hash_compact    occupied slots -> dense list of (group key, slot)
//   [xchg_sort_column on the group keys: output order = numbers < NaN < null]
//   k_assign_gid      slot -> dense group id in OUTPUT order
//   k_lookup_gid      row -> group id
//   k_acc_*           per value column, accumulate into DENSE per-group arrays
//                     (10 M groups x 8 B = 80 MB per array: L2-resident on B200,
//                     where the sparse table is not) with warp-aggregated atomics:
//                     lanes of a warp that hit the same group are combined by one
//                     leader (match_any + shuffles) before a single atomic per
//                     statistic -- count int64, sum fp64 / int64, min / max on
//                     order-preserving encodings, NaN count; second pass for
//                     sum((x-mean)^2) when Std is requested.
//   k_merge_*         the same over PARTIAL rows (K9: map-side partials from several
//                     blocks / ranks; M2 combined as in Chan et al., the formula of
//                     aggregate.py:275-289 summed over all partials at once)
//   k_finalize        statistics -> output values + validity (mean = sum/count,
//                     std = sqrt(M2/(n-ddof)), < 2 values -> 0.0, all-null -> null)
// Algorithmic bytes: 8 B/row for the key scan + 4 (slot) + per value column its
// width + 4 (group id); the accumulators live in L2.
#include <algorithm>
#include <type_traits>

#include "common.cuh"

namespace {

constexpr uint64_t EMPTY = 0xFFFFFFFFFFFFFFFFull;
constexpr int AG_THREADS = 256;

__device__ __forceinline__ uint64_t mix64(uint64_t k) {
  k ^= k >> 33;
  k *= 0xff51afd7ed558ccdull;
  k ^= k >> 33;
  k *= 0xc4ceb9fe1a85ec53ull;
  k ^= k >> 33;
  return k;
}

int ag_grid(uint64_t n) {
  return (int)std::max<uint64_t>(1, std::min<uint64_t>((n + AG_THREADS - 1) / AG_THREADS,
                                                       (uint64_t)xchg_sm_count() * 8));
}

// ---- table ---------------------------------------------------------------------

__global__ void __launch_bounds__(AG_THREADS)
k_hash_build(const uint64_t* __restrict__ norm, const uint8_t* __restrict__ validity, int64_t vbit,
             uint64_t n, unsigned long long* table, uint64_t cap, uint32_t* __restrict__ slot_out,
             uint32_t* special_present, uint32_t* overflow) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  const uint64_t mask = cap - 1;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const uint64_t key = norm[i];
    const bool valid = validity ? bit_is_set(validity, (int64_t)i + vbit) : true;
    uint32_t slot;
    if (!valid) {
      slot = (uint32_t)cap;  // the null group
      special_present[0] = 1;
    } else if (key == EMPTY) {
      slot = (uint32_t)cap + 1;  // the one key that equals the sentinel
      special_present[1] = 1;
    } else {
      uint64_t h = mix64(key) & mask;
      uint64_t probes = 0;
      while (true) {
        unsigned long long cur = table[h];
        if (cur == EMPTY) cur = atomicCAS(&table[h], (unsigned long long)EMPTY, (unsigned long long)key);
        if (cur == EMPTY || cur == key) break;
        h = (h + 1) & mask;
        if (++probes > cap) {
          *overflow = 1;
          break;
        }
      }
      slot = (uint32_t)h;
    }
    slot_out[i] = slot;
  }
}

__global__ void __launch_bounds__(AG_THREADS)
k_hash_compact(const unsigned long long* __restrict__ table, uint64_t cap,
               uint64_t* __restrict__ out_keys, uint32_t* __restrict__ out_slots,
               unsigned long long* counter) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  const int lane = threadIdx.x & 31;
  const uint64_t cap_up = (cap + 31) & ~31ull;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < cap_up; i += stride) {
    const unsigned long long k = i < cap ? table[i] : EMPTY;
    const bool occ = k != EMPTY;
    const unsigned m = __ballot_sync(0xffffffffu, occ);
    if (m == 0) continue;
    unsigned long long base = 0;
    if (lane == 0) base = atomicAdd(counter, (unsigned long long)__popc(m));
    base = __shfl_sync(0xffffffffu, base, 0);
    if (occ) {
      const uint64_t o = base + __popc(m & ((1u << lane) - 1));
      out_keys[o] = k;
      out_slots[o] = (uint32_t)i;
    }
  }
}

// the two out-of-table groups go last: [key == sentinel value] then [null]; both get
// key UINT64_MAX and the later sort is stable
__global__ void k_hash_append_specials(const uint32_t* special_present, uint64_t cap,
                                       uint64_t* out_keys, uint32_t* out_slots,
                                       unsigned long long* counter) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    unsigned long long c = *counter;
    if (special_present[1]) {
      out_keys[c] = EMPTY;
      out_slots[c] = (uint32_t)cap + 1;
      ++c;
    }
    if (special_present[0]) {
      out_keys[c] = EMPTY;
      out_slots[c] = (uint32_t)cap;
      ++c;
    }
    *counter = c;
  }
}

__global__ void __launch_bounds__(AG_THREADS)
k_assign_gid(const uint32_t* __restrict__ order, const uint32_t* __restrict__ slots, uint64_t g,
             uint32_t* __restrict__ gid_of_slot) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; j < g; j += stride)
    gid_of_slot[slots[order[j]]] = (uint32_t)j;
}

__global__ void __launch_bounds__(AG_THREADS)
k_lookup_gid(const uint32_t* __restrict__ slot, const uint32_t* __restrict__ gid_of_slot, uint64_t n,
             uint32_t* __restrict__ rgid) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    rgid[i] = ld_random(gid_of_slot + slot[i]);
}

// ---- order-preserving encodings for min / max -------------------------------------

template <typename T>
__device__ __forceinline__ uint64_t enc_value(T x) {
  if constexpr (std::is_same<T, float>::value || std::is_same<T, double>::value) {
    const double d = (double)x;
    const uint64_t b = (uint64_t)__double_as_longlong(d);
    return (b >> 63) ? ~b : (b | 0x8000000000000000ull);
  } else if constexpr (std::is_signed<T>::value) {
    return (uint64_t)(int64_t)x ^ 0x8000000000000000ull;
  } else {
    return (uint64_t)x;
  }
}

// ---- accumulate raw rows ------------------------------------------------------------

struct AccArrays {
  unsigned long long* cnt;   // non-null values
  void* sum;                 // int64 (integer columns) or double
  unsigned long long* minv;  // encodings; init UINT64_MAX
  unsigned long long* maxv;  // init 0
  unsigned long long* nan;   // NaN values (float columns)
};

// With many groups two lanes of a warp almost never share one, and the match / fold
// machinery below is pure overhead (match.any alone costs ~64 cycles per warp on the
// SM's single ADU pipe when all 32 values differ): the DIRECT kernels issue one atomic
// per row and statistic.  The host picks them when num_groups > AG_DIRECT_MIN_GROUPS.
constexpr uint64_t AG_DIRECT_MIN_GROUPS = 1u << 16;

__global__ void __launch_bounds__(AG_THREADS)
k_acc_rows_direct(const uint32_t* __restrict__ rgid, uint64_t n, unsigned long long* rows) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    atomicAdd(&rows[rgid[i]], 1ull);
}

template <typename T>
__global__ void __launch_bounds__(AG_THREADS)
k_acc_values_direct(const uint32_t* __restrict__ rgid, const T* __restrict__ vals,
                    const uint8_t* __restrict__ validity, int64_t vbit, uint64_t n, AccArrays a,
                    int want_minmax, int want_cnt);

__global__ void __launch_bounds__(AG_THREADS)
k_acc_rows(const uint32_t* __restrict__ rgid, uint64_t n, unsigned long long* rows) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  const uint64_t n_up = (n + 31) & ~31ull;
  const int lane = threadIdx.x & 31;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_up; i += stride) {
    const bool active = i < n;
    const unsigned act = __ballot_sync(0xffffffffu, active);
    if (!active) continue;
    const uint32_t g = rgid ? rgid[i] : 0u;
    const unsigned m = __match_any_sync(act, g);
    if (lane == __ffs(m) - 1) atomicAdd(&rows[g], (unsigned long long)__popc(m));
  }
}

template <typename T>
__global__ void __launch_bounds__(AG_THREADS)
k_acc_values(const uint32_t* __restrict__ rgid, const T* __restrict__ vals,
             const uint8_t* __restrict__ validity, int64_t vbit, uint64_t n, AccArrays a,
             int want_minmax) {
  constexpr bool IS_FP = std::is_floating_point<T>::value;
  using S = typename std::conditional<IS_FP, double, long long>::type;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  const uint64_t n_up = (n + 31) & ~31ull;
  const int lane = threadIdx.x & 31;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_up; i += stride) {
    const bool active = i < n;
    const unsigned act = __ballot_sync(0xffffffffu, active);
    if (!active) continue;
    const uint32_t g = rgid ? rgid[i] : 0u;
    const bool valid = validity ? bit_is_set(validity, (int64_t)i + vbit) : true;
    const T x = valid ? vals[i] : T(0);
    const unsigned m = __match_any_sync(act, g);
    const int leader = __ffs(m) - 1;
    // The leader of every same-group lane set folds the set's values, then issues one
    // atomic per statistic.  All lanes walk the loop together with FULL-warp shuffles
    // (trip count = the largest set in the warp: 1 when every lane has its own group);
    // per-set shuffle masks would make the hardware converge and run the sets one after
    // the other.
    S sum = 0;
    unsigned cnt = 0, nan = 0;
    uint64_t mn = EMPTY, mx = 0;
    const int steps = (int)__reduce_max_sync(act, (unsigned)__popc(m));
    unsigned rem = m;
    for (int st = 0; st < steps; ++st) {
      const int src = rem ? __ffs(rem) - 1 : lane;
      const T xv = __shfl_sync(act, x, src);
      const int vv = __shfl_sync(act, (int)valid, src);
      if (lane == leader && rem && vv) {
        ++cnt;
        sum += (S)xv;
        if (want_minmax) {
          bool is_nan = false;
          if constexpr (IS_FP) is_nan = xv != xv;
          if (is_nan) {
            ++nan;
          } else {
            const uint64_t e = enc_value<T>(xv);
            mn = e < mn ? e : mn;
            mx = e > mx ? e : mx;
          }
        }
      }
      rem &= rem - 1;
    }
    if (lane == leader && cnt) {
      atomicAdd(&a.cnt[g], (unsigned long long)cnt);
      if constexpr (IS_FP) atomicAdd(&((double*)a.sum)[g], sum);
      else atomicAdd(&((unsigned long long*)a.sum)[g], (unsigned long long)sum);
      if (want_minmax) {
        if (nan) atomicAdd(&a.nan[g], (unsigned long long)nan);
        if (cnt > nan) {
          atomicMin(&a.minv[g], (unsigned long long)mn);
          atomicMax(&a.maxv[g], (unsigned long long)mx);
        }
      }
    }
  }
}

template <typename T>
__global__ void __launch_bounds__(AG_THREADS)
k_acc_values_direct(const uint32_t* __restrict__ rgid, const T* __restrict__ vals,
                    const uint8_t* __restrict__ validity, int64_t vbit, uint64_t n, AccArrays a,
                    int want_minmax, int want_cnt) {
  constexpr bool IS_FP = std::is_floating_point<T>::value;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    if (validity && !bit_is_set(validity, (int64_t)i + vbit)) continue;
    const uint32_t g = rgid[i];
    const T x = vals[i];
    if (want_cnt) atomicAdd(&a.cnt[g], 1ull);
    if constexpr (IS_FP) atomicAdd(&((double*)a.sum)[g], (double)x);
    else atomicAdd(&((unsigned long long*)a.sum)[g], (unsigned long long)(long long)x);
    if (want_minmax) {
      bool is_nan = false;
      if constexpr (IS_FP) is_nan = x != x;
      if (is_nan) {
        atomicAdd(&a.nan[g], 1ull);
      } else {
        const uint64_t e = enc_value<T>(x);
        atomicMin(&a.minv[g], (unsigned long long)e);
        atomicMax(&a.maxv[g], (unsigned long long)e);
      }
    }
  }
}

template <typename T>
__global__ void __launch_bounds__(AG_THREADS)
k_acc_m2(const uint32_t* __restrict__ rgid, const T* __restrict__ vals,
         const uint8_t* __restrict__ validity, int64_t vbit, uint64_t n,
         const unsigned long long* __restrict__ cnt, const void* __restrict__ sum, int sum_is_int,
         double* m2) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  const uint64_t n_up = (n + 31) & ~31ull;
  const int lane = threadIdx.x & 31;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_up; i += stride) {
    const bool active = i < n;
    const unsigned act = __ballot_sync(0xffffffffu, active);
    if (!active) continue;
    const uint32_t g = rgid ? rgid[i] : 0u;
    const bool valid = validity ? bit_is_set(validity, (int64_t)i + vbit) : true;
    double d2 = 0.0;
    if (valid) {
      const double c = (double)cnt[g];
      const double s = sum_is_int ? (double)((const long long*)sum)[g] : ((const double*)sum)[g];
      const double d = (double)vals[i] - s / c;
      d2 = d * d;
    }
    const unsigned m = __match_any_sync(act, g);
    const int leader = __ffs(m) - 1;
    double tot = 0.0;
    const int steps = (int)__reduce_max_sync(act, (unsigned)__popc(m));
    unsigned rem = m;
    for (int st = 0; st < steps; ++st) {
      const double v = __shfl_sync(act, d2, rem ? __ffs(rem) - 1 : lane);
      if (lane == leader && rem) tot += v;
      rem &= rem - 1;
    }
    if (lane == leader && tot != 0.0) atomicAdd(&m2[g], tot);  // (NaN != 0: it propagates)
  }
}

// ---- merge partial rows (K9) ---------------------------------------------------------
// A partial row of a value column is six 8-byte words:
//   [0] cnt  [1] sum (int64 or double)  [2] min enc  [3] max enc  [4] nan cnt  [5] M2
constexpr int ST_WORDS = 6;

__global__ void __launch_bounds__(AG_THREADS)
k_merge_rows(const uint32_t* __restrict__ rgid, const long long* __restrict__ part_rows, uint64_t n,
             unsigned long long* rows) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    atomicAdd(&rows[rgid ? rgid[i] : 0u], (unsigned long long)part_rows[i]);
}

__global__ void __launch_bounds__(AG_THREADS)
k_merge_states(const uint32_t* __restrict__ rgid, const uint64_t* __restrict__ st, uint64_t n,
               int sum_is_int, AccArrays a) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const uint32_t g = rgid ? rgid[i] : 0u;
    const uint64_t* s = st + i * ST_WORDS;
    const uint64_t c = s[0];
    if (c == 0) continue;
    atomicAdd(&a.cnt[g], (unsigned long long)c);
    if (sum_is_int) atomicAdd(&((unsigned long long*)a.sum)[g], (unsigned long long)s[1]);
    else atomicAdd(&((double*)a.sum)[g], __longlong_as_double((long long)s[1]));
    atomicMin(&a.minv[g], (unsigned long long)s[2]);
    atomicMax(&a.maxv[g], (unsigned long long)s[3]);
    if (s[4]) atomicAdd(&a.nan[g], (unsigned long long)s[4]);
  }
}

// M2 of the union = sum_i [ M2_i + cnt_i * (mean_i - mean)^2 ]   (Chan et al.; the
// pairwise form in aggregate.py:275-289 telescopes to this)
__global__ void __launch_bounds__(AG_THREADS)
k_merge_m2(const uint32_t* __restrict__ rgid, const uint64_t* __restrict__ st, uint64_t n,
           int sum_is_int, const unsigned long long* __restrict__ cnt,
           const void* __restrict__ sum, double* m2) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const uint32_t g = rgid ? rgid[i] : 0u;
    const uint64_t* s = st + i * ST_WORDS;
    const uint64_t c = s[0];
    if (c == 0) continue;
    const double si = sum_is_int ? (double)(long long)s[1] : __longlong_as_double((long long)s[1]);
    const double tot = sum_is_int ? (double)((const long long*)sum)[g] : ((const double*)sum)[g];
    const double d = si / (double)c - tot / (double)cnt[g];
    atomicAdd(&m2[g], __longlong_as_double((long long)s[5]) + (double)c * d * d);
  }
}

__global__ void __launch_bounds__(AG_THREADS)
k_pack_states(uint64_t g, AccArrays a, const double* __restrict__ m2, uint64_t* __restrict__ out) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < g; i += stride) {
    uint64_t* o = out + i * ST_WORDS;
    o[0] = a.cnt[i];
    o[1] = ((const uint64_t*)a.sum)[i];
    o[2] = a.minv[i];
    o[3] = a.maxv[i];
    o[4] = a.nan[i];
    o[5] = m2 ? (uint64_t)__double_as_longlong(m2[i]) : 0ull;
  }
}

// ---- finalize ---------------------------------------------------------------------------

__device__ __forceinline__ double dec_double(uint64_t e) {
  const uint64_t b = (e >> 63) ? (e & 0x7FFFFFFFFFFFFFFFull) : ~e;
  return __longlong_as_double((long long)b);
}

// kind: 0 count, 1 sum, 2 min, 3 max, 4 mean, 5 std
__global__ void __launch_bounds__(AG_THREADS)
k_finalize(int kind, int ddof, int ignore_nulls, int sum_is_int, int val_is_signed, uint64_t g,
           const unsigned long long* __restrict__ rows, AccArrays a, const double* __restrict__ m2,
           uint64_t* __restrict__ out, uint8_t* __restrict__ valid_out) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < g; i += stride) {
    if (kind == 0) {
      out[i] = rows[i];
      valid_out[i] = 1;
      continue;
    }
    uint64_t c = a.cnt[i];
    // AbsMax accumulates row by row in the reference, and its null test (pandas.isnull,
    // null_aggregate.py:269-276) takes NaN for a null: only the other values count
    if (kind == 6 && !sum_is_int) c -= a.nan[i] < c ? a.nan[i] : c;
    const bool ok = c > 0 && (ignore_nulls || c == rows[i]);
    valid_out[i] = ok ? 1 : 0;
    uint64_t r = 0;
    if (ok) {
      const double s = sum_is_int ? (double)((const long long*)a.sum)[i] : ((const double*)a.sum)[i];
      if (kind == 1) {
        r = ((const uint64_t*)a.sum)[i];
      } else if (kind == 2 || kind == 3) {
        const uint64_t e = kind == 2 ? a.minv[i] : a.maxv[i];
        if (sum_is_int) {
          r = val_is_signed ? (e ^ 0x8000000000000000ull) : e;
        } else {
          // Arrow's min_max skips NaN; a group of NaNs only yields NaN
          const bool none = a.nan[i] >= c;
          r = (uint64_t)__double_as_longlong(none ? __longlong_as_double(0x7FF8000000000000ll) : dec_double(e));
        }
      } else if (kind == 4) {
        r = (uint64_t)__double_as_longlong(s / (double)c);
      } else if (kind == 6) {
        // AbsMax (python/ray/data/aggregate.py:330-354): max(0, |x|) over the group's values =
        // the larger magnitude of its min and max (both skip NaN, which counts as a null here)
        if (sum_is_int) {
          if (val_is_signed) {
            const long long mn = (long long)(a.minv[i] ^ 0x8000000000000000ull);
            const long long mx = (long long)(a.maxv[i] ^ 0x8000000000000000ull);
            const uint64_t an = mn < 0 ? (uint64_t)(-(mn + 1)) + 1ull : (uint64_t)mn;
            const uint64_t ax = mx < 0 ? (uint64_t)(-(mx + 1)) + 1ull : (uint64_t)mx;
            r = an > ax ? an : ax;
          } else {
            r = a.maxv[i];
          }
        } else {
          r = (uint64_t)__double_as_longlong(fmax(fabs(dec_double(a.minv[i])), fabs(dec_double(a.maxv[i]))));
        }
      } else {
        double v = 0.0;
        if (c >= 2) v = sqrt(m2[i] / ((double)c - (double)ddof));
        r = (uint64_t)__double_as_longlong(v);
      }
    }
    out[i] = r;
  }
}

// group keys (radix keys, ascending) -> output key column: integers -> int64, floats ->
// double (the reference rebuilds rows from Python scalars: _internal/numpy_support.py)
__global__ void __launch_bounds__(AG_THREADS)
k_group_keys_out(const uint64_t* __restrict__ norm, int key_type, uint64_t g, uint64_t* __restrict__ out) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < g; i += stride) {
    const uint64_t k = norm[i];
    uint64_t r;
    switch (key_type) {
      case XCHG_I8: r = (uint64_t)(int64_t)(int8_t)((uint8_t)k ^ 0x80u); break;
      case XCHG_I16: r = (uint64_t)(int64_t)(int16_t)((uint16_t)k ^ 0x8000u); break;
      case XCHG_I32: r = (uint64_t)(int64_t)(int32_t)((uint32_t)k ^ 0x80000000u); break;
      case XCHG_I64: r = k ^ 0x8000000000000000ull; break;
      case XCHG_F32: {
        const uint32_t t = (uint32_t)k;
        const uint32_t b = (t & 0x80000000u) ? (t & 0x7FFFFFFFu) : ~t;
        const float f = (t == 0xFFFFFFFEu) ? __int_as_float(0x7FC00000) : __uint_as_float(b);
        r = (uint64_t)__double_as_longlong((double)f);
        break;
      }
      case XCHG_F64:
        r = (k == 0xFFFFFFFFFFFFFFFEull) ? 0x7FF8000000000000ull
                                         : (uint64_t)__double_as_longlong(dec_double(k));
        break;
      default: r = k; break;  // unsigned
    }
    out[i] = r;
  }
}

__global__ void k_add_u64(unsigned long long* dst, const unsigned long long* __restrict__ src, uint64_t n) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) dst[i] += src[i];
}

__global__ void k_fill_u64(unsigned long long* p, uint64_t n, unsigned long long v) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) p[i] = v;
}

template <typename T>
void launch_acc(const uint32_t* rgid, const void* vals, const uint8_t* validity, int64_t vbit,
                uint64_t n, uint64_t g, AccArrays a, int want_minmax, int want_cnt, cudaStream_t s) {
  if (rgid && g > AG_DIRECT_MIN_GROUPS) {
    XCHG_LAUNCH(XK_AGGREGATE, s, k_acc_values_direct<T><<<ag_grid(n), AG_THREADS, 0, s>>>(
        rgid, (const T*)vals, validity, vbit, n, a, want_minmax, want_cnt));
  } else {
    XCHG_LAUNCH(XK_AGGREGATE, s, k_acc_values<T><<<ag_grid(n), AG_THREADS, 0, s>>>(
        rgid, (const T*)vals, validity, vbit, n, a, want_minmax));
  }
}
template <typename T>
void launch_m2(const uint32_t* rgid, const void* vals, const uint8_t* validity, int64_t vbit,
               uint64_t n, const unsigned long long* cnt, const void* sum, int sum_is_int, double* m2,
               cudaStream_t s) {
  XCHG_LAUNCH(XK_AGGREGATE, s, k_acc_m2<T><<<ag_grid(n), AG_THREADS, 0, s>>>(
      rgid, (const T*)vals, validity, vbit, n, cnt, sum, sum_is_int, m2));
}

AccArrays arrays_of(uint64_t* stats, uint64_t g) {
  AccArrays a;
  a.cnt = (unsigned long long*)stats;
  a.sum = stats + g;
  a.minv = (unsigned long long*)(stats + 2 * g);
  a.maxv = (unsigned long long*)(stats + 3 * g);
  a.nan = (unsigned long long*)(stats + 4 * g);
  return a;
}

// ---- dense (direct-address) grouping ---------------------------------------------------
// When the radix keys of the valid rows span a small range [kmin, kmin + range), the
// groups need no hash table, no sort and no slot indirection: a presence byte per key
// value, a scan of the presence bytes (rank of every present value = its group id, already
// in key order) and one lookup per row.  GroupBy on an int32 key with 10 M distinct values
// over 250 M rows: ~5 ms instead of ~22 ms for build + compact + sort + assign + lookup.
constexpr int DG_TILE = 4096;  // presence bytes per CTA in the scan (256 threads x 16)

__global__ void __launch_bounds__(AG_THREADS)
k_key_minmax(const uint64_t* __restrict__ norm, const uint8_t* __restrict__ validity, int64_t vbit,
             uint64_t n, unsigned long long* out /* [0] min, [1] max, [2] valid rows */) {
  __shared__ unsigned long long smn[AG_THREADS / 32], smx[AG_THREADS / 32], scn[AG_THREADS / 32];
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  unsigned long long mn = EMPTY, mx = 0, cn = 0;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    if (validity && !bit_is_set(validity, (int64_t)i + vbit)) continue;
    const unsigned long long k = norm[i];
    mn = k < mn ? k : mn;
    mx = k > mx ? k : mx;
    ++cn;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const unsigned long long a = __shfl_xor_sync(0xffffffffu, mn, o);
    const unsigned long long b = __shfl_xor_sync(0xffffffffu, mx, o);
    cn += __shfl_xor_sync(0xffffffffu, cn, o);
    mn = a < mn ? a : mn;
    mx = b > mx ? b : mx;
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) { smn[warp] = mn; smx[warp] = mx; scn[warp] = cn; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < AG_THREADS / 32; ++w) {
      mn = smn[w] < mn ? smn[w] : mn;
      mx = smx[w] > mx ? smx[w] : mx;
      cn += scn[w];
    }
    if (cn) {
      atomicMin(&out[0], mn);
      atomicMax(&out[1], mx);
      atomicAdd(&out[2], cn);
    }
  }
}

__global__ void __launch_bounds__(AG_THREADS)
k_dense_mark(const uint64_t* __restrict__ norm, const uint8_t* __restrict__ validity, int64_t vbit,
             uint64_t n, uint64_t kmin, uint8_t* __restrict__ present, uint32_t* null_present) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    if (validity && !bit_is_set(validity, (int64_t)i + vbit)) {
      *null_present = 1;
      continue;
    }
    present[norm[i] - kmin] = 1;  // benign race: every writer stores the same byte
  }
}

// Row count per key value instead of a presence byte: the same pass then also yields the
// groups' row counts (Count, and the rows column of partial rows), which would otherwise
// cost a second atomic pass over the rows (k_acc_rows_direct).
__global__ void __launch_bounds__(AG_THREADS)
k_dense_count(const uint64_t* __restrict__ norm, const uint8_t* __restrict__ validity, int64_t vbit,
              uint64_t n, uint64_t kmin, uint32_t* __restrict__ counts, uint32_t* null_rows) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  const uint64_t n_up = (n + 31) & ~31ull;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_up; i += stride) {
    const bool active = i < n;
    const bool valid = active && (validity ? bit_is_set(validity, (int64_t)i + vbit) : true);
    const unsigned nulls = __ballot_sync(0xffffffffu, active && !valid);
    if (nulls && (threadIdx.x & 31) == 0) atomicAdd(null_rows, (uint32_t)__popc(nulls));
    if (valid) atomicAdd(&counts[norm[i] - kmin], 1u);
  }
}

__global__ void __launch_bounds__(AG_THREADS)
k_dense_presence(const uint32_t* __restrict__ counts, uint64_t range, uint8_t* __restrict__ present) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < range; i += stride)
    present[i] = counts[i] != 0u;
}

// tile_sums[t] = number of present values in tile t
__global__ void __launch_bounds__(256)
k_dense_tile_sums(const uint8_t* __restrict__ present, uint64_t range, uint32_t* __restrict__ tile_sums) {
  __shared__ uint32_t ws[8];
  const uint64_t base = (uint64_t)blockIdx.x * DG_TILE + (uint64_t)threadIdx.x * 16;
  uint32_t c = 0;
  if (base + 16 <= range) {
    const uint4 v = *reinterpret_cast<const uint4*>(present + base);
    c = __popc(v.x) + __popc(v.y) + __popc(v.z) + __popc(v.w);  // bytes are 0 or 1
  } else {
    for (uint64_t j = base; j < range && j < base + 16; ++j) c += present[j];
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
  if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = c;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t t = 0;
    for (int w = 0; w < 8; ++w) t += ws[w];
    tile_sums[blockIdx.x] = t;
  }
}

// one CTA: exclusive scan of the tile sums in place; total -> *total_out
__global__ void __launch_bounds__(1024)
k_dense_scan_sums(uint32_t* tile_sums, uint32_t ntiles, unsigned long long* total_out) {
  __shared__ uint32_t ws[32];
  __shared__ uint32_t carry;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (uint32_t b = 0; b < ntiles; b += 1024) {
    const uint32_t i = b + threadIdx.x;
    const uint32_t c = i < ntiles ? tile_sums[i] : 0u;
    uint32_t x = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= o) x += y;
    }
    if (lane == 31) ws[warp] = x;
    __syncthreads();
    if (warp == 0) {
      uint32_t w = ws[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t y = __shfl_up_sync(0xffffffffu, w, o);
        if (lane >= o) w += y;
      }
      ws[lane] = w;  // inclusive over warps
    }
    __syncthreads();
    const uint32_t before = carry + (warp ? ws[warp - 1] : 0u) + x - c;
    if (i < ntiles) tile_sums[i] = before;
    __syncthreads();
    if (threadIdx.x == 1023) carry = before + c;
    __syncthreads();
  }
  if (threadIdx.x == 0) *total_out = carry;
}

// rank[d] = number of present values below d (the group id of value d when present);
// group_keys[rank[d]] = kmin + d for every present d: the groups' radix keys in key order
__global__ void __launch_bounds__(256)
k_dense_emit(const uint8_t* __restrict__ present, uint64_t range, uint64_t kmin,
             const uint32_t* __restrict__ tile_excl, uint32_t* __restrict__ rank,
             uint64_t* __restrict__ group_keys, const uint32_t* __restrict__ counts,
             uint64_t* __restrict__ group_rows) {
  __shared__ uint32_t ws[8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint64_t base = (uint64_t)blockIdx.x * DG_TILE + (uint64_t)threadIdx.x * 16;
  uint8_t b[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) b[j] = 0;
  if (base + 16 <= range) {
    *reinterpret_cast<uint4*>(b) = *reinterpret_cast<const uint4*>(present + base);
  } else {
    for (int j = 0; j < 16; ++j)
      if (base + j < range) b[j] = present[base + j];
  }
  uint32_t c = 0;
#pragma unroll
  for (int j = 0; j < 16; ++j) c += b[j];
  uint32_t x = c;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
    if (lane >= o) x += y;
  }
  if (lane == 31) ws[warp] = x;
  __syncthreads();
  uint32_t r = tile_excl[blockIdx.x] + x - c;
  for (int w = 0; w < warp; ++w) r += ws[w];
#pragma unroll
  for (int j = 0; j < 16; ++j) {
    if (base + j < range) {
      rank[base + j] = r;
      if (b[j]) {
        group_keys[r] = kmin + base + j;
        if (group_rows) group_rows[r] = counts[base + j];
      }
      r += b[j];
    }
  }
}

__global__ void __launch_bounds__(AG_THREADS)
k_dense_lookup(const uint64_t* __restrict__ norm, const uint8_t* __restrict__ validity, int64_t vbit,
               uint64_t n, uint64_t kmin, const uint32_t* __restrict__ rank, uint32_t null_gid,
               uint32_t* __restrict__ rgid) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const bool valid = validity ? bit_is_set(validity, (int64_t)i + vbit) : true;
    rgid[i] = valid ? ld_random(rank + (norm[i] - kmin)) : null_gid;
  }
}

// ---- grouping of SORTED rows (any key: several columns, strings) ---------------------------
// The general path: rows are sorted by the key columns (stable chained K2/K3 sorts, string
// keys included), every key column marks the sorted positions where it changes, and a scan
// of the marks numbers the groups -- the run detection of ArrowBlockAccessor.combine
// (arrow_block.py:478-497, keys_equal: NaN == NaN, null its own group) and of
// _get_block_boundaries (python/ray/data/block.py) as data-parallel kernels.

// flags[i] |= sorted row i differs from sorted row i-1 in this column (flags[0] = 1).
// norm: radix keys in SORTED order; validity is indexed by original row id through perm.
__global__ void __launch_bounds__(AG_THREADS)
k_group_flags_fixed(const uint64_t* __restrict__ norm, const uint8_t* __restrict__ validity, int64_t vbit,
                    const uint32_t* __restrict__ perm, uint64_t n, uint8_t* __restrict__ flags) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    bool diff = i == 0;
    if (!diff) {
      diff = norm[i] != norm[i - 1];
      if (!diff && validity)
        diff = bit_is_set(validity, (int64_t)(perm ? perm[i] : (uint32_t)i) + vbit) !=
               bit_is_set(validity, (int64_t)(perm ? perm[i - 1] : (uint32_t)(i - 1)) + vbit);
    }
    if (diff) flags[i] = 1;
  }
}

// The same for a variable-width column (int64 offsets + bytes): equal = same validity and,
// when valid, same length and bytes.
__global__ void __launch_bounds__(AG_THREADS)
k_group_flags_bytes(const int64_t* __restrict__ offsets, const uint8_t* __restrict__ data,
                    const uint8_t* __restrict__ validity, int64_t vbit,
                    const uint32_t* __restrict__ perm, uint64_t n, uint8_t* __restrict__ flags) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    bool diff = i == 0;
    if (!diff) {
      const uint32_t a = perm ? perm[i - 1] : (uint32_t)(i - 1), b = perm ? perm[i] : (uint32_t)i;
      const bool va = validity ? bit_is_set(validity, (int64_t)a + vbit) : true;
      const bool vb = validity ? bit_is_set(validity, (int64_t)b + vbit) : true;
      if (va != vb) {
        diff = true;
      } else if (va) {
        const int64_t a0 = offsets[a], a1 = offsets[a + 1], b0 = offsets[b], b1 = offsets[b + 1];
        if (a1 - a0 != b1 - b0) {
          diff = true;
        } else {
          for (int64_t j = 0; j < a1 - a0; ++j)
            if (data[a0 + j] != data[b0 + j]) {
              diff = true;
              break;
            }
        }
      }
    }
    if (diff) flags[i] = 1;
  }
}

// group id of every ORIGINAL row and the original row id of every group's first row
__global__ void __launch_bounds__(256)
k_groups_emit(const uint8_t* __restrict__ flags, uint64_t n, const uint32_t* __restrict__ tile_excl,
              const uint32_t* __restrict__ perm, uint32_t* __restrict__ row_gid,
              uint32_t* __restrict__ group_first) {
  __shared__ uint32_t ws[8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint64_t base = (uint64_t)blockIdx.x * DG_TILE + (uint64_t)threadIdx.x * 16;
  uint8_t b[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) b[j] = (base + j < n) ? flags[base + j] : 0;
  uint32_t c = 0;
#pragma unroll
  for (int j = 0; j < 16; ++j) c += b[j];
  uint32_t x = c;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
    if (lane >= o) x += y;
  }
  if (lane == 31) ws[warp] = x;
  __syncthreads();
  uint32_t r = tile_excl[blockIdx.x] + x - c;  // marks before this thread's first position
  for (int w = 0; w < warp; ++w) r += ws[w];
#pragma unroll
  for (int j = 0; j < 16; ++j) {
    if (base + j < n) {
      r += b[j];
      const uint32_t row = perm ? perm[base + j] : (uint32_t)(base + j);
      row_gid[row] = r - 1;
      if (b[j]) group_first[r - 1] = row;
    }
  }
}

__global__ void __launch_bounds__(AG_THREADS)
k_group_first_rows(const uint32_t* __restrict__ rgid, uint64_t n, uint32_t* __restrict__ first) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    atomicMin(&first[rgid[i]], (uint32_t)i);
}

bool key_type_is_int(int t) { return t >= XCHG_I8 && t <= XCHG_U64; }

}  // namespace

extern "C" int xchg_hash_table_bytes(uint64_t capacity, uint64_t* bytes_out) {
  XCHG_REQUIRE(bytes_out != nullptr, "bytes_out is NULL");
  XCHG_REQUIRE(capacity >= 2 && (capacity & (capacity - 1)) == 0, "capacity must be a power of two");
  if (capacity > (1ull << 31)) {
    xchg_set_error("hash table capacity %llu exceeds 2^31 slots", (unsigned long long)capacity);
    return XCHG_ERR_LIMIT;
  }
  // keys[cap] | gid_of_slot[cap + 2] | special_present[2] + overflow + pad | counter
  *bytes_out = capacity * 8 + xchg_align_up((capacity + 2) * 4, 256) + 256 + 256;
  return XCHG_OK;
}

extern "C" int xchg_hash_groups(const uint64_t* norm_keys, const uint8_t* validity,
                                int64_t validity_bit_offset, uint64_t n, uint64_t capacity,
                                void* table_ws, uint64_t table_ws_bytes, uint32_t* slot_out,
                                uint64_t* group_keys_out, uint32_t* group_slots_out,
                                uint64_t* num_groups_out, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  XCHG_REQUIRE(table_ws && slot_out && group_keys_out && group_slots_out && num_groups_out,
               "NULL argument");
  uint64_t need = 0;
  int rc = xchg_hash_table_bytes(capacity, &need);
  if (rc) return rc;
  if (table_ws_bytes < need) {
    xchg_set_error("hash table workspace too small: have %llu, need %llu",
                   (unsigned long long)table_ws_bytes, (unsigned long long)need);
    return XCHG_ERR_WORKSPACE;
  }
  XCHG_REQUIRE(n <= XCHG_MAX_ROWS, "n exceeds XCHG_MAX_ROWS");
  XCHG_REQUIRE(n == 0 || norm_keys, "NULL keys");
  uint8_t* base = (uint8_t*)table_ws;
  unsigned long long* table = (unsigned long long*)base;
  uint32_t* misc = (uint32_t*)(base + capacity * 8 + xchg_align_up((capacity + 2) * 4, 256));
  unsigned long long* counter = (unsigned long long*)((uint8_t*)misc + 256);
  XCHG_CUDA_TRY(cudaMemsetAsync(table, 0xFF, capacity * 8, s));
  XCHG_CUDA_TRY(cudaMemsetAsync(misc, 0, 512, s));
  if (n) {
    XCHG_LAUNCH(XK_AGGREGATE, s, k_hash_build<<<ag_grid(n), AG_THREADS, 0, s>>>(
        norm_keys, validity, validity_bit_offset, n, table, capacity, slot_out, misc, misc + 2));
    XCHG_LAUNCH_CHECK();
  }
  XCHG_LAUNCH(XK_AGGREGATE, s, k_hash_compact<<<ag_grid(capacity), AG_THREADS, 0, s>>>(
      table, capacity, group_keys_out, group_slots_out, counter));
  XCHG_LAUNCH_CHECK();
  XCHG_LAUNCH(XK_AGGREGATE, s, k_hash_append_specials<<<1, 32, 0, s>>>(misc, capacity, group_keys_out,
                                                                    group_slots_out, counter));
  XCHG_LAUNCH_CHECK();
  uint32_t host_misc[4] = {0, 0, 0, 0};
  unsigned long long host_count = 0;
  XCHG_CUDA_TRY(cudaMemcpyAsync(host_misc, misc, sizeof(host_misc), cudaMemcpyDeviceToHost, s));
  XCHG_CUDA_TRY(cudaMemcpyAsync(&host_count, counter, 8, cudaMemcpyDeviceToHost, s));
  XCHG_CUDA_TRY(cudaStreamSynchronize(s));
  if (host_misc[2]) {
    xchg_set_error("hash table of %llu slots overflowed; retry with a larger capacity",
                   (unsigned long long)capacity);
    return XCHG_ERR_LIMIT;
  }
  *num_groups_out = host_count;
  return XCHG_OK;
}

extern "C" int xchg_key_minmax(const uint64_t* norm_keys, const uint8_t* validity,
                               int64_t validity_bit_offset, uint64_t n, void* scratch32,
                               uint64_t* out3_host, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  XCHG_REQUIRE(scratch32 && out3_host, "NULL scratch / out");
  out3_host[0] = EMPTY;
  out3_host[1] = 0;
  out3_host[2] = 0;
  if (n == 0) return XCHG_OK;
  XCHG_REQUIRE(norm_keys != nullptr, "NULL keys");
  unsigned long long init[3] = {EMPTY, 0ull, 0ull};
  XCHG_CUDA_TRY(cudaMemcpyAsync(scratch32, init, sizeof(init), cudaMemcpyHostToDevice, s));
  XCHG_LAUNCH(XK_AGGREGATE, s, k_key_minmax<<<ag_grid(n), AG_THREADS, 0, s>>>(
      norm_keys, validity, validity_bit_offset, n, (unsigned long long*)scratch32));
  XCHG_LAUNCH_CHECK();
  XCHG_CUDA_TRY(cudaMemcpyAsync(out3_host, scratch32, 24, cudaMemcpyDeviceToHost, s));
  XCHG_CUDA_TRY(cudaStreamSynchronize(s));
  return XCHG_OK;
}

namespace {
struct DenseWs {
  uint8_t* present;
  uint32_t* counts;
  uint32_t* rank;
  uint32_t* tile_sums;
  uint32_t* null_present;
  unsigned long long* total;
  uint64_t bytes;
};
DenseWs dense_carve(void* base, uint64_t range) {
  DenseWs w;
  uint64_t off = 0;
  auto take = [&](uint64_t b) {
    uint64_t o = off;
    off = xchg_align_up(off + b, 256);
    return (uint8_t*)base + o;
  };
  const uint64_t ntiles = (range + DG_TILE - 1) / DG_TILE;
  w.present = take(range + 16);
  w.counts = (uint32_t*)take(range * 4);
  w.rank = (uint32_t*)take(range * 4);
  w.tile_sums = (uint32_t*)take((ntiles + 1) * 4);
  w.null_present = (uint32_t*)take(64);
  w.total = (unsigned long long*)take(64);
  w.bytes = off;
  return w;
}
}  // namespace

extern "C" int xchg_dense_groups_bytes(uint64_t range, uint64_t* bytes_out) {
  XCHG_REQUIRE(bytes_out != nullptr, "bytes_out is NULL");
  if (range == 0 || range > XCHG_DENSE_MAX_RANGE) {
    xchg_set_error("dense key range %llu outside [1, %llu]", (unsigned long long)range,
                   (unsigned long long)XCHG_DENSE_MAX_RANGE);
    return XCHG_ERR_LIMIT;
  }
  *bytes_out = dense_carve(nullptr, range).bytes + 256;
  return XCHG_OK;
}

extern "C" int xchg_dense_groups(const uint64_t* norm_keys, const uint8_t* validity,
                                 int64_t validity_bit_offset, uint64_t n, uint64_t key_min,
                                 uint64_t range, void* workspace, uint64_t workspace_bytes,
                                 uint32_t* row_gid_out, uint64_t* group_keys_out,
                                 uint64_t* group_rows_out, uint64_t* num_groups_out, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  XCHG_REQUIRE(workspace && row_gid_out && group_keys_out && num_groups_out, "NULL argument");
  XCHG_REQUIRE(n <= XCHG_MAX_ROWS, "n exceeds XCHG_MAX_ROWS");
  uint64_t need = 0;
  int rc = xchg_dense_groups_bytes(range, &need);
  if (rc) return rc;
  if (workspace_bytes < need) {
    xchg_set_error("dense grouping workspace too small: have %llu, need %llu",
                   (unsigned long long)workspace_bytes, (unsigned long long)need);
    return XCHG_ERR_WORKSPACE;
  }
  *num_groups_out = 0;
  if (n == 0) return XCHG_OK;
  XCHG_REQUIRE(norm_keys != nullptr, "NULL keys");
  DenseWs w = dense_carve((void*)xchg_align_up((uint64_t)(uintptr_t)workspace, 256), range);
  const uint32_t ntiles = (uint32_t)((range + DG_TILE - 1) / DG_TILE);
  XCHG_CUDA_TRY(cudaMemsetAsync(w.present, 0, range + 16, s));
  XCHG_CUDA_TRY(cudaMemsetAsync(w.null_present, 0, 64, s));
  if (group_rows_out) {
    XCHG_CUDA_TRY(cudaMemsetAsync(w.counts, 0, range * 4, s));
    XCHG_LAUNCH(XK_AGGREGATE, s, k_dense_count<<<ag_grid(n), AG_THREADS, 0, s>>>(
        norm_keys, validity, validity_bit_offset, n, key_min, w.counts, w.null_present));
    XCHG_LAUNCH(XK_AGGREGATE, s, k_dense_presence<<<ag_grid(range), AG_THREADS, 0, s>>>(w.counts, range,
                                                                                   w.present));
  } else {
    XCHG_LAUNCH(XK_AGGREGATE, s, k_dense_mark<<<ag_grid(n), AG_THREADS, 0, s>>>(
        norm_keys, validity, validity_bit_offset, n, key_min, w.present, w.null_present));
  }
  XCHG_LAUNCH(XK_AGGREGATE, s, k_dense_tile_sums<<<ntiles, 256, 0, s>>>(w.present, range, w.tile_sums));
  XCHG_LAUNCH(XK_AGGREGATE, s, k_dense_scan_sums<<<1, 1024, 0, s>>>(w.tile_sums, ntiles, w.total));
  XCHG_LAUNCH(XK_AGGREGATE, s, k_dense_emit<<<ntiles, 256, 0, s>>>(w.present, range, key_min, w.tile_sums,
                                                               w.rank, group_keys_out, w.counts,
                                                               group_rows_out));
  XCHG_LAUNCH_CHECK();
  unsigned long long total = 0;
  uint32_t has_null = 0;
  XCHG_CUDA_TRY(cudaMemcpyAsync(&total, w.total, 8, cudaMemcpyDeviceToHost, s));
  XCHG_CUDA_TRY(cudaMemcpyAsync(&has_null, w.null_present, 4, cudaMemcpyDeviceToHost, s));
  XCHG_CUDA_TRY(cudaStreamSynchronize(s));
  if (has_null) {  // the null group goes last, with the sentinel key the hash path gives it
    const unsigned long long e = EMPTY;
    XCHG_CUDA_TRY(cudaMemcpyAsync(group_keys_out + total, &e, 8, cudaMemcpyHostToDevice, s));
    if (group_rows_out) {  // has_null is the number of null rows on the counting path
      const unsigned long long nr = has_null;
      XCHG_CUDA_TRY(cudaMemcpyAsync(group_rows_out + total, &nr, 8, cudaMemcpyHostToDevice, s));
    }
  }
  XCHG_LAUNCH(XK_AGGREGATE, s, k_dense_lookup<<<ag_grid(n), AG_THREADS, 0, s>>>(
      norm_keys, validity, validity_bit_offset, n, key_min, w.rank, (uint32_t)total, row_gid_out));
  XCHG_LAUNCH_CHECK();
  *num_groups_out = total + (has_null ? 1 : 0);
  return XCHG_OK;
}

extern "C" int xchg_group_flags_fixed(const uint64_t* norm_sorted, const uint8_t* validity,
                                      int64_t validity_bit_offset, const uint32_t* perm, uint64_t n,
                                      uint8_t* flags, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  if (n == 0) return XCHG_OK;
  XCHG_REQUIRE(norm_sorted && flags, "NULL keys / flags");
  XCHG_LAUNCH(XK_AGGREGATE, s, k_group_flags_fixed<<<ag_grid(n), AG_THREADS, 0, s>>>(
      norm_sorted, validity, validity_bit_offset, perm, n, flags));
  XCHG_LAUNCH_CHECK();
  return XCHG_OK;
}

extern "C" int xchg_group_flags_bytes(const int64_t* offsets, const uint8_t* data, const uint8_t* validity,
                                      int64_t validity_bit_offset, const uint32_t* perm, uint64_t n,
                                      uint8_t* flags, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  if (n == 0) return XCHG_OK;
  XCHG_REQUIRE(offsets && flags, "NULL offsets / flags");
  XCHG_LAUNCH(XK_AGGREGATE, s, k_group_flags_bytes<<<ag_grid(n), AG_THREADS, 0, s>>>(
      offsets, data, validity, validity_bit_offset, perm, n, flags));
  XCHG_LAUNCH_CHECK();
  return XCHG_OK;
}

extern "C" int xchg_group_first_rows(const uint32_t* row_gid, uint64_t n, uint64_t num_groups,
                                     uint32_t* first_out, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  if (num_groups == 0) return XCHG_OK;
  XCHG_REQUIRE(first_out != nullptr, "NULL first_out");
  XCHG_CUDA_TRY(cudaMemsetAsync(first_out, 0xFF, num_groups * 4, s));
  if (n == 0) return XCHG_OK;
  XCHG_REQUIRE(row_gid != nullptr, "NULL row_gid");
  XCHG_LAUNCH(XK_AGGREGATE, s, k_group_first_rows<<<ag_grid(n), AG_THREADS, 0, s>>>(row_gid, n, first_out));
  XCHG_LAUNCH_CHECK();
  return XCHG_OK;
}

extern "C" int xchg_groups_from_flags_bytes(uint64_t n, uint64_t* bytes_out) {
  XCHG_REQUIRE(bytes_out != nullptr, "bytes_out is NULL");
  XCHG_REQUIRE(n <= XCHG_MAX_ROWS, "n exceeds XCHG_MAX_ROWS");
  *bytes_out = xchg_align_up(((n + DG_TILE - 1) / DG_TILE + 1) * 4, 256) + 512;
  return XCHG_OK;
}

extern "C" int xchg_groups_from_flags(const uint8_t* flags, const uint32_t* perm, uint64_t n,
                                      void* workspace, uint64_t workspace_bytes, uint32_t* row_gid_out,
                                      uint32_t* group_first_out, uint64_t* num_groups_out, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  XCHG_REQUIRE(num_groups_out != nullptr, "NULL num_groups_out");
  *num_groups_out = 0;
  if (n == 0) return XCHG_OK;
  XCHG_REQUIRE(flags && workspace && row_gid_out && group_first_out, "NULL argument");
  uint64_t need = 0;
  int rc = xchg_groups_from_flags_bytes(n, &need);
  if (rc) return rc;
  if (workspace_bytes < need) {
    xchg_set_error("group scan workspace too small: have %llu, need %llu",
                   (unsigned long long)workspace_bytes, (unsigned long long)need);
    return XCHG_ERR_WORKSPACE;
  }
  uint8_t* base = (uint8_t*)xchg_align_up((uint64_t)(uintptr_t)workspace, 256);
  const uint32_t ntiles = (uint32_t)((n + DG_TILE - 1) / DG_TILE);
  uint32_t* tile_sums = (uint32_t*)base;
  unsigned long long* total = (unsigned long long*)(base + xchg_align_up((uint64_t)(ntiles + 1) * 4, 256));
  XCHG_LAUNCH(XK_AGGREGATE, s, k_dense_tile_sums<<<ntiles, 256, 0, s>>>(flags, n, tile_sums));
  XCHG_LAUNCH(XK_AGGREGATE, s, k_dense_scan_sums<<<1, 1024, 0, s>>>(tile_sums, ntiles, total));
  XCHG_LAUNCH(XK_AGGREGATE, s, k_groups_emit<<<ntiles, 256, 0, s>>>(flags, n, tile_sums, perm, row_gid_out,
                                                                group_first_out));
  XCHG_LAUNCH_CHECK();
  unsigned long long g = 0;
  XCHG_CUDA_TRY(cudaMemcpyAsync(&g, total, 8, cudaMemcpyDeviceToHost, s));
  XCHG_CUDA_TRY(cudaStreamSynchronize(s));
  *num_groups_out = g;
  return XCHG_OK;
}

extern "C" int xchg_group_ids(const uint32_t* order, const uint32_t* group_slots, uint64_t num_groups,
                              const uint32_t* slot, uint64_t n, uint64_t capacity, void* table_ws,
                              uint32_t* row_gid_out, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  if (num_groups == 0 || n == 0) return XCHG_OK;
  XCHG_REQUIRE(order && group_slots && slot && table_ws && row_gid_out, "NULL argument");
  uint32_t* gid_of_slot = (uint32_t*)((uint8_t*)table_ws + capacity * 8);
  XCHG_LAUNCH(XK_AGGREGATE, s, k_assign_gid<<<ag_grid(num_groups), AG_THREADS, 0, s>>>(
      order, group_slots, num_groups, gid_of_slot));
  XCHG_LAUNCH_CHECK();
  XCHG_LAUNCH(XK_AGGREGATE, s, k_lookup_gid<<<ag_grid(n), AG_THREADS, 0, s>>>(slot, gid_of_slot, n,
                                                                            row_gid_out));
  XCHG_LAUNCH_CHECK();
  return XCHG_OK;
}

extern "C" int xchg_agg_stats_init(uint64_t* stats, double* m2, uint64_t num_groups, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  if (num_groups == 0) return XCHG_OK;
  XCHG_REQUIRE(stats != nullptr, "NULL stats");
  XCHG_CUDA_TRY(cudaMemsetAsync(stats, 0, num_groups * 8 * 5, s));
  XCHG_LAUNCH(XK_AGGREGATE, s, k_fill_u64<<<ag_grid(num_groups), AG_THREADS, 0, s>>>(
      (unsigned long long*)(stats + 2 * num_groups), num_groups, (unsigned long long)EMPTY));
  XCHG_LAUNCH_CHECK();
  if (m2) XCHG_CUDA_TRY(cudaMemsetAsync(m2, 0, num_groups * 8, s));
  return XCHG_OK;
}

extern "C" int xchg_agg_count_rows(const uint32_t* row_gid, uint64_t n, uint64_t num_groups,
                                   uint64_t* rows, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  if (n == 0) return XCHG_OK;
  XCHG_REQUIRE(rows != nullptr, "NULL rows");
  if (row_gid && num_groups > AG_DIRECT_MIN_GROUPS) {
    XCHG_LAUNCH(XK_AGGREGATE, s, k_acc_rows_direct<<<ag_grid(n), AG_THREADS, 0, s>>>(
        row_gid, n, (unsigned long long*)rows));
  } else {
    XCHG_LAUNCH(XK_AGGREGATE, s, k_acc_rows<<<ag_grid(n), AG_THREADS, 0, s>>>(
        row_gid, n, (unsigned long long*)rows));
  }
  XCHG_LAUNCH_CHECK();
  return XCHG_OK;
}

extern "C" int xchg_agg_accumulate(const uint32_t* row_gid, const void* values, int value_type,
                                   const uint8_t* validity, int64_t validity_bit_offset, uint64_t n,
                                   uint64_t num_groups, uint64_t* stats, int want_minmax,
                                   const uint64_t* rows, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  if (n == 0) return XCHG_OK;
  XCHG_REQUIRE(values && stats, "NULL values / stats");
  AccArrays a = arrays_of(stats, num_groups);
  // a column without nulls: its per-group value counts are the row counts the caller
  // already has -- one atomic per row less
  const bool cnt_from_rows = rows != nullptr && validity == nullptr && row_gid != nullptr &&
                             num_groups > AG_DIRECT_MIN_GROUPS;
  const int want_cnt = cnt_from_rows ? 0 : 1;
  if (cnt_from_rows) {
    XCHG_LAUNCH(XK_AGGREGATE, s, k_add_u64<<<ag_grid(num_groups), AG_THREADS, 0, s>>>(
        a.cnt, (const unsigned long long*)rows, num_groups));
    XCHG_LAUNCH_CHECK();
  }
  switch (value_type) {
    case XCHG_I8: launch_acc<int8_t>(row_gid, values, validity, validity_bit_offset, n, num_groups, a, want_minmax, want_cnt, s); break;
    case XCHG_I16: launch_acc<int16_t>(row_gid, values, validity, validity_bit_offset, n, num_groups, a, want_minmax, want_cnt, s); break;
    case XCHG_I32: launch_acc<int32_t>(row_gid, values, validity, validity_bit_offset, n, num_groups, a, want_minmax, want_cnt, s); break;
    case XCHG_I64: launch_acc<long long>(row_gid, values, validity, validity_bit_offset, n, num_groups, a, want_minmax, want_cnt, s); break;
    case XCHG_U8: launch_acc<uint8_t>(row_gid, values, validity, validity_bit_offset, n, num_groups, a, want_minmax, want_cnt, s); break;
    case XCHG_U16: launch_acc<uint16_t>(row_gid, values, validity, validity_bit_offset, n, num_groups, a, want_minmax, want_cnt, s); break;
    case XCHG_U32: launch_acc<uint32_t>(row_gid, values, validity, validity_bit_offset, n, num_groups, a, want_minmax, want_cnt, s); break;
    case XCHG_U64: launch_acc<unsigned long long>(row_gid, values, validity, validity_bit_offset, n, num_groups, a, want_minmax, want_cnt, s); break;
    case XCHG_F32: launch_acc<float>(row_gid, values, validity, validity_bit_offset, n, num_groups, a, want_minmax, want_cnt, s); break;
    case XCHG_F64: launch_acc<double>(row_gid, values, validity, validity_bit_offset, n, num_groups, a, want_minmax, want_cnt, s); break;
    default:
      xchg_set_error("unknown value_type %d", value_type);
      return XCHG_ERR_INVALID;
  }
  XCHG_LAUNCH_CHECK();
  return XCHG_OK;
}

extern "C" int xchg_agg_accumulate_m2(const uint32_t* row_gid, const void* values, int value_type,
                                      const uint8_t* validity, int64_t validity_bit_offset,
                                      uint64_t n, uint64_t num_groups, const uint64_t* stats,
                                      double* m2, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  if (n == 0) return XCHG_OK;
  XCHG_REQUIRE(values && stats && m2, "NULL values / stats / m2");
  const unsigned long long* cnt = (const unsigned long long*)stats;
  const void* sum = stats + num_groups;
  const int isint = key_type_is_int(value_type) ? 1 : 0;
  switch (value_type) {
    case XCHG_I8: launch_m2<int8_t>(row_gid, values, validity, validity_bit_offset, n, cnt, sum, isint, m2, s); break;
    case XCHG_I16: launch_m2<int16_t>(row_gid, values, validity, validity_bit_offset, n, cnt, sum, isint, m2, s); break;
    case XCHG_I32: launch_m2<int32_t>(row_gid, values, validity, validity_bit_offset, n, cnt, sum, isint, m2, s); break;
    case XCHG_I64: launch_m2<long long>(row_gid, values, validity, validity_bit_offset, n, cnt, sum, isint, m2, s); break;
    case XCHG_U8: launch_m2<uint8_t>(row_gid, values, validity, validity_bit_offset, n, cnt, sum, isint, m2, s); break;
    case XCHG_U16: launch_m2<uint16_t>(row_gid, values, validity, validity_bit_offset, n, cnt, sum, isint, m2, s); break;
    case XCHG_U32: launch_m2<uint32_t>(row_gid, values, validity, validity_bit_offset, n, cnt, sum, isint, m2, s); break;
    case XCHG_U64: launch_m2<unsigned long long>(row_gid, values, validity, validity_bit_offset, n, cnt, sum, isint, m2, s); break;
    case XCHG_F32: launch_m2<float>(row_gid, values, validity, validity_bit_offset, n, cnt, sum, isint, m2, s); break;
    case XCHG_F64: launch_m2<double>(row_gid, values, validity, validity_bit_offset, n, cnt, sum, isint, m2, s); break;
    default:
      xchg_set_error("unknown value_type %d", value_type);
      return XCHG_ERR_INVALID;
  }
  XCHG_LAUNCH_CHECK();
  return XCHG_OK;
}

extern "C" int xchg_agg_merge_rows(const uint32_t* row_gid, const int64_t* partial_rows, uint64_t n,
                                   uint64_t* rows, void* stream) {
  if (n == 0) return XCHG_OK;
  XCHG_REQUIRE(partial_rows && rows, "NULL argument");
  XCHG_LAUNCH(XK_AGGREGATE, (cudaStream_t)stream, k_merge_rows<<<ag_grid(n), AG_THREADS, 0, (cudaStream_t)stream>>>(
      row_gid, (const long long*)partial_rows, n, (unsigned long long*)rows));
  XCHG_LAUNCH_CHECK();
  return XCHG_OK;
}

extern "C" int xchg_agg_merge_states(const uint32_t* row_gid, const uint64_t* partial_states,
                                     uint64_t n, int sum_is_int, uint64_t num_groups, uint64_t* stats,
                                     double* m2, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  if (n == 0) return XCHG_OK;
  XCHG_REQUIRE(partial_states && stats, "NULL argument");
  AccArrays a = arrays_of(stats, num_groups);
  XCHG_LAUNCH(XK_AGGREGATE, s, k_merge_states<<<ag_grid(n), AG_THREADS, 0, s>>>(
      row_gid, partial_states, n, sum_is_int, a));
  XCHG_LAUNCH_CHECK();
  if (m2) {
    XCHG_LAUNCH(XK_AGGREGATE, s, k_merge_m2<<<ag_grid(n), AG_THREADS, 0, s>>>(
        row_gid, partial_states, n, sum_is_int, a.cnt, a.sum, m2));
    XCHG_LAUNCH_CHECK();
  }
  return XCHG_OK;
}

extern "C" int xchg_agg_pack_states(uint64_t num_groups, const uint64_t* stats, const double* m2,
                                    uint64_t* partial_states_out, void* stream) {
  if (num_groups == 0) return XCHG_OK;
  XCHG_REQUIRE(stats && partial_states_out, "NULL argument");
  XCHG_LAUNCH(XK_AGGREGATE, (cudaStream_t)stream, k_pack_states<<<ag_grid(num_groups), AG_THREADS, 0, (cudaStream_t)stream>>>(
      num_groups, arrays_of((uint64_t*)stats, num_groups), m2, partial_states_out));
  XCHG_LAUNCH_CHECK();
  return XCHG_OK;
}

extern "C" int xchg_agg_finalize(int kind, int ddof, int ignore_nulls, int value_type,
                                 uint64_t num_groups, const uint64_t* rows, const uint64_t* stats,
                                 const double* m2, void* out, uint8_t* valid_out, void* stream) {
  if (num_groups == 0) return XCHG_OK;
  XCHG_REQUIRE(kind >= 0 && kind <= 6, "kind must be 0..6");
  XCHG_REQUIRE(rows && out && valid_out, "NULL rows / out / valid_out");
  XCHG_REQUIRE(kind == 0 || stats, "NULL stats");
  XCHG_REQUIRE(kind != 5 || m2, "std needs m2");
  AccArrays a = arrays_of((uint64_t*)stats, num_groups);
  const int isint = key_type_is_int(value_type) ? 1 : 0;
  const int is_signed = value_type >= XCHG_I8 && value_type <= XCHG_I64;
  XCHG_LAUNCH(XK_AGGREGATE, (cudaStream_t)stream, k_finalize<<<ag_grid(num_groups), AG_THREADS, 0, (cudaStream_t)stream>>>(
      kind, ddof, ignore_nulls, isint, is_signed, num_groups, (const unsigned long long*)rows, a, m2,
      (uint64_t*)out, valid_out));
  XCHG_LAUNCH_CHECK();
  return XCHG_OK;
}

extern "C" int xchg_group_keys_out(const uint64_t* sorted_norm_keys, int key_type, uint64_t num_groups,
                                   void* out, void* stream) {
  if (num_groups == 0) return XCHG_OK;
  XCHG_REQUIRE(sorted_norm_keys && out, "NULL argument");
  XCHG_LAUNCH(XK_AGGREGATE, (cudaStream_t)stream, k_group_keys_out<<<ag_grid(num_groups), AG_THREADS, 0, (cudaStream_t)stream>>>(
      sorted_norm_keys, key_type, num_groups, (uint64_t*)out));
  XCHG_LAUNCH_CHECK();
  return XCHG_OK;
}
