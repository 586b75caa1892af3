// K2 (key normalisation + digit histograms) and K3 (stable onesweep LSD radix sort
// of 64-bit radix keys with 32-bit row ids) for sm_100a.
//
// Replaces pyarrow.compute.sort_indices as the reference calls it from
// python/ray/data/_internal/arrow_ops/transform_pyarrow.py:41 and :546.
//
// Data layout in HBM: SoA -- one u64 normalised key array and one u32 row-id
// array, ping-ponged between two buffer sets; 8-bit digits, only the digit places
// whose histogram has more than one non-empty bin are run (range(10M) in int64
// costs 3 passes, random 63-bit keys 8).  Every pass is ONE kernel: each CTA takes
// an 8192-row tile (dynamic tile id), ranks it with warp match_any, publishes its
// per-digit counts and resolves its global offsets by decoupled look-back, stages
// the tile in shared memory in digit order and writes it out in coalesced runs.
// Algorithmic bytes per pass: 24 B/row (12 read + 12 written).
#include <algorithm>
#include <atomic>
#include <cstdlib>
#include <type_traits>

#include "common.cuh"
#include "keynorm.cuh"

namespace {

constexpr int BINS = 256;
constexpr int KEY_PLACES = 8;
constexpr int NUM_PLACES = 9;  // 8 key digit places + the validity place
// CTA shapes of the onesweep pass: 512 threads x 2 CTAs/SM (8192-row tiles) or 256
// threads x 4 CTAs/SM (4096-row tiles).  Same registers and shared memory per SM; the
// smaller CTA keeps every thread busy in the 256-wide scan / look-back phases and lets
// four tiles per SM interleave their barrier-separated phases.
constexpr int RS_MIN_TILE = 256 * 8;
constexpr int RS_DEFAULT_THREADS = 256;
constexpr int RS_DEFAULT_ITEMS = 16;
constexpr int RS_BATCH = 4;  // items ranked together (independent match / atomic / shuffle chains)
constexpr size_t rs_smem(int threads, int items) { return (size_t)threads * items * 12 + (BINS + BINS + 32) * 4; }
// MODE 2 (destination partition) keeps its thresholds behind the pass's own shared memory
constexpr size_t RS_RULE_SMEM = (size_t)BINS * 8 + BINS;
// CTAs per SM by items per thread (register budget: 2 regs per key + rank + ~24)
constexpr int rs_min_blocks(int threads, int items) {
  return (items >= 16 ? 1024 : (items >= 12 ? 1280 : 1536)) / threads;
}

constexpr int LB_WINDOW = 8;  // predecessors fetched per look-back round trip

constexpr uint32_t FLAG_MASK = 3u << 30;
constexpr uint32_t FLAG_LOCAL = 1u << 30;
constexpr uint32_t FLAG_INCL = 2u << 30;
constexpr uint32_t VAL_MASK = (1u << 30) - 1;

// ------------------------------------------------------------------ K2 -----

template <typename T>
__global__ void __launch_bounds__(256)
k_normalize_hist(const T* __restrict__ keys, const uint32_t* __restrict__ perm,
                 const uint8_t* __restrict__ validity, int64_t vbit, uint32_t n, int desc,
                 uint64_t* __restrict__ out, uint32_t* __restrict__ ghist) {
  __shared__ uint32_t sh[NUM_PLACES * BINS];
  if (ghist) {
    for (int j = threadIdx.x; j < NUM_PLACES * BINS; j += blockDim.x) sh[j] = 0;
    __syncthreads();
  }
  const unsigned lane = threadIdx.x & 31;
  const uint32_t stride = gridDim.x * blockDim.x;
  const uint32_t n_up = (n + 31u) & ~31u;  // whole warps stay in the loop for the votes
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n_up; i += stride) {
    const bool active = i < n;
    uint64_t key = 0;
    bool valid = true;
    if (active) {
      const uint32_t row = perm ? perm[i] : i;
      valid = validity ? bit_is_set(validity, (int64_t)row + vbit) : true;
      key = valid ? norm_value<T>(keys[row], desc != 0) : null_key<T>();
      out[i] = key;
    }
    if (ghist) {
      const unsigned act = __ballot_sync(0xffffffffu, active);
      const int first = __ffs(act) - 1;
#pragma unroll
      for (int p = 0; p < NUM_PLACES; ++p) {
        const uint32_t d = (p < KEY_PLACES) ? (uint32_t)(key >> (8 * p)) & 255u : (valid ? 0u : 1u);
        const uint32_t d0 = __shfl_sync(0xffffffffu, d, first);
        const bool uni = __all_sync(0xffffffffu, !active || d == d0);
        if (uni) {
          if ((int)lane == first) atomicAdd(&sh[p * BINS + d0], (uint32_t)__popc(act));
        } else if (active) {
          atomicAdd(&sh[p * BINS + d], 1u);
        }
      }
    }
  }
  if (ghist) {
    __syncthreads();
    for (int j = threadIdx.x; j < NUM_PLACES * BINS; j += blockDim.x)
      if (sh[j]) atomicAdd(&ghist[j], sh[j]);
  }
}

// One CTA: exclusive scan of each place's 256-bin histogram and the "place is
// non-trivial" flags (more than one non-empty bin).
__global__ void __launch_bounds__(BINS)
k_scan_hist(const uint32_t* __restrict__ ghist, uint32_t* __restrict__ gexcl,
            uint32_t* __restrict__ active, uint32_t n) {
  __shared__ uint32_t wsum[BINS / 32];
  __shared__ uint32_t s_full;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int p = 0; p < NUM_PLACES; ++p) {
    if (tid == 0) s_full = 0;
    __syncthreads();
    const uint32_t c = ghist[p * BINS + tid];
    if (c == n) s_full = 1;
    uint32_t x = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= o) x += y;
    }
    if (lane == 31) wsum[warp] = x;
    __syncthreads();
    uint32_t base = 0;
    for (int w = 0; w < warp; ++w) base += wsum[w];
    gexcl[p * BINS + tid] = base + x - c;
    if (tid == 0) active[p] = s_full ? 0u : 1u;
    __syncthreads();
  }
}

// First look at a key column that is sorted straight from its raw values (no nulls, no
// incoming permutation): OR and AND of all radix keys -- the keys differ exactly in the bits
// of OR ^ AND, so the digit places are laid over [lowest differing bit, highest differing
// bit] instead of over fixed bytes -- and the histograms of the digit places in `ks` (the
// place the first pass runs on, guessed from the previous sort of this key type; all places
// for an all-LSD sort).  Nothing is written per row: the passes read the raw column again
// and normalise in registers (RAW passes).
struct KeyShifts {
  int n;                  // digit places to build histograms for (slot h -> ghist + h * BINS)
  int shift[KEY_PLACES];  // any bit position: a digit is (key >> shift) & 255
};

template <typename T>
__global__ void __launch_bounds__(256)
k_key_scan(const T* __restrict__ keys, uint32_t n, int desc, KeyShifts ks, int want_orand,
           unsigned long long* __restrict__ orand, uint32_t* __restrict__ ghist) {
  __shared__ uint32_t sh[KEY_PLACES * BINS];
  __shared__ unsigned long long s_or[8], s_and[8];
  for (int j = threadIdx.x; j < KEY_PLACES * BINS; j += blockDim.x) sh[j] = 0;
  __syncthreads();
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t stride = gridDim.x * blockDim.x;
  const uint32_t n_up = (n + 31u) & ~31u;
  uint64_t acc_or = 0, acc_and = ~0ull;
  // one shared-memory atomic per key and place; a warp whose 32 keys share the digit (sorted
  // or narrow inputs) adds once instead -- that test is only made while it keeps succeeding
  bool try_uniform = true;
  // four keys per thread and trip, all loads issued before the first is used: with one load
  // in flight per thread the pass ran at 2.4 TB/s (30 % of DRAM bandwidth, latency-bound)
  constexpr int KS_UNROLL = 4;
  for (uint32_t i0 = blockIdx.x * blockDim.x + threadIdx.x; i0 < n_up; i0 += stride * KS_UNROLL) {
    uint64_t kq[KS_UNROLL];
#pragma unroll
    for (int u = 0; u < KS_UNROLL; ++u) {
      const uint64_t iu = (uint64_t)i0 + (uint64_t)u * stride;
      kq[u] = iu < n ? norm_value<T>(keys[iu], desc != 0) : 0;
    }
#pragma unroll
    for (int u = 0; u < KS_UNROLL; ++u) {
    const uint64_t i = (uint64_t)i0 + (uint64_t)u * stride;
    if (i >= n_up) break;  // warp-uniform: a warp's 32 indices share their 32-aligned block
    const bool active = i < n;
    const uint64_t key = kq[u];
    if (active) {
      acc_or |= key;
      acc_and &= key;
    }
    if (try_uniform) {
      const unsigned act = __ballot_sync(0xffffffffu, active);
      const int first = __ffs(act) - 1;
      bool all_uni = true;
      for (int p = 0; p < ks.n; ++p) {
        const uint32_t d = (uint32_t)(key >> ks.shift[p]) & 255u;
        const uint32_t d0 = __shfl_sync(0xffffffffu, d, first);
        const bool uni = __all_sync(0xffffffffu, !active || d == d0);
        if (uni) {
          if ((int)lane == first) atomicAdd(&sh[p * BINS + d0], (uint32_t)__popc(act));
        } else {
          all_uni = false;
          if (active) atomicAdd(&sh[p * BINS + d], 1u);
        }
      }
      try_uniform = all_uni;  // warp-uniform by construction
    } else if (active) {
      for (int p = 0; p < ks.n; ++p) atomicAdd(&sh[p * BINS + ((uint32_t)(key >> ks.shift[p]) & 255u)], 1u);
    }
    }
  }
  if (want_orand) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      acc_or |= __shfl_xor_sync(0xffffffffu, acc_or, o);
      acc_and &= __shfl_xor_sync(0xffffffffu, acc_and, o);
    }
    if (lane == 0) {
      s_or[warp] = acc_or;
      s_and[warp] = acc_and;
    }
  }
  __syncthreads();
  if (want_orand && threadIdx.x == 0) {
    uint64_t o = 0, a = ~0ull;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) {
      o |= s_or[w];
      a &= s_and[w];
    }
    atomicOr(&orand[0], (unsigned long long)o);
    atomicAnd(&orand[1], (unsigned long long)a);
  }
  for (int p = 0; p < ks.n; ++p) {
    for (int j = threadIdx.x; j < BINS; j += blockDim.x)
      if (sh[p * BINS + j]) atomicAdd(&ghist[p * BINS + j], sh[p * BINS + j]);
  }
}

// ------------------------------------------------------------------ K3 -----

template <int MODE>
__device__ __forceinline__ uint32_t digit_of(uint64_t key, uint32_t val, int shift,
                                             const uint8_t* validity, int64_t vbit) {
  if constexpr (MODE == 0) {
    return (uint32_t)(key >> shift) & 255u;
  } else {
    // validity place: valid rows -> bin 0, null rows -> bin 1; padding (val ==
    // UINT32_MAX) -> bin 255 so that it ranks after every real row.
    if (val == 0xFFFFFFFFu) return 255u;
    return bit_is_set(validity, (int64_t)val + vbit) ? 0u : 1u;
  }
}

// Lanes of the warp holding the same 8-bit digit.  Eight ballots and a select/and per
// bit instead of one match.any: MATCH.ANY runs on the SM's single ADU pipe at ~2 cycles
// per distinct value in the warp (~64 cycles per instruction on random digits; ncu showed
// that pipe 70-74 % busy and bounding both the onesweep pass and the local sort at
// ~1.2 ms per 125 M rows), whereas the votes and logic ops spread over the four
// schedulers' ordinary pipes.
__device__ __forceinline__ unsigned match_digit(uint32_t d) {
  // Per bit: test, ballot, select the all-ones / zero mask, one LOP3 for m &= v ^ mask
  // (ptxas emits six instructions per bit for this; a hand-written four-instruction PTX
  // sequence assembles in isolation but fails register allocation inside these kernels).
  unsigned m = 0xffffffffu;
#pragma unroll
  for (int b = 0; b < 8; ++b) {
    const bool p = (d & (1u << b)) != 0u;
    const unsigned v = __ballot_sync(0xffffffffu, p);
    m &= v ^ (p ? 0u : 0xffffffffu);
  }
  return m;
}

// The same in four instructions per bit instead of the six ptxas emits for the C++ form
// (bit test into a predicate, ballot, select of the all-ones / zero mask, one LOP3):
// 15 % fewer instructions in a onesweep pass.  ptxas fails register allocation ("register
// count of '7'", the predicate file) when 16 of these per thread sit inside the local
// sort's pass loop, so the 16-row-per-thread local sorts keep match_digit.
__device__ __forceinline__ unsigned match_digit_asm(uint32_t d) {
  unsigned m = 0xffffffffu;
#define XCHG_MATCH_BIT(MASK)                                                                     \
  asm volatile("{\n\t.reg .pred p;\n\t.reg .b32 t, v, x;\n\t"                                   \
               "and.b32 t, %1, " #MASK ";\n\tsetp.ne.u32 p, t, 0;\n\t"                           \
               "vote.sync.ballot.b32 v, p, 0xffffffff;\n\tselp.b32 x, 0, 0xffffffff, p;\n\t"      \
               "lop3.b32 %0, %0, v, x, 0x60;\n\t}"                                               \
               : "+r"(m)                                                                         \
               : "r"(d));
  XCHG_MATCH_BIT(1) XCHG_MATCH_BIT(2) XCHG_MATCH_BIT(4) XCHG_MATCH_BIT(8)
  XCHG_MATCH_BIT(16) XCHG_MATCH_BIT(32) XCHG_MATCH_BIT(64) XCHG_MATCH_BIT(128)
#undef XCHG_MATCH_BIT
  return m;
}

// Bump of a warp-PRIVATE shared-memory counter by the lanes whose predicate is set:
// returns the old value.  Plain load + add + store, predicated (no divergent branch):
// within one call the active lanes hold distinct digits, hence distinct addresses, and
// no other warp touches this warp's counter row, so no atomic is needed -- a shared
// atomic costs ~2 cycles per active lane on this part (ATOMS, 60 cycles per warp and
// item with random digits: it bounded both the onesweep pass and the local sort), the
// load/store pair a few cycles per warp.  Successive calls are ordered by the
// __syncwarp() the caller issues between them.
__device__ __forceinline__ uint32_t warp_counter_bump(uint32_t saddr, uint32_t v, bool pred) {
  uint32_t old = 0;
  asm volatile(
      "{\n\t.reg .pred p;\n\t.reg .u32 t;\n\tsetp.ne.u32 p, %3, 0;\n\t"
      "@p ld.shared.u32 %0, [%1];\n\t@p add.u32 t, %0, %2;\n\t@p st.shared.u32 [%1], t;\n\t}"
      : "+r"(old)
      : "r"(saddr), "r"(v), "r"((uint32_t)pred)
      : "memory");
  return old;
}

// One segment of an MSD level: rows [start, start + count) share every digit above the
// level's place; its tiles are tbase .. tbase + ceil(count / tile) - 1.
struct Seg {
  uint32_t start, count, tbase, pad;
};

// SEG = false: one LSD / top-level pass over the whole array.  SEG = true: the same pass
// run independently inside every segment of an MSD level (tile -> segment through
// tile_seg; digit offsets, look-back and output positions are relative to the segment).
// RAW passes (the first pass over a column): the keys are read in their own type and
// normalised in registers instead of being read from a materialised radix-key array.
// MODE 2: the "digit" of a row is its destination under the thresholds (keynorm.cuh): the
// same stable one-pass partition, by range instead of by byte.
struct RawKeys {
  const void* keys;       // raw key column (RAW passes), key_type elements
  int key_type, desc;
  const uint64_t* thr;    // MODE 2: nthr thresholds in sort order + their strict flags
  const uint8_t* strict;
  int nthr;
};

template <typename T, int RS_ITEMS>
__device__ __forceinline__ void load_raw_keys(const void* keys_, bool desc, uint32_t warp_base, uint32_t n,
                                              bool full, uint64_t (&key)[RS_ITEMS]) {
  const T* keys = reinterpret_cast<const T*>(keys_);
  if (full) {
#pragma unroll
    for (int i = 0; i < RS_ITEMS; ++i) key[i] = norm_value<T>(keys[warp_base + 32 * i], desc);
  } else {
#pragma unroll
    for (int i = 0; i < RS_ITEMS; ++i) {
      const uint32_t idx = warp_base + 32 * i;
      key[i] = idx < n ? norm_value<T>(keys[idx], desc) : 0xFFFFFFFFFFFFFFFFull;
    }
  }
}

template <int MODE, int RS_THREADS, int RS_ITEMS, bool SEG, bool RAW = false>
__global__ void __launch_bounds__(RS_THREADS, rs_min_blocks(RS_THREADS, RS_ITEMS))
k_onesweep(const uint64_t* __restrict__ keys_in_, const uint32_t* __restrict__ vals_in_,
           uint64_t* __restrict__ keys_out_, uint32_t* __restrict__ vals_out_, uint32_t n_,
           int shift, const uint32_t* __restrict__ gexcl_, uint32_t* status,
           uint32_t* tile_counter, const uint8_t* __restrict__ validity, int64_t vbit,
           const Seg* __restrict__ segs, const uint32_t* __restrict__ tile_seg, RawKeys raw) {
  constexpr int RS_WARPS = RS_THREADS / 32;
  constexpr int RS_TILE = RS_THREADS * RS_ITEMS;
  extern __shared__ __align__(16) uint8_t smem[];
  uint64_t* skeys = reinterpret_cast<uint64_t*>(smem);
  uint32_t* svals = reinterpret_cast<uint32_t*>(smem + (size_t)RS_TILE * 8);
  uint32_t* whist = reinterpret_cast<uint32_t*>(smem);  // aliases skeys until ranks are final
  uint32_t* sbase = reinterpret_cast<uint32_t*>(smem + (size_t)RS_TILE * 12);
  uint32_t* sstart = sbase + BINS;
  uint32_t* smisc = sstart + BINS;  // [0..7] warp sums of the digit scan, [16] tile id
  uint64_t* thr_s = reinterpret_cast<uint64_t*>(smisc + 32);  // MODE 2 only
  uint8_t* strict_s = reinterpret_cast<uint8_t*>(thr_s + BINS);

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

  if (tid == 0) smisc[16] = atomicAdd(tile_counter, 1u);
  for (int j = tid; j < RS_WARPS * BINS; j += RS_THREADS) whist[j] = 0;
  if constexpr (MODE == 2) {
    for (int j = tid; j < raw.nthr; j += RS_THREADS) {
      thr_s[j] = raw.thr[j];
      strict_s[j] = raw.strict[j];
    }
  }
  __syncthreads();
  const uint32_t tile = smisc[16];
  const uint64_t* keys_in = keys_in_;
  const uint32_t* vals_in = vals_in_;
  uint64_t* keys_out = keys_out_;
  uint32_t* vals_out = vals_out_;
  const uint32_t* gexcl = gexcl_;
  uint32_t n = n_, tile0 = 0;
  if constexpr (SEG) {
    const uint32_t sj = tile_seg[tile];
    const Seg sg = segs[sj];
    keys_in += sg.start;
    vals_in += sg.start;
    keys_out += sg.start;
    vals_out += sg.start;
    gexcl += (size_t)sj * BINS;
    n = sg.count;
    tile0 = sg.tbase;
  }
  const uint32_t tile_base = (tile - tile0) * (uint32_t)RS_TILE;
  const uint32_t tile_items = min((uint32_t)RS_TILE, n - tile_base);
  const bool full = tile_items == (uint32_t)RS_TILE;  // every tile but the last
  const uint32_t warp_base = tile_base + warp * (32 * RS_ITEMS) + lane;

  // ---- load (warp-striped: item i of lane l sits at warp_base + 32*i) ----
  uint64_t key[RS_ITEMS];
  uint32_t val[RS_ITEMS];  // only live in MODE 1 (the compiler drops it otherwise)
  if constexpr (RAW) {
    const bool dsc = raw.desc != 0;
    switch (raw.key_type) {  // uniform over the grid
      case XCHG_I8: load_raw_keys<int8_t>(raw.keys, dsc, warp_base, n, full, key); break;
      case XCHG_I16: load_raw_keys<int16_t>(raw.keys, dsc, warp_base, n, full, key); break;
      case XCHG_I32: load_raw_keys<int32_t>(raw.keys, dsc, warp_base, n, full, key); break;
      case XCHG_I64: load_raw_keys<int64_t>(raw.keys, dsc, warp_base, n, full, key); break;
      case XCHG_U8: load_raw_keys<uint8_t>(raw.keys, dsc, warp_base, n, full, key); break;
      case XCHG_U16: load_raw_keys<uint16_t>(raw.keys, dsc, warp_base, n, full, key); break;
      case XCHG_U32: load_raw_keys<uint32_t>(raw.keys, dsc, warp_base, n, full, key); break;
      case XCHG_U64: load_raw_keys<uint64_t>(raw.keys, dsc, warp_base, n, full, key); break;
      case XCHG_F32: load_raw_keys<float>(raw.keys, dsc, warp_base, n, full, key); break;
      default: load_raw_keys<double>(raw.keys, dsc, warp_base, n, full, key); break;
    }
  } else if (full) {
#pragma unroll
    for (int i = 0; i < RS_ITEMS; ++i) key[i] = keys_in[warp_base + 32 * i];
  } else {
#pragma unroll
    for (int i = 0; i < RS_ITEMS; ++i) {
      const uint32_t idx = warp_base + 32 * i;
      key[i] = idx < n ? keys_in[idx] : 0xFFFFFFFFFFFFFFFFull;
    }
  }
#pragma unroll
  for (int i = 0; i < RS_ITEMS; ++i) {
    const uint32_t idx = warp_base + 32 * i;
    if constexpr (MODE == 1) val[i] = idx < n ? (vals_in ? vals_in[idx] : idx) : 0xFFFFFFFFu;
    else val[i] = 0u;
  }

  // MODE 2: destinations are found once (a binary search each) and kept, four to a register;
  // rows past the end of the last tile rank behind every real row (bin 255)
  uint32_t dg[MODE == 2 ? RS_ITEMS / 4 : 1];
  if constexpr (MODE == 2) {
#pragma unroll
    for (int i = 0; i < RS_ITEMS; ++i) {
      const uint32_t idx = warp_base + 32 * i;
      const uint32_t d = (full || idx < n) ? dest_of(key[i], thr_s, strict_s, raw.nthr) : 255u;
      if ((i & 3) == 0) dg[i >> 2] = d;
      else dg[i >> 2] |= d << (8 * (i & 3));
    }
  }
  auto digit_at = [&](int i) -> uint32_t {
    if constexpr (MODE == 2) return (dg[i >> 2] >> (8 * (i & 3))) & 255u;
    else return digit_of<MODE>(key[i], val[i], shift, validity, vbit);
  };

  // ---- rank inside the warp, in (i, lane) = memory order: stable ----
  // RS_BATCH items at a time: match masks first (independent), then the leaders' counter
  // bumps on the warp's private counter row one item after the other (ordered by
  // __syncwarp, so equal digits receive increasing bases in item order), then the
  // broadcasts.  The bumps are plain shared-memory load/add/store, not atomics (see
  // warp_counter_bump).
  const uint32_t myhist_s = (uint32_t)__cvta_generic_to_shared(whist + warp * BINS);
  uint32_t* myhist = whist + warp * BINS;
  uint32_t rank2[RS_ITEMS / 2];  // two 16-bit in-tile ranks per register (tile <= 8192 rows)
  const unsigned lt_mask = lanemask_lt();
#pragma unroll
  for (int h = 0; h < RS_ITEMS; h += RS_BATCH) {
    uint32_t d[RS_BATCH];
    unsigned m[RS_BATCH];
#pragma unroll
    for (int u = 0; u < RS_BATCH; ++u) {
      d[u] = digit_at(h + u);
      m[u] = match_digit_asm(d[u]);
    }
    uint32_t prev[RS_BATCH];
#pragma unroll
    for (int u = 0; u < RS_BATCH; ++u) {  // the lowest lane of every same-digit set adds the set's size
      prev[u] = warp_counter_bump(myhist_s + d[u] * 4, (uint32_t)__popc(m[u]), (m[u] & lt_mask) == 0);
      __syncwarp();
    }
#pragma unroll
    for (int u = 0; u < RS_BATCH; ++u) {
      const uint32_t r = __shfl_sync(0xffffffffu, prev[u], __ffs(m[u]) - 1) + (uint32_t)__popc(m[u] & lt_mask);
      if (((h + u) & 1) == 0) rank2[(h + u) >> 1] = r;
      else rank2[(h + u) >> 1] |= r << 16;
    }
  }
  __syncthreads();

  // ---- per digit: exclusive offsets across warps, tile totals, digit scan ----
  uint32_t total = 0, incl = 0;
  if (tid < BINS) {
#pragma unroll
    for (int w0 = 0; w0 < RS_WARPS; w0 += 4) {  // four loads in flight, few live registers
      uint32_t c[4];
#pragma unroll
      for (int w = 0; w < 4; ++w) c[w] = whist[(w0 + w) * BINS + tid];
#pragma unroll
      for (int w = 0; w < 4; ++w) {
        whist[(w0 + w) * BINS + tid] = total;
        total += c[w];
      }
    }
    incl = total;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t y = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += y;
    }
    if (lane == 31) smisc[warp] = incl;
  }
  __syncthreads();
  if (tid < BINS) {
    uint32_t start = incl - total;
    for (int w = 0; w < warp; ++w) start += smisc[w];
    sstart[tid] = start;

    // decoupled look-back on this digit's column of the tile-status matrix
    uint32_t cnt = total;
    if (tid == BINS - 1) cnt -= (uint32_t)RS_TILE - tile_items;  // padding sits in bin 255
    volatile uint32_t* st = status + (size_t)tile * BINS + tid;
    uint32_t excl = 0;
    if (tile == tile0) {
      *st = FLAG_INCL | cnt;
    } else {
      *st = FLAG_LOCAL | cnt;
      // Windowed walk: the LB_WINDOW predecessors' words are fetched with independent
      // loads (one L2 round trip for the window, not one per tile), then consumed in
      // order up to the first word that is an inclusive prefix or not published yet.
      int64_t t = (int64_t)tile - 1;
      bool done = false;
      while (!done) {
        uint32_t w[LB_WINDOW];
#pragma unroll
        for (int j = 0; j < LB_WINDOW; ++j) {
          const int64_t tj = t - j;
          w[j] = tj >= (int64_t)tile0 ? *(volatile uint32_t*)(status + (size_t)tj * BINS + tid)
                                      : (2u << 30);  // before the first tile: an inclusive prefix of zero
        }
#pragma unroll
        for (int j = 0; j < LB_WINDOW; ++j) {
          if (done) break;
          const uint32_t sj = w[j];
          if ((sj & FLAG_MASK) == 0) break;  // not published yet: re-read from here
          excl += sj & VAL_MASK;
          --t;
          if (sj & FLAG_INCL) done = true;
        }
      }
      *st = FLAG_INCL | (excl + cnt);
    }
    sbase[tid] = gexcl[tid] + excl - start;
  }
  __syncthreads();

  // ---- final in-tile ranks (both halves of a register at once: no carry, ranks < 65536) ----
#pragma unroll
  for (int i = 0; i < RS_ITEMS; i += 2) {
    const uint32_t d0 = digit_at(i);
    const uint32_t d1 = digit_at(i + 1);
    rank2[i >> 1] += (myhist[d0] + sstart[d0]) + ((myhist[d1] + sstart[d1]) << 16);
  }
  __syncthreads();  // whist is dead from here: skeys may overwrite it

  // ---- stage the tile in digit order ----
#pragma unroll
  for (int i = 0; i < RS_ITEMS; ++i) skeys[(rank2[i >> 1] >> (16 * (i & 1))) & 0xFFFFu] = key[i];
#pragma unroll
  for (int i = 0; i < RS_ITEMS; ++i) {
    uint32_t v;
    if constexpr (MODE == 1) {
      v = val[i];
    } else {
      const uint32_t idx = warp_base + 32 * i;
      v = (full || idx < n) ? (vals_in ? vals_in[idx] : idx) : 0xFFFFFFFFu;
    }
    svals[(rank2[i >> 1] >> (16 * (i & 1))) & 0xFFFFu] = v;
  }
  __syncthreads();

  // ---- coalesced write-out: consecutive threads, consecutive ranks ----
  if (full) {
#pragma unroll
    for (int i = 0; i < RS_ITEMS; ++i) {
      const uint32_t pos = i * RS_THREADS + tid;
      const uint64_t k = skeys[pos];
      const uint32_t v = svals[pos];
      uint32_t dd;
      if constexpr (MODE == 2) dd = dest_of(k, thr_s, strict_s, raw.nthr);
      else dd = digit_of<MODE>(k, v, shift, validity, vbit);
      const uint32_t g = sbase[dd] + pos;
      keys_out[g] = k;
      vals_out[g] = v;
    }
  } else {
#pragma unroll
    for (int i = 0; i < RS_ITEMS; ++i) {
      const uint32_t pos = i * RS_THREADS + tid;
      if (pos < tile_items) {
        const uint64_t k = skeys[pos];
        const uint32_t v = svals[pos];
        uint32_t dd;
        if constexpr (MODE == 2) dd = dest_of(k, thr_s, strict_s, raw.nthr);
        else dd = digit_of<MODE>(k, v, shift, validity, vbit);
        const uint32_t g = sbase[dd] + pos;
        keys_out[g] = k;
        vals_out[g] = v;
      }
    }
  }
}

// Inverse of norm_value for integer keys: sorted radix keys -> sorted raw keys
// (sequential, instead of a random gather of the key column through the permutation).
template <typename T>
__global__ void __launch_bounds__(256)
k_denormalize(const uint64_t* __restrict__ norm, uint32_t n, int desc, T* __restrict__ out) {
  using U = typename std::make_unsigned<T>::type;
  const uint32_t stride = gridDim.x * blockDim.x;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    U u = (U)norm[i];
    if (desc) u = (U)~u;
    if constexpr (std::is_signed<T>::value) u ^= (U)((U)1 << (sizeof(T) * 8 - 1));
    out[i] = (T)u;
  }
}

__global__ void k_iota_or_copy(const uint32_t* __restrict__ in, uint32_t* __restrict__ out, uint32_t n) {
  const uint32_t stride = gridDim.x * blockDim.x;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) out[i] = in ? in[i] : i;
}

// ------------------------------------------------- MSD levels + local sort -----
//
// Sorting 125 M random 64-bit keys by eight LSD passes moves 8 x 24 B per row through
// HBM.  The hybrid below runs the passes from the TOP place down instead, each level
// splitting every segment (rows equal in all higher places) into 256 children with a
// segmented onesweep pass, and stops moving a segment through HBM as soon as it fits in
// shared memory: a CTA then loads it once, finishes all remaining places with a local
// LSD sort and writes it to its final position.  Random 64-bit keys take two global
// levels (65 536 segments of ~1 900 rows) and one local sort: 3 x 24 + 8 B per row.
// Every pass is stable, so the result is identical to the LSD sort's.

constexpr int LS_CLASSES = 5;
constexpr uint32_t LS_CAP = 8192;  // largest segment a CTA sorts locally
constexpr uint32_t LS_GROUP_MEMBER_MAX = 512;  // children below this size are packed into shared jobs
constexpr uint32_t LS_RUN_LIMIT = 8;  // longest out-of-order run one thread fixes by insertion

struct MsdCounters {
  uint32_t nsmall[LS_CLASSES];
  uint32_t nbig, ntiles, pad0;
};

__host__ __device__ constexpr uint32_t ls_class_cap(int c) { return 512u << c; }

// Histogram of one digit place inside every segment of a level (8 B read per row).
template <int TILE>
__global__ void __launch_bounds__(256)
k_seg_hist(const uint64_t* __restrict__ keys, int shift, const Seg* __restrict__ segs,
           const uint32_t* __restrict__ tile_seg, uint32_t* __restrict__ seg_hist) {
  __shared__ uint32_t sh[BINS];
  const uint32_t tile = blockIdx.x;
  const uint32_t sj = tile_seg[tile];
  const Seg sg = segs[sj];
  const uint32_t base = (tile - sg.tbase) * (uint32_t)TILE;
  const uint32_t items = min((uint32_t)TILE, sg.count - base);
  sh[threadIdx.x] = 0;
  __syncthreads();
  const uint64_t* k = keys + sg.start + base;
  for (uint32_t i = threadIdx.x; i < items; i += 256) atomicAdd(&sh[(uint32_t)(k[i] >> shift) & 255u], 1u);
  __syncthreads();
  if (sh[threadIdx.x]) atomicAdd(&seg_hist[(size_t)sj * BINS + threadIdx.x], sh[threadIdx.x]);
}

// One CTA per segment: exclusive scan of its 256 counts.
__global__ void __launch_bounds__(BINS)
k_seg_scan(const uint32_t* __restrict__ seg_hist, uint32_t* __restrict__ seg_excl) {
  __shared__ uint32_t wsum[BINS / 32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t c = seg_hist[(size_t)blockIdx.x * BINS + tid];
  uint32_t x = c;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
    if (lane >= o) x += y;
  }
  if (lane == 31) wsum[warp] = x;
  __syncthreads();
  uint32_t base = 0;
  for (int w = 0; w < warp; ++w) base += wsum[w];
  seg_excl[(size_t)blockIdx.x * BINS + tid] = base + x - c;
}

// Local-sort job: rows [start, start + count); GROUP set = the rows are several sibling
// children (consecutive digits of this level's place), so the sort also covers that place.
constexpr uint64_t LS_JOB_GROUP = 1ull << 48;
__device__ __forceinline__ uint64_t ls_job(uint32_t start, uint32_t count, bool group) {
  return (uint64_t)start | ((uint64_t)count << 32) | (group ? LS_JOB_GROUP : 0ull);
}

// Children of a level, one thread per parent segment walking its 256 digit counts:
// a child too large for a local sort becomes a segment of the next level and claims its
// tiles; consecutive children of fewer than LS_GROUP_MEMBER_MAX rows are packed greedily
// into local-sort jobs of at most LS_CAP rows (so a parent just above LS_CAP does not
// turn into 256 jobs of a few rows each); the others are jobs of their own.  can_group = 0 (validity place: no key digit tells the children apart)
// keeps every child a job of its own.  parents == NULL: the top-level segment [0, n).
struct SmallLists {
  uint64_t* p[LS_CLASSES];
};

__global__ void __launch_bounds__(32)
k_msd_classify(const Seg* __restrict__ parents, uint32_t nparents, const uint32_t* __restrict__ hist,
               const uint32_t* __restrict__ excl, uint32_t tile_rows, int can_group,
               Seg* __restrict__ big_next, SmallLists lists, MsdCounters* __restrict__ ctr) {
  // one warp per parent: the lanes stage its counts and offsets in shared memory, lane 0
  // walks them
  __shared__ uint32_t h[BINS], x[BINS];
  const uint32_t j = blockIdx.x;
  for (int d = threadIdx.x; d < BINS; d += 32) {
    h[d] = hist[(size_t)j * BINS + d];
    x[d] = excl[(size_t)j * BINS + d];
  }
  __syncwarp();
  if (threadIdx.x != 0) return;
  const uint32_t base = parents ? parents[j].start : 0u;
  auto cls_of = [](uint32_t c) {
    int k = 0;
    while (c > ls_class_cap(k)) ++k;
    return k;
  };
  // two walks over the digit counts: count the jobs / segments / tiles this parent
  // produces, reserve list space with one atomic per list, then write the entries
  uint32_t njob[LS_CLASSES] = {0}, nb = 0, nt = 0;
  uint32_t ojob[LS_CLASSES] = {0}, ob = 0, ot = 0;
  for (int pass = 0; pass < 2; ++pass) {
    uint32_t gstart = 0, gcount = 0, gmembers = 0;
    auto emit = [&](uint32_t st, uint32_t c, bool group) {
      const int cls = cls_of(c);
      if (pass == 0) {
        ++njob[cls];
      } else {
        lists.p[cls][ojob[cls]++] = ls_job(st, c, group);
      }
    };
    for (int d = 0; d < BINS; ++d) {
      const uint32_t c = h[d];
      if (c == 0) continue;
      const uint32_t st = base + x[d];
      const bool big = c > LS_CAP;
      const bool alone = !can_group || c >= LS_GROUP_MEMBER_MAX;  // large enough for a job of its own
      if (gmembers && (big || alone || gcount + c > LS_CAP)) {
        emit(gstart, gcount, gmembers > 1);
        gmembers = 0;
      }
      if (big) {
        const uint32_t tiles = (c + tile_rows - 1) / tile_rows;
        if (pass == 0) {
          ++nb;
          nt += tiles;
        } else {
          big_next[ob++] = Seg{st, c, ot, 0u};
          ot += tiles;
        }
      } else if (alone) {
        emit(st, c, false);
      } else if (gmembers == 0) {
        gstart = st;
        gcount = c;
        gmembers = 1;
      } else {
        gcount += c;
        ++gmembers;
      }
    }
    if (gmembers) emit(gstart, gcount, gmembers > 1);
    if (pass == 0) {
#pragma unroll
      for (int c = 0; c < LS_CLASSES; ++c)
        if (njob[c]) ojob[c] = atomicAdd(&ctr->nsmall[c], njob[c]);
      if (nb) {
        ob = atomicAdd(&ctr->nbig, nb);
        ot = atomicAdd(&ctr->ntiles, nt);
      }
    }
  }
}

__global__ void __launch_bounds__(128)
k_fill_tile_map(const Seg* __restrict__ segs, uint32_t tile_rows, uint32_t* __restrict__ tile_seg) {
  const Seg sg = segs[blockIdx.x];
  const uint32_t nt = (sg.count + tile_rows - 1) / tile_rows;
  for (uint32_t t = threadIdx.x; t < nt; t += blockDim.x) tile_seg[sg.tbase + t] = blockIdx.x;
}

struct LocalPlaces {
  int n;                      // places below the level that produced the jobs
  int shift[KEY_PLACES + 1];  // their shifts, ascending; shift[n] = that level's own place
};

constexpr size_t ls_smem(int threads, int items) {
  return (size_t)threads * items * 12 + (size_t)(2 * (threads / 32) * BINS + BINS + 32) * 4;
}
constexpr int ls_min_blocks(int threads, int items) {
  // registers: 3 per item + ranks + ~28; shared memory: 12 B per row + counters
  return items >= 16 ? (threads >= 512 ? 1 : (threads >= 256 ? 2 : 4))
                     : (threads >= 512 ? 2 : (threads >= 256 ? 4 : 8));
}

// One CTA per small segment: load it, LSD-sort it by the remaining (lower) places in
// shared memory, write it to its final position.  Rows past the segment's end are padded
// with all-ones keys, which stay behind every real row through every stable pass.
template <int LS_THREADS, int LS_ITEMS>
__global__ void __launch_bounds__(LS_THREADS, ls_min_blocks(LS_THREADS, LS_ITEMS))
k_local_sort(const uint64_t* __restrict__ src_keys, const uint32_t* __restrict__ src_vals,
             uint64_t* __restrict__ dst_keys, uint32_t* __restrict__ dst_vals,
             const uint64_t* __restrict__ small, LocalPlaces places) {
  constexpr int WARPS = LS_THREADS / 32;
  constexpr int ROWS = LS_THREADS * LS_ITEMS;
  constexpr int BPT = BINS / LS_THREADS > 0 ? BINS / LS_THREADS : 1;  // bins per scanning thread
  constexpr int SCAN_THREADS = BINS / BPT;
  extern __shared__ __align__(16) uint8_t smem[];
  uint64_t* skeys = reinterpret_cast<uint64_t*>(smem);
  uint32_t* svals = reinterpret_cast<uint32_t*>(smem + (size_t)ROWS * 8);
  uint32_t* whist = reinterpret_cast<uint32_t*>(smem + (size_t)ROWS * 12);  // [WARPS][BINS], zero between passes
  uint32_t* woff = whist + WARPS * BINS;                                    // [WARPS][BINS]
  uint32_t* sstart = woff + WARPS * BINS;                                   // [BINS]
  uint32_t* smisc = sstart + BINS;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint64_t ent = small[blockIdx.x];
  const uint32_t st = (uint32_t)ent, cnt = (uint32_t)(ent >> 32) & 0xFFFFu;
  const int nplaces = places.n + ((ent & LS_JOB_GROUP) ? 1 : 0);  // a group also sorts by shift[n]
  const uint32_t wb = warp * (32 * LS_ITEMS) + lane;

  for (int j = tid; j < WARPS * BINS; j += LS_THREADS) whist[j] = 0;

  uint64_t key[LS_ITEMS];
  uint32_t val[LS_ITEMS];
#pragma unroll
  for (int i = 0; i < LS_ITEMS; ++i) {
    const uint32_t idx = wb + 32 * i;
    key[i] = idx < cnt ? src_keys[st + idx] : 0xFFFFFFFFFFFFFFFFull;
    val[i] = idx < cnt ? src_vals[st + idx] : 0xFFFFFFFFu;
  }
  __syncthreads();

  const uint32_t myhist_s = (uint32_t)__cvta_generic_to_shared(whist + warp * BINS);
  const uint32_t* myoff = woff + warp * BINS;
  const unsigned lt_mask = lanemask_lt();
  constexpr int BATCH = LS_ITEMS < RS_BATCH ? LS_ITEMS : RS_BATCH;

  // With more than two places left, two passes on the TOP two of them nearly finish the
  // job: a segment has at most 4096 rows, so few rows share those 16 bits.  The rows that
  // do form short runs, which one thread each puts in order by insertion (stable:
  // strictly-greater moves only).  A run longer than LS_RUN_LIMIT that is out of order
  // sends the whole segment through the plain LSD passes over every place instead
  // (correct from any stable arrangement of the rows).
  // (a group adds its own place as a third pass: its rows differ there by construction)
  const int qsteps = (ent & LS_JOB_GROUP) ? 3 : 2;
  bool quick = nplaces > qsteps;
  int step = 0, nsteps = quick ? qsteps : nplaces;
  while (true) {
  for (; step < nsteps; ++step) {
    const int shift = quick ? places.shift[nplaces - qsteps + step] : places.shift[step];
    uint32_t rank2[(LS_ITEMS + 1) / 2];
#pragma unroll
    for (int h = 0; h < LS_ITEMS; h += BATCH) {
      uint32_t d[BATCH];
      unsigned m[BATCH];
#pragma unroll
      for (int u = 0; u < BATCH; ++u) {
        d[u] = (uint32_t)(key[h + u] >> shift) & 255u;
        if constexpr (LS_ITEMS <= 8) m[u] = match_digit_asm(d[u]);
        else m[u] = match_digit(d[u]);
      }
      uint32_t prev[BATCH];
#pragma unroll
      for (int u = 0; u < BATCH; ++u) {
        prev[u] = warp_counter_bump(myhist_s + d[u] * 4, (uint32_t)__popc(m[u]), (m[u] & lt_mask) == 0);
        __syncwarp();
      }
#pragma unroll
      for (int u = 0; u < BATCH; ++u) {
        const uint32_t r = __shfl_sync(0xffffffffu, prev[u], __ffs(m[u]) - 1) + (uint32_t)__popc(m[u] & lt_mask);
        if (((h + u) & 1) == 0) rank2[(h + u) >> 1] = r;
        else rank2[(h + u) >> 1] |= r << 16;
      }
    }
    __syncthreads();

    // per digit: offsets of the warps' counts, digit totals, exclusive digit scan; the
    // counters are left at zero for the next pass
    uint32_t total = 0, incl = 0;
    if (tid < SCAN_THREADS) {
      uint32_t tb[BPT];
#pragma unroll
      for (int b = 0; b < BPT; ++b) {
        const int bin = tid * BPT + b;
        uint32_t t = 0;
#pragma unroll
        for (int w = 0; w < WARPS; ++w) {
          const uint32_t c = whist[w * BINS + bin];
          whist[w * BINS + bin] = 0;
          woff[w * BINS + bin] = t;
          t += c;
        }
        tb[b] = t;
        total += t;
      }
      incl = total;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        uint32_t y = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += y;
      }
      if (lane == 31) smisc[warp] = incl;
      // stash the per-thread bin totals for the second half (registers survive the barrier)
      uint32_t run = incl - total;
#pragma unroll
      for (int b = 0; b < BPT; ++b) {
        sstart[tid * BPT + b] = run;  // still lacks the preceding warps' sums
        run += tb[b];
      }
    }
    __syncthreads();
    if (tid < SCAN_THREADS && warp > 0) {
      uint32_t add = 0;
      for (int w = 0; w < warp; ++w) add += smisc[w];
#pragma unroll
      for (int b = 0; b < BPT; ++b) sstart[tid * BPT + b] += add;
    }
    __syncthreads();

#pragma unroll
    for (int i = 0; i < LS_ITEMS; ++i) {
      const uint32_t d = (uint32_t)(key[i] >> shift) & 255u;
      const uint32_t r = ((rank2[i >> 1] >> (16 * (i & 1))) & 0xFFFFu) + myoff[d] + sstart[d];
      skeys[r] = key[i];
      svals[r] = val[i];
    }
    __syncthreads();
    if (step + 1 < nsteps) {
#pragma unroll
      for (int i = 0; i < LS_ITEMS; ++i) {
        key[i] = skeys[wb + 32 * i];
        val[i] = svals[wb + 32 * i];
      }
      // the next scatter comes after two more barriers: no reader is overtaken
    }
  }
  if (!quick) break;

  // ---- runs of rows equal in the top two places: find the ones out of order ----
  const int s2 = places.shift[nplaces - qsteps];
  uint32_t* fixlist = woff;  // free between the scatter and the next pass's scan
  if (tid == 0) smisc[24] = 0;
  __syncthreads();
  bool bad = false;
  for (uint32_t i = tid + 1; i < cnt; i += LS_THREADS) {
    const uint64_t k = skeys[i], kp = skeys[i - 1];
    if ((kp >> s2) != (k >> s2) || k >= kp) continue;
    // (i-1, i) is an inversion inside a run: the FIRST inversion of a run owns the run
    uint32_t ra = i - 1;
    uint64_t prev = kp;
    bool first = true, ok = true;
    while (ra > 0) {
      const uint64_t q = skeys[ra - 1];
      if ((q >> s2) != (k >> s2)) break;
      if (prev < q) { first = false; break; }
      prev = q;
      --ra;
      if (i - ra > LS_RUN_LIMIT) { ok = false; break; }
    }
    if (!ok) { bad = true; continue; }
    if (!first) continue;
    uint32_t rb = i + 1;
    while (rb < cnt && (skeys[rb] >> s2) == (k >> s2)) {
      ++rb;
      if (rb - ra > LS_RUN_LIMIT) { ok = false; break; }
    }
    if (!ok) { bad = true; continue; }
    fixlist[atomicAdd(&smisc[24], 1u)] = ra | (rb << 16);
  }
  if (!__syncthreads_or(bad)) {
    const uint32_t nfix = smisc[24];
    for (uint32_t e = tid; e < nfix; e += LS_THREADS) {
      const uint32_t ra = fixlist[e] & 0xFFFFu, rb = fixlist[e] >> 16;
      for (uint32_t x = ra + 1; x < rb; ++x) {
        const uint64_t kx = skeys[x];
        const uint32_t vx = svals[x];
        uint32_t j = x;
        while (j > ra && skeys[j - 1] > kx) {
          skeys[j] = skeys[j - 1];
          svals[j] = svals[j - 1];
          --j;
        }
        skeys[j] = kx;
        svals[j] = vx;
      }
    }
    __syncthreads();
    break;
  }
  quick = false;
  step = 0;
  nsteps = nplaces;
#pragma unroll
  for (int i = 0; i < LS_ITEMS; ++i) {
    key[i] = skeys[wb + 32 * i];
    val[i] = svals[wb + 32 * i];
  }
  }
  for (uint32_t pos = tid; pos < cnt; pos += LS_THREADS) {
    dst_keys[st + pos] = skeys[pos];
    dst_vals[st + pos] = svals[pos];
  }
}

struct SortWorkspace {
  uint64_t* norm;    // N: normalised keys in input order
  uint64_t* keys_b;  // ping-pong side B
  uint32_t* vals_b;
  uint64_t* keys_a;  // side A keys when the caller does not want them
  uint32_t* ghist;   // NUM_PLACES * BINS
  uint32_t* gexcl;   // NUM_PLACES * BINS
  uint32_t* active;  // NUM_PLACES (+ padding)
  uint32_t* tile_counter;  // 16
  uint32_t* status;  // ntiles * BINS
  // MSD levels
  Seg* big[2];             // segments of the current / next level
  uint32_t* seg_hist;      // nbig * BINS
  uint32_t* seg_excl;      // nbig * BINS
  uint32_t* tile_seg;      // tile -> segment
  uint64_t* small[LS_CLASSES];  // local-sort lists by size class
  MsdCounters* ctr;
  uint64_t total;
};

// Bounds that size the MSD scratch: a level has at most n / (LS_CAP + 1) segments too
// large for a local sort, each with at most ceil(count / tile) <= count / tile + 1 tiles.
uint64_t msd_max_big(uint64_t n) { return n / (LS_CAP + 1) + 1; }
uint64_t msd_max_tiles(uint64_t n) { return n / RS_MIN_TILE + 1 + msd_max_big(n); }
uint64_t msd_max_small(uint64_t n, int cls) {
  // children of one level: 256 per parent; a class of sizes > lo also has < n / lo members
  const uint64_t by_parent = BINS * msd_max_big(n);
  const uint64_t by_size = cls == 0 ? n : n / ls_class_cap(cls - 1);
  return std::min<uint64_t>(by_parent, by_size) + 1;
}

SortWorkspace carve(void* base, uint64_t n) {
  SortWorkspace w;
  uint64_t off = 0;
  auto take = [&](uint64_t bytes) {
    uint64_t o = off;
    off = xchg_align_up(off + bytes, 256);
    return (uint8_t*)base + o;
  };
  const uint64_t ntiles = msd_max_tiles(n);
  w.norm = (uint64_t*)take(n * 8);
  w.keys_b = (uint64_t*)take(n * 8);
  w.keys_a = (uint64_t*)take(n * 8);
  w.vals_b = (uint32_t*)take(n * 4);
  w.ghist = (uint32_t*)take(NUM_PLACES * BINS * 4);
  w.gexcl = (uint32_t*)take(NUM_PLACES * BINS * 4);
  w.active = (uint32_t*)take(64);
  w.tile_counter = (uint32_t*)take(64);
  w.status = (uint32_t*)take(ntiles * BINS * 4);
  w.big[0] = (Seg*)take(msd_max_big(n) * sizeof(Seg));
  w.big[1] = (Seg*)take(msd_max_big(n) * sizeof(Seg));
  w.seg_hist = (uint32_t*)take(msd_max_big(n) * BINS * 4);
  w.seg_excl = (uint32_t*)take(msd_max_big(n) * BINS * 4);
  w.tile_seg = (uint32_t*)take(ntiles * 4);
  for (int c = 0; c < LS_CLASSES; ++c) w.small[c] = (uint64_t*)take(msd_max_small(n, c) * 8);
  w.ctr = (MsdCounters*)take(sizeof(MsdCounters));
  w.total = off;
  return w;
}

template <typename T>
void launch_normalize(const void* keys, const uint32_t* perm, const uint8_t* validity, int64_t vbit,
                      uint32_t n, int desc, uint64_t* out, uint32_t* ghist, cudaStream_t s) {
  int blocks = (int)std::min<uint64_t>((n + 255) / 256, (uint64_t)xchg_sm_count() * 8);
  XCHG_LAUNCH(XK_NORMALIZE, s, k_normalize_hist<T><<<blocks, 256, 0, s>>>((const T*)keys, perm, validity, vbit, n, desc, out, ghist));
}

int dispatch_normalize(const void* keys, int key_type, const uint32_t* perm, const uint8_t* validity,
                       int64_t vbit, uint32_t n, int desc, uint64_t* out, uint32_t* ghist,
                       cudaStream_t s) {
  switch (key_type) {
    case XCHG_I8: launch_normalize<int8_t>(keys, perm, validity, vbit, n, desc, out, ghist, s); break;
    case XCHG_I16: launch_normalize<int16_t>(keys, perm, validity, vbit, n, desc, out, ghist, s); break;
    case XCHG_I32: launch_normalize<int32_t>(keys, perm, validity, vbit, n, desc, out, ghist, s); break;
    case XCHG_I64: launch_normalize<int64_t>(keys, perm, validity, vbit, n, desc, out, ghist, s); break;
    case XCHG_U8: launch_normalize<uint8_t>(keys, perm, validity, vbit, n, desc, out, ghist, s); break;
    case XCHG_U16: launch_normalize<uint16_t>(keys, perm, validity, vbit, n, desc, out, ghist, s); break;
    case XCHG_U32: launch_normalize<uint32_t>(keys, perm, validity, vbit, n, desc, out, ghist, s); break;
    case XCHG_U64: launch_normalize<uint64_t>(keys, perm, validity, vbit, n, desc, out, ghist, s); break;
    case XCHG_F32: launch_normalize<float>(keys, perm, validity, vbit, n, desc, out, ghist, s); break;
    case XCHG_F64: launch_normalize<double>(keys, perm, validity, vbit, n, desc, out, ghist, s); break;
    default:
      xchg_set_error("unknown key_type %d", key_type);
      return XCHG_ERR_INVALID;
  }
  XCHG_LAUNCH_CHECK();
  return XCHG_OK;
}

int dispatch_key_scan(const void* keys, int key_type, uint32_t n, int desc, const KeyShifts& place_mask,
                      int want_orand, unsigned long long* orand, uint32_t* ghist, cudaStream_t s) {
  const int blocks = (int)std::min<uint64_t>((n + 255) / 256, (uint64_t)xchg_sm_count() * 8);
  const bool ok = xchg_dispatch_key_type(key_type, [&](auto tag) {
    using T = typename decltype(tag)::type;
    XCHG_LAUNCH(XK_NORMALIZE, s,
                k_key_scan<T><<<blocks, 256, 0, s>>>((const T*)keys, n, desc, place_mask, want_orand, orand, ghist));
  });
  XCHG_REQUIRE(ok, "unknown key_type %d", key_type);
  XCHG_LAUNCH_CHECK();
  return XCHG_OK;
}

int key_type_bytes(int key_type) {
  switch (key_type) {
    case XCHG_I8: case XCHG_U8: return 1;
    case XCHG_I16: case XCHG_U16: return 2;
    case XCHG_I32: case XCHG_U32: case XCHG_F32: return 4;
    default: return 8;
  }
}

}  // namespace

// CTA shape of the onesweep pass: XCHG_ONESWEEP_THREADS=256|512 and
// XCHG_ONESWEEP_ITEMS=8|12|16 override the defaults (read once; tuning knobs for
// profiling runs).
static void onesweep_shape(int* threads_out, int* items_out) {
  static int threads = 0, items = 0;
  if (threads == 0) {
    const char* e = getenv("XCHG_ONESWEEP_THREADS");
    threads = (e && atoi(e) == 512) ? 512 : ((e && atoi(e) == 256) ? 256 : RS_DEFAULT_THREADS);
    const char* f = getenv("XCHG_ONESWEEP_ITEMS");
    const int fi = f ? atoi(f) : 0;
    items = (fi == 8 || fi == 12 || fi == 16) ? fi : RS_DEFAULT_ITEMS;
    if (threads == 512) items = 16;
  }
  *threads_out = threads;
  *items_out = items;
}

// One onesweep pass.  segs == NULL: over the whole array of n rows (mode 1 = validity
// place).  segs != NULL: inside every segment of an MSD level, ntiles_seg tiles in all.
// raw != NULL: a RAW pass (keys read in their own type, see RawKeys); mode 2 = partition
// by destination (always RAW).  RAW passes use the default CTA shape.
static int onesweep_pass(int mode, const uint64_t* kin, const uint32_t* vin, uint64_t* kout,
                         uint32_t* vout, uint32_t n, int shift, const uint32_t* gexcl,
                         uint32_t* status, uint32_t* tile_counter, const uint8_t* validity,
                         int64_t vbit, const Seg* segs, const uint32_t* tile_seg,
                         uint32_t ntiles_seg, cudaStream_t s, const RawKeys* raw = nullptr) {
  int threads, items;
  onesweep_shape(&threads, &items);
  const RawKeys no_raw = {nullptr, 0, 0, nullptr, nullptr, 0};
#define XCHG_ONESWEEP_CASE(MODE, SEG, THREADS, ITEMS, RAW, EXTRA)                                       \
  {                                                                                                    \
    XCHG_ONCE_PER_DEVICE(XCHG_CUDA_TRY(cudaFuncSetAttribute(k_onesweep<MODE, THREADS, ITEMS, SEG, RAW>, \
                                         cudaFuncAttributeMaxDynamicSharedMemorySize,                  \
                                         (int)(rs_smem(THREADS, ITEMS) + (EXTRA)))));                  \
    const uint32_t ntiles = (SEG) ? ntiles_seg : (n + THREADS * ITEMS - 1) / (THREADS * ITEMS);        \
    XCHG_LAUNCH(XK_ONESWEEP, s,                                                                        \
                k_onesweep<MODE, THREADS, ITEMS, SEG, RAW>                                             \
                <<<ntiles, THREADS, rs_smem(THREADS, ITEMS) + (EXTRA), s>>>(                           \
                    kin, vin, kout, vout, n, shift, gexcl, status, tile_counter, validity, vbit, segs,  \
                    tile_seg, raw ? *raw : no_raw));                                                   \
  }
#define XCHG_ONESWEEP_MODE(MODE, SEG)                                          \
  if (threads == 512) XCHG_ONESWEEP_CASE(MODE, SEG, 512, 16, false, 0)         \
  else if (items == 16) XCHG_ONESWEEP_CASE(MODE, SEG, 256, 16, false, 0)       \
  else if (items == 12) XCHG_ONESWEEP_CASE(MODE, SEG, 256, 12, false, 0)       \
  else XCHG_ONESWEEP_CASE(MODE, SEG, 256, 8, false, 0)
  if (raw && mode == 2) { XCHG_ONESWEEP_CASE(2, false, RS_DEFAULT_THREADS, RS_DEFAULT_ITEMS, true, RS_RULE_SMEM) }
  else if (raw) { XCHG_ONESWEEP_CASE(0, false, RS_DEFAULT_THREADS, RS_DEFAULT_ITEMS, true, 0) }
  else if (segs) { XCHG_ONESWEEP_MODE(0, true) }
  else if (mode == 0) { XCHG_ONESWEEP_MODE(0, false) }
  else { XCHG_ONESWEEP_MODE(1, false) }
#undef XCHG_ONESWEEP_MODE
#undef XCHG_ONESWEEP_CASE
  XCHG_LAUNCH_CHECK();
  return XCHG_OK;
}

static int seg_hist_pass(const uint64_t* keys, int shift, const Seg* segs, const uint32_t* tile_seg,
                         uint32_t ntiles, uint32_t* seg_hist, cudaStream_t s) {
  int threads, items;
  onesweep_shape(&threads, &items);
  const int tile = threads * items;
#define XCHG_SEG_HIST(TILE)                                                                          \
  XCHG_LAUNCH(XK_MSD_AUX, s, k_seg_hist<TILE><<<ntiles, 256, 0, s>>>(keys, shift, segs, tile_seg, seg_hist))
  if (tile == 8192) XCHG_SEG_HIST(8192);
  else if (tile == 4096) XCHG_SEG_HIST(4096);
  else if (tile == 3072) XCHG_SEG_HIST(3072);
  else XCHG_SEG_HIST(2048);
#undef XCHG_SEG_HIST
  XCHG_LAUNCH_CHECK();
  return XCHG_OK;
}

// Local sorts of one level's small segments, one launch per non-empty size class.
static int local_sort_level(const uint64_t* src_k, const uint32_t* src_v, uint64_t* dst_k, uint32_t* dst_v,
                            uint64_t* const* small, const uint32_t* nsmall, const LocalPlaces& lp,
                            cudaStream_t s) {
  // 8 rows per thread (256 / 512 threads) for the 2048- and 4096-row classes: same speed as
  // 16 rows per thread on fewer threads, and the shape that can use match_digit_asm;
  // XCHG_LS_WIDE2=0 / XCHG_LS_WIDE3=0 select 128 x 16 / 256 x 16
  static int wide2 = -1;
  static int wide3 = -1;
  if (wide2 < 0) {
    const char* e = getenv("XCHG_LS_WIDE2");
    wide2 = (e && atoi(e) == 0) ? 0 : 1;
    const char* f = getenv("XCHG_LS_WIDE3");
    wide3 = (f && atoi(f) == 0) ? 0 : 1;
  }
  XCHG_ONCE_PER_DEVICE({
    XCHG_CUDA_TRY(cudaFuncSetAttribute(k_local_sort<128, 16>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)ls_smem(128, 16)));
    XCHG_CUDA_TRY(cudaFuncSetAttribute(k_local_sort<256, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)ls_smem(256, 8)));
    XCHG_CUDA_TRY(cudaFuncSetAttribute(k_local_sort<256, 16>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)ls_smem(256, 16)));
    XCHG_CUDA_TRY(cudaFuncSetAttribute(k_local_sort<512, 16>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)ls_smem(512, 16)));
    XCHG_CUDA_TRY(cudaFuncSetAttribute(k_local_sort<512, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)ls_smem(512, 8)));
  });
#define XCHG_LOCAL_SORT(CLS, THREADS, ITEMS)                                                         \
  if (nsmall[CLS])                                                                                   \
    XCHG_LAUNCH(XK_LOCAL_SORT, s,                                                                    \
                k_local_sort<THREADS, ITEMS><<<nsmall[CLS], THREADS, ls_smem(THREADS, ITEMS), s>>>(   \
                    src_k, src_v, dst_k, dst_v, small[CLS], lp))
  XCHG_LOCAL_SORT(0, 128, 4);
  XCHG_LOCAL_SORT(1, 128, 8);
  if (wide2) { XCHG_LOCAL_SORT(2, 256, 8); } else { XCHG_LOCAL_SORT(2, 128, 16); }
  if (wide3) { XCHG_LOCAL_SORT(3, 512, 8); } else { XCHG_LOCAL_SORT(3, 256, 16); }
  XCHG_LOCAL_SORT(4, 512, 16);
#undef XCHG_LOCAL_SORT
  XCHG_LAUNCH_CHECK();
  return XCHG_OK;
}

// 0 = hybrid (MSD levels + local sort), 1 = every pass global (LSD).  XCHG_SORT_ALGO=lsd
// selects 1 at load; xchg_set_sort_algorithm overrides.
static std::atomic<int> g_sort_algo{-1};
static int sort_algorithm() {
  int a = g_sort_algo.load(std::memory_order_relaxed);
  if (a < 0) {
    const char* e = getenv("XCHG_SORT_ALGO");
    a = (e && e[0] == 'l') ? 1 : 0;
    g_sort_algo.store(a);
  }
  return a;
}

extern "C" int xchg_set_sort_algorithm(int algorithm) {
  XCHG_REQUIRE(algorithm == 0 || algorithm == 1, "algorithm must be 0 (hybrid) or 1 (LSD)");
  g_sort_algo.store(algorithm);
  return XCHG_OK;
}

extern "C" int xchg_sort_workspace_bytes(uint64_t n, uint64_t* bytes_out) {
  XCHG_REQUIRE(bytes_out != nullptr, "bytes_out is NULL");
  if (n > XCHG_MAX_ROWS) {
    xchg_set_error("n=%llu exceeds XCHG_MAX_ROWS", (unsigned long long)n);
    return XCHG_ERR_LIMIT;
  }
  *bytes_out = carve(nullptr, n).total + 256;
  return XCHG_OK;
}

extern "C" int xchg_normalize_keys(const void* keys, int key_type, int descending,
                                   const uint8_t* validity, int64_t validity_bit_offset,
                                   const uint32_t* perm, uint64_t n, uint64_t* norm_out,
                                   void* stream) {
  if (n == 0) return XCHG_OK;
  XCHG_REQUIRE(keys && norm_out, "NULL keys / norm_out");
  if (n > XCHG_MAX_ROWS) {
    xchg_set_error("n=%llu exceeds XCHG_MAX_ROWS", (unsigned long long)n);
    return XCHG_ERR_LIMIT;
  }
  return dispatch_normalize(keys, key_type, perm, validity, validity_bit_offset, (uint32_t)n,
                            descending, norm_out, nullptr, (cudaStream_t)stream);
}

extern "C" int xchg_sort_column(const void* keys, int key_type, int descending,
                                const uint8_t* validity, int64_t validity_bit_offset,
                                const uint32_t* perm_in, uint64_t n, uint32_t* perm_out,
                                uint64_t* norm_keys_out, void* workspace, uint64_t workspace_bytes,
                                void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  if (n == 0) return XCHG_OK;
  XCHG_REQUIRE(keys && perm_out && workspace, "NULL keys / perm_out / workspace");
  if (n > XCHG_MAX_ROWS) {
    xchg_set_error("n=%llu exceeds XCHG_MAX_ROWS", (unsigned long long)n);
    return XCHG_ERR_LIMIT;
  }
  uint8_t* base = (uint8_t*)xchg_align_up((uint64_t)(uintptr_t)workspace, 256);
  SortWorkspace w = carve(base, n);
  if ((uint64_t)(base - (uint8_t*)workspace) + w.total > workspace_bytes) {
    xchg_set_error("sort workspace too small: have %llu, need %llu",
                   (unsigned long long)workspace_bytes, (unsigned long long)(w.total + 256));
    return XCHG_ERR_WORKSPACE;
  }
  const uint32_t n32 = (uint32_t)n;
  const uint32_t ntiles = (n32 + RS_MIN_TILE - 1) / RS_MIN_TILE;  // status rows cleared per pass

  XCHG_CUDA_TRY(cudaMemsetAsync(w.ghist, 0, NUM_PLACES * BINS * 4, s));
  int rc;
  // digit places in ascending significance: shift, histogram slot (ghist / gexcl row)
  struct Place {
    int shift, slot;
    bool validity;
  };
  Place places[NUM_PLACES];
  int np = 0;
  // A column without nulls that is sorted in its stored order needs no materialised radix
  // keys: one scan tells which bits matter (and builds the first pass's histogram), the
  // first pass reads the raw values (RAW).  XCHG_SORT_RAW=0 keeps the normalise pass.
  static const bool raw_enabled = [] {
    const char* e = getenv("XCHG_SORT_RAW");
    return !(e && atoi(e) == 0);
  }();
  const bool raw_able = raw_enabled && perm_in == nullptr && validity == nullptr;
  const RawKeys raw0 = {keys, key_type, descending, nullptr, nullptr, 0};
  const RawKeys* raw = raw_able ? &raw0 : nullptr;
  if (raw_able) {
    XCHG_REQUIRE(key_type >= XCHG_I8 && key_type <= XCHG_F64, "unknown key_type %d", key_type);
    unsigned long long* orand = (unsigned long long*)w.active;  // 64 bytes: {OR, AND}
    XCHG_CUDA_TRY(cudaMemsetAsync(orand, 0, 8, s));
    XCHG_CUDA_TRY(cudaMemsetAsync(orand + 1, 0xFF, 8, s));
    // the top place's shift is only known after the scan: guess it (what the last sort of
    // this key type used; the type's top byte the first time) and scan again on a miss
    static std::atomic<int> last_top_shift[XCHG_F64 + 1];
    static std::atomic<bool> last_init{false};
    if (!last_init.exchange(true))
      for (int t = 0; t <= XCHG_F64; ++t) last_top_shift[t].store(8 * (key_type_bytes(t) - 1));
    KeyShifts ks;
    ks.n = 1;
    ks.shift[0] = last_top_shift[key_type].load();
    rc = dispatch_key_scan(keys, key_type, n32, descending, ks, 1, orand, w.ghist, s);
    if (rc) return rc;
    unsigned long long oa[2];
    XCHG_CUDA_TRY(cudaMemcpyAsync(oa, orand, sizeof(oa), cudaMemcpyDeviceToHost, s));
    XCHG_CUDA_TRY(cudaStreamSynchronize(s));
    const uint64_t diff = oa[0] ^ oa[1];
    if (diff) {
      const int top = 63 - __builtin_clzll(diff), low = __builtin_ctzll(diff);
      // places from the top bit down, 8 bits each, the lowest one clamped to `low` (it may
      // then share bits with its neighbour: harmless for a stable pass in either direction)
      int shifts[KEY_PLACES];
      int cnt = 0;
      for (int sh = std::max(low, top - 7);; sh = std::max(low, sh - 8)) {
        shifts[cnt++] = sh;
        if (sh == low) break;
      }
      np = cnt;
      for (int j = 0; j < np; ++j) places[j] = Place{shifts[np - 1 - j], 0, false};
      last_top_shift[key_type].store(places[np - 1].shift);
      const bool all_places = sort_algorithm() == 1 || np <= 2;
      const bool hit = ks.shift[0] == places[np - 1].shift;
      if (all_places ? !(hit && np == 1) : !hit) {
        KeyShifts need;
        need.n = 0;
        if (all_places) {
          for (int j = 0; j < np; ++j) {
            places[j].slot = need.n;
            need.shift[need.n++] = places[j].shift;
          }
        } else {
          need.shift[need.n++] = places[np - 1].shift;
        }
        XCHG_CUDA_TRY(cudaMemsetAsync(w.ghist, 0, NUM_PLACES * BINS * 4, s));
        rc = dispatch_key_scan(keys, key_type, n32, descending, need, 0, orand, w.ghist, s);
        if (rc) return rc;
      }
      XCHG_LAUNCH(XK_SCAN_HIST, s, k_scan_hist<<<1, BINS, 0, s>>>(w.ghist, w.gexcl, w.active, n32));
      XCHG_LAUNCH_CHECK();
    }
  } else {
    rc = dispatch_normalize(keys, key_type, perm_in, validity, validity_bit_offset, n32, descending,
                            w.norm, w.ghist, s);
    if (rc) return rc;
    XCHG_LAUNCH(XK_SCAN_HIST, s, k_scan_hist<<<1, BINS, 0, s>>>(w.ghist, w.gexcl, w.active, n32));
    XCHG_LAUNCH_CHECK();
    uint32_t active[NUM_PLACES];
    XCHG_CUDA_TRY(cudaMemcpyAsync(active, w.active, sizeof(active), cudaMemcpyDeviceToHost, s));
    XCHG_CUDA_TRY(cudaStreamSynchronize(s));
    for (int p = 0; p < KEY_PLACES; ++p)
      if (active[p]) places[np++] = Place{8 * p, p, false};
    const bool wide_int = key_type == XCHG_I64 || key_type == XCHG_U64;
    if (wide_int && validity && active[KEY_PLACES]) places[np++] = Place{0, KEY_PLACES, true};
  }

  uint64_t* keys_a = norm_keys_out ? norm_keys_out : w.keys_a;
  if (np == 0) {
    // every key equal: the stable order is the input order
    int blocks = (int)std::min<uint64_t>((n + 255) / 256, (uint64_t)xchg_sm_count() * 8);
    XCHG_LAUNCH(XK_MISC, s, k_iota_or_copy<<<blocks, 256, 0, s>>>(perm_in, perm_out, n32));
    XCHG_LAUNCH_CHECK();
    if (norm_keys_out) {
      if (raw_able) {
        rc = dispatch_normalize(keys, key_type, nullptr, nullptr, 0, n32, descending, norm_keys_out, nullptr, s);
        if (rc) return rc;
      } else {
        XCHG_CUDA_TRY(cudaMemcpyAsync(norm_keys_out, w.norm, n * 8, cudaMemcpyDeviceToDevice, s));
      }
    }
    return XCHG_OK;
  }
  // XCHG_SORT_ALGO=lsd keeps every pass global (A/B runs); the default finishes small
  // segments in shared memory.  With one or two places the LSD passes are already minimal.
  if (sort_algorithm() == 1 || np <= 2) {
    const uint64_t* kin = w.norm;
    const uint32_t* vin = perm_in;
    for (int j = 0; j < np; ++j) {
      const bool to_a = ((np - 1 - j) % 2) == 0;  // the last pass lands on side A
      uint64_t* kout = to_a ? keys_a : w.keys_b;
      uint32_t* vout = to_a ? perm_out : w.vals_b;
      const Place p = places[j];
      XCHG_CUDA_TRY(cudaMemsetAsync(w.tile_counter, 0, 64, s));
      XCHG_CUDA_TRY(cudaMemsetAsync(w.status, 0, (size_t)ntiles * BINS * 4, s));
      rc = onesweep_pass(p.validity ? 1 : 0, kin, vin, kout, vout, n32, p.shift, w.gexcl + p.slot * BINS,
                         w.status, w.tile_counter, validity, validity_bit_offset, nullptr, nullptr, 0, s,
                         j == 0 ? raw : nullptr);
      if (rc) return rc;
      kin = kout;
      vin = vout;
    }
    return XCHG_OK;
  }

  // ---- MSD levels from the top place down; level l writes side X(l), X(np-1) = A ----
  int os_threads, os_items;
  onesweep_shape(&os_threads, &os_items);
  const uint32_t tile_rows = (uint32_t)(os_threads * os_items);
  auto side_keys = [&](int l) { return ((np - 1 - l) % 2) == 0 ? keys_a : w.keys_b; };
  auto side_vals = [&](int l) { return ((np - 1 - l) % 2) == 0 ? perm_out : w.vals_b; };
  MsdCounters h;
  uint32_t nbig = 1, ntiles_l = 0;
  for (int l = 0; l < np; ++l) {
    const Place p = places[np - 1 - l];
    const bool last = l == np - 1;
    Seg* cur = w.big[l & 1];
    Seg* nxt = w.big[(l + 1) & 1];
    const uint32_t* hist;
    const uint32_t* excl;
    XCHG_CUDA_TRY(cudaMemsetAsync(w.tile_counter, 0, 64, s));
    if (l == 0) {
      XCHG_CUDA_TRY(cudaMemsetAsync(w.status, 0, (size_t)((n32 + tile_rows - 1) / tile_rows) * BINS * 4, s));
      rc = onesweep_pass(p.validity ? 1 : 0, w.norm, perm_in, side_keys(0), side_vals(0), n32, p.shift,
                         w.gexcl + p.slot * BINS, w.status, w.tile_counter, validity, validity_bit_offset,
                         nullptr, nullptr, 0, s, raw);
      if (rc) return rc;
      hist = w.ghist + p.slot * BINS;
      excl = w.gexcl + p.slot * BINS;
    } else {
      XCHG_LAUNCH(XK_MSD_AUX, s, k_fill_tile_map<<<nbig, 128, 0, s>>>(cur, tile_rows, w.tile_seg));
      XCHG_CUDA_TRY(cudaMemsetAsync(w.seg_hist, 0, (size_t)nbig * BINS * 4, s));
      XCHG_CUDA_TRY(cudaMemsetAsync(w.status, 0, (size_t)ntiles_l * BINS * 4, s));
      rc = seg_hist_pass(side_keys(l - 1), p.shift, cur, w.tile_seg, ntiles_l, w.seg_hist, s);
      if (rc) return rc;
      XCHG_LAUNCH(XK_MSD_AUX, s, k_seg_scan<<<nbig, BINS, 0, s>>>(w.seg_hist, w.seg_excl));
      XCHG_LAUNCH_CHECK();
      rc = onesweep_pass(0, side_keys(l - 1), side_vals(l - 1), side_keys(l), side_vals(l), 0, p.shift,
                         w.seg_excl, w.status, w.tile_counter, nullptr, 0, cur, w.tile_seg, ntiles_l, s);
      if (rc) return rc;
      hist = w.seg_hist;
      excl = w.seg_excl;
    }
    if (last) break;  // every row now sits at its final position on side A
    XCHG_CUDA_TRY(cudaMemsetAsync(w.ctr, 0, sizeof(MsdCounters), s));
    const int can_group = !p.validity;
    SmallLists lists;
    for (int c = 0; c < LS_CLASSES; ++c) lists.p[c] = w.small[c];
    XCHG_LAUNCH(XK_MSD_AUX, s,
                k_msd_classify<<<nbig, 32, 0, s>>>(l == 0 ? nullptr : cur, nbig, hist, excl, tile_rows, can_group,
                                                   nxt, lists, w.ctr));
    XCHG_LAUNCH_CHECK();
    XCHG_CUDA_TRY(cudaMemcpyAsync(&h, w.ctr, sizeof(h), cudaMemcpyDeviceToHost, s));
    XCHG_CUDA_TRY(cudaStreamSynchronize(s));
    // small children of this level: finish them now, by the places below this one
    LocalPlaces lp;
    lp.n = np - 1 - l;
    for (int j = 0; j < lp.n; ++j) lp.shift[j] = places[j].shift;
    lp.shift[lp.n] = can_group ? p.shift : 0;
    rc = local_sort_level(side_keys(l), side_vals(l), keys_a, perm_out, w.small, h.nsmall, lp, s);
    if (rc) return rc;
    nbig = h.nbig;
    ntiles_l = h.ntiles;
    if (nbig == 0) break;
  }
  return XCHG_OK;
}

// ---- stable partition of a block's rows by destination (the exchange's map side) ----

int xchg_internal_fill_rule(DestRule* rule, const uint32_t* pick_host, const uint8_t* strict_host, int nthr);

namespace {
struct PartitionWorkspace {
  uint32_t* gexcl;         // BINS
  uint32_t* tile_counter;  // 16
  uint64_t* thr;           // BINS
  uint8_t* strict;         // BINS
  uint32_t* status;        // ntiles * BINS
  uint64_t total;
};
PartitionWorkspace carve_partition(void* base, uint64_t n) {
  PartitionWorkspace w;
  uint64_t off = 0;
  auto take = [&](uint64_t bytes) {
    uint64_t o = off;
    off = xchg_align_up(off + bytes, 256);
    return (uint8_t*)base + o;
  };
  const uint64_t ntiles = n / (RS_DEFAULT_THREADS * RS_DEFAULT_ITEMS) + 1;
  w.gexcl = (uint32_t*)take(BINS * 4);
  w.tile_counter = (uint32_t*)take(64);
  w.thr = (uint64_t*)take(BINS * 8);
  w.strict = (uint8_t*)take(BINS);
  w.status = (uint32_t*)take(ntiles * BINS * 4);
  w.total = off;
  return w;
}

// thresholds picked from the boundary list, and the exclusive scan of the destination counts
__global__ void __launch_bounds__(BINS)
k_partition_prepare(const uint64_t* __restrict__ bounds, DestRule rule, const unsigned long long* __restrict__ counts,
                    uint64_t* __restrict__ thr, uint8_t* __restrict__ strict, uint32_t* __restrict__ gexcl) {
  __shared__ uint32_t wsum[BINS / 32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid < rule.nthr) {
    thr[tid] = bounds[rule.pick[tid]];
    strict[tid] = rule.strict[tid];
  }
  const uint32_t c = tid <= rule.nthr ? (uint32_t)counts[tid] : 0u;
  uint32_t x = c;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
    if (lane >= o) x += y;
  }
  if (lane == 31) wsum[warp] = x;
  __syncthreads();
  uint32_t base = 0;
  for (int w = 0; w < warp; ++w) base += wsum[w];
  gexcl[tid] = base + x - c;
}
}  // namespace

extern "C" int xchg_partition_workspace_bytes(uint64_t n, uint64_t* bytes_out) {
  XCHG_REQUIRE(bytes_out != nullptr, "bytes_out is NULL");
  if (n > XCHG_MAX_ROWS) {
    xchg_set_error("n=%llu exceeds XCHG_MAX_ROWS", (unsigned long long)n);
    return XCHG_ERR_LIMIT;
  }
  *bytes_out = carve_partition(nullptr, n).total + 256;
  return XCHG_OK;
}

extern "C" int xchg_partition_rows(const void* keys, int key_type, int descending, uint64_t n,
                                   const uint64_t* bounds, const uint32_t* pick_host,
                                   const uint8_t* strict_host, int nthr, const uint64_t* counts,
                                   uint64_t* keys_out, uint32_t* rows_out, void* workspace,
                                   uint64_t workspace_bytes, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  if (n == 0) return XCHG_OK;
  XCHG_REQUIRE(keys && counts && keys_out && rows_out && workspace, "NULL argument");
  XCHG_REQUIRE(key_type >= XCHG_I8 && key_type <= XCHG_F64, "unknown key_type %d", key_type);
  XCHG_REQUIRE(nthr == 0 || bounds, "NULL bounds");
  if (n > XCHG_MAX_ROWS) {
    xchg_set_error("n=%llu exceeds XCHG_MAX_ROWS", (unsigned long long)n);
    return XCHG_ERR_LIMIT;
  }
  DestRule rule;
  int rc = xchg_internal_fill_rule(&rule, pick_host, strict_host, nthr);
  if (rc) return rc;
  uint8_t* base = (uint8_t*)xchg_align_up((uint64_t)(uintptr_t)workspace, 256);
  PartitionWorkspace w = carve_partition(base, n);
  if ((uint64_t)(base - (uint8_t*)workspace) + w.total > workspace_bytes) {
    xchg_set_error("partition workspace too small: have %llu, need %llu",
                   (unsigned long long)workspace_bytes, (unsigned long long)(w.total + 256));
    return XCHG_ERR_WORKSPACE;
  }
  const uint32_t n32 = (uint32_t)n;
  const uint32_t ntiles = (n32 + RS_DEFAULT_THREADS * RS_DEFAULT_ITEMS - 1) / (RS_DEFAULT_THREADS * RS_DEFAULT_ITEMS);
  XCHG_CUDA_TRY(cudaMemsetAsync(w.tile_counter, 0, 64, s));
  XCHG_CUDA_TRY(cudaMemsetAsync(w.status, 0, (size_t)ntiles * BINS * 4, s));
  XCHG_LAUNCH(XK_PARTITION, s,
              k_partition_prepare<<<1, BINS, 0, s>>>(bounds, rule, (const unsigned long long*)counts, w.thr,
                                                    w.strict, w.gexcl));
  XCHG_LAUNCH_CHECK();
  const RawKeys raw = {keys, key_type, descending, w.thr, w.strict, nthr};
  return onesweep_pass(2, nullptr, nullptr, keys_out, rows_out, n32, 0, w.gexcl, w.status, w.tile_counter,
                       nullptr, 0, nullptr, nullptr, 0, s, &raw);
}

extern "C" int xchg_denormalize_keys(const uint64_t* norm, int key_type, int descending, uint64_t n,
                                     void* out, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  if (n == 0) return XCHG_OK;
  XCHG_REQUIRE(norm && out, "NULL norm / out");
  if (n > XCHG_MAX_ROWS) {
    xchg_set_error("n=%llu exceeds XCHG_MAX_ROWS", (unsigned long long)n);
    return XCHG_ERR_LIMIT;
  }
  const uint32_t n32 = (uint32_t)n;
  const int blocks = (int)std::min<uint64_t>((n + 255) / 256, (uint64_t)xchg_sm_count() * 8);
#define XCHG_DENORM(T) \
  XCHG_LAUNCH(XK_NORMALIZE, s, k_denormalize<T><<<blocks, 256, 0, s>>>(norm, n32, descending, (T*)out))
  switch (key_type) {
    case XCHG_I8: XCHG_DENORM(int8_t); break;
    case XCHG_I16: XCHG_DENORM(int16_t); break;
    case XCHG_I32: XCHG_DENORM(int32_t); break;
    case XCHG_I64: XCHG_DENORM(int64_t); break;
    case XCHG_U8: XCHG_DENORM(uint8_t); break;
    case XCHG_U16: XCHG_DENORM(uint16_t); break;
    case XCHG_U32: XCHG_DENORM(uint32_t); break;
    case XCHG_U64: XCHG_DENORM(uint64_t); break;
    default:
      xchg_set_error("xchg_denormalize_keys: integer key types only (float keys keep NaN payloads and "
                     "signed zeros only through a gather), got %d", key_type);
      return XCHG_ERR_INVALID;
  }
#undef XCHG_DENORM
  XCHG_LAUNCH_CHECK();
  return XCHG_OK;
}
