// K5: payload gather ("take") kernels for sm_100a.
//
// Replaces pyarrow Table.take as the reference calls it from
// python/ray/data/_internal/arrow_ops/transform_pyarrow.py:45-71 (take_table) for
// primitive, validity-bitmap and variable-width (binary/string/list) buffers.
//
// The row gather (k_gather_rows) is the payload hot loop: rows are random reads,
// so each CTA stages a tile of rows in shared memory with 16-byte cp.async copies
// that keep the SOURCE alignment (no registers held while loads are in flight =
// deep memory-level parallelism), then each warp re-aligns a row to its
// DESTINATION with funnel shifts and writes 4-byte words; consecutive rows are
// consecutive in the output, so the writes of a tile form one dense stream.
// Algorithmic bytes: 2 x row bytes (+4 B index, +16 B offsets for var-width).
#include <algorithm>
#include <mutex>

#include "common.cuh"

namespace {

// ------------------------------------------------------------ fixed width ---

template <typename T>
__global__ void __launch_bounds__(256)
k_gather_fixed(const T* __restrict__ in, const uint32_t* __restrict__ idx, uint64_t n,
               T* __restrict__ out) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    out[i] = ld_random(in + idx[i]);
}

__global__ void __launch_bounds__(256)
k_gather_bitmap(const uint8_t* __restrict__ in, int64_t in_bit, const uint32_t* __restrict__ idx,
                uint64_t n, uint32_t* __restrict__ out_words, unsigned long long* null_count) {
  // one warp builds one 32-bit output word per iteration
  const uint64_t nwords = (n + 31) / 32;
  const int lane = threadIdx.x & 31;
  const uint64_t warp0 = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const uint64_t nwarps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
  unsigned long long nulls = 0;
  for (uint64_t w = warp0; w < nwords; w += nwarps) {
    const uint64_t i = w * 32 + lane;
    bool bit = false;
    if (i < n) bit = bit_is_set(in, (int64_t)idx[i] + in_bit);
    const unsigned word = __ballot_sync(0xffffffffu, bit);
    if (lane == 0) {
      out_words[w] = word;
      const unsigned live = (w * 32 + 32 <= n) ? 0xffffffffu : ((1u << (n - w * 32)) - 1u);
      nulls += __popc(~word & live);
    }
  }
  if (null_count && lane == 0 && nulls) atomicAdd(null_count, nulls);
}

// ---------------------------------------------------------------- offsets ---

__global__ void k_offsets_widen(const int32_t* __restrict__ in, uint64_t count, int64_t base,
                                int64_t* __restrict__ out) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  const int64_t first = in[0];
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride)
    out[i] = (int64_t)in[i] - first + base;
}

__global__ void k_offsets_narrow(const int64_t* __restrict__ in, uint64_t count,
                                 int32_t* __restrict__ out) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  const int64_t first = in[0];
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride)
    out[i] = (int32_t)(in[i] - first);
}

__global__ void k_offsets_uniform32(uint64_t count, uint32_t width, int32_t* __restrict__ out) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride)
    out[i] = (int32_t)(i * width);
}

__global__ void k_offsets_uniform64(uint64_t count, uint64_t width, int64_t base,
                                    int64_t* __restrict__ out) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride)
    out[i] = base + (int64_t)(i * width);
}

__global__ void k_offsets_check_uniform(const int64_t* __restrict__ off, uint64_t n_rows,
                                        uint64_t width, uint32_t* flag) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  bool bad = false;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_rows; i += stride)
    bad |= (uint64_t)(off[i + 1] - off[i]) != width;
  if (__any_sync(0xffffffffu, bad) && (threadIdx.x & 31) == 0) atomicOr(flag, 1u);
}

// gathered lengths -> exclusive offsets: reduce / scan-of-tiles / scan (3 launches)
constexpr int SC_THREADS = 256;
constexpr int SC_ITEMS = 8;
constexpr int SC_TILE = SC_THREADS * SC_ITEMS;  // 2048 rows

__device__ __forceinline__ int64_t block_exclusive_scan(int64_t x, int64_t* total, int64_t* swarp) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int64_t incl = x;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    int64_t y = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += y;
  }
  if (lane == 31) swarp[warp] = incl;
  __syncthreads();
  int64_t base = 0, tot = 0;
  for (int w = 0; w < SC_THREADS / 32; ++w) {
    const int64_t s = swarp[w];
    if (w < warp) base += s;
    tot += s;
  }
  __syncthreads();
  *total = tot;
  return base + incl - x;
}

__global__ void __launch_bounds__(SC_THREADS)
k_lengths_reduce(const int64_t* __restrict__ in_off, const uint32_t* __restrict__ idx, uint64_t n,
                 int64_t* __restrict__ tile_sums) {
  __shared__ int64_t swarp[SC_THREADS / 32];
  const uint64_t base = (uint64_t)blockIdx.x * SC_TILE;
  int64_t sum = 0;
#pragma unroll
  for (int k = 0; k < SC_ITEMS; ++k) {
    const uint64_t i = base + k * SC_THREADS + threadIdx.x;
    if (i < n) {
      const uint32_t r = idx[i];
      sum += in_off[r + 1] - in_off[r];
    }
  }
  int64_t tot;
  block_exclusive_scan(sum, &tot, swarp);
  if (threadIdx.x == 0) tile_sums[blockIdx.x] = tot;
}

__global__ void __launch_bounds__(SC_THREADS)
k_scan_tiles(int64_t* tile_sums, uint64_t ntiles) {
  __shared__ int64_t swarp[SC_THREADS / 32];
  int64_t carry = 0;
  for (uint64_t b = 0; b < ntiles; b += SC_THREADS) {
    const uint64_t i = b + threadIdx.x;
    const int64_t x = i < ntiles ? tile_sums[i] : 0;
    int64_t tot;
    const int64_t ex = block_exclusive_scan(x, &tot, swarp);
    if (i < ntiles) tile_sums[i] = carry + ex;
    carry += tot;
  }
}

__global__ void __launch_bounds__(SC_THREADS)
k_lengths_scan(const int64_t* __restrict__ in_off, const uint32_t* __restrict__ idx, uint64_t n,
               const int64_t* __restrict__ tile_excl, int64_t* __restrict__ out_off) {
  __shared__ int64_t swarp[SC_THREADS / 32];
  // blocked arrangement: thread t owns rows base + t*SC_ITEMS .. +SC_ITEMS-1
  const uint64_t base = (uint64_t)blockIdx.x * SC_TILE + (uint64_t)threadIdx.x * SC_ITEMS;
  int64_t len[SC_ITEMS];
  int64_t sum = 0;
#pragma unroll
  for (int k = 0; k < SC_ITEMS; ++k) {
    const uint64_t i = base + k;
    len[k] = 0;
    if (i < n) {
      const uint32_t r = idx[i];
      len[k] = in_off[r + 1] - in_off[r];
    }
    sum += len[k];
  }
  int64_t tot;
  int64_t run = tile_excl[blockIdx.x] + block_exclusive_scan(sum, &tot, swarp);
  if (blockIdx.x == 0 && threadIdx.x == 0) out_off[0] = 0;
#pragma unroll
  for (int k = 0; k < SC_ITEMS; ++k) {
    const uint64_t i = base + k;
    run += len[k];
    if (i < n) out_off[i + 1] = run;
  }
}

// The same two passes for callers that also want the rows' SOURCE spans kept: the first pass
// stores in_off[idx[i]] and the row's length (into out_off[i + 1], which the second pass scans
// in place), so in_off -- a random 16-byte read per row that costs a 64-128-byte DRAM fill -- is
// touched once per take instead of three times (lengths, scan, data gather).
__global__ void __launch_bounds__(SC_THREADS)
k_spans_reduce(const int64_t* __restrict__ in_off, const uint32_t* __restrict__ idx, uint64_t n,
               int64_t* __restrict__ src_start, int64_t* __restrict__ out_off, int64_t* __restrict__ tile_sums) {
  __shared__ int64_t swarp[SC_THREADS / 32];
  const uint64_t base = (uint64_t)blockIdx.x * SC_TILE;
  int64_t sum = 0;
#pragma unroll
  for (int k = 0; k < SC_ITEMS; ++k) {
    const uint64_t i = base + k * SC_THREADS + threadIdx.x;
    if (i < n) {
      const uint32_t r = idx[i];
      const int64_t b = in_off[r], len = in_off[r + 1] - b;
      src_start[i] = b;
      out_off[i + 1] = len;
      sum += len;
    }
  }
  int64_t tot;
  block_exclusive_scan(sum, &tot, swarp);
  if (threadIdx.x == 0) tile_sums[blockIdx.x] = tot;
}

// out_off[i + 1] holds row i's length on entry and the end of row i on exit (every element is
// read and written by the same thread)
__global__ void __launch_bounds__(SC_THREADS)
k_spans_scan(uint64_t n, const int64_t* __restrict__ tile_excl, int64_t* out_off) {
  __shared__ int64_t swarp[SC_THREADS / 32];
  const uint64_t base = (uint64_t)blockIdx.x * SC_TILE + (uint64_t)threadIdx.x * SC_ITEMS;
  int64_t len[SC_ITEMS];
  int64_t sum = 0;
#pragma unroll
  for (int k = 0; k < SC_ITEMS; ++k) {
    const uint64_t i = base + k;
    len[k] = i < n ? out_off[i + 1] : 0;
    sum += len[k];
  }
  int64_t tot;
  int64_t run = tile_excl[blockIdx.x] + block_exclusive_scan(sum, &tot, swarp);
  if (blockIdx.x == 0 && threadIdx.x == 0) out_off[0] = 0;
#pragma unroll
  for (int k = 0; k < SC_ITEMS; ++k) {
    const uint64_t i = base + k;
    run += len[k];
    if (i < n) out_off[i + 1] = run;
  }
}

// ------------------------------------------------------------- row gather ---

constexpr int GR_THREADS = 256;
constexpr int GR_WARPS = GR_THREADS / 32;
constexpr int GR_MAX_ROWS = 512;             // rows per tile
constexpr int GR_STAGE = 56 * 1024;          // staging bytes per CTA
constexpr size_t GR_SMEM = GR_STAGE + 32 + (size_t)GR_MAX_ROWS * (8 + 8 + 4 + 4) + 64;

__device__ __forceinline__ void cp_async16(uint32_t smem_addr, const void* g) {
  asm volatile("cp.async.cg.shared.global.L2::64B [%0], [%1], 16;" ::"r"(smem_addr), "l"(g) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() {
  asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory");
}

// whole-CTA global->global copy for a row too large for the staging buffer
__device__ void copy_big_row(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, int64_t len) {
  const bool a16 = ((((uintptr_t)src) | ((uintptr_t)dst)) & 15) == 0;
  int64_t done = 0;
  if (a16) {
    const int64_t nv = len / 16;
    for (int64_t k = threadIdx.x; k < nv; k += GR_THREADS)
      st_stream_u4((uint4*)dst + k, ld_stream_u4((const uint4*)src + k));
    done = nv * 16;
  }
  for (int64_t k = done + threadIdx.x; k < len; k += GR_THREADS) dst[k] = src[k];
}

// SPANS: in_off holds the source START of every OUTPUT row (k_spans_reduce), idx is not read
template <bool UNIFORM, bool SPANS = false>
__global__ void __launch_bounds__(GR_THREADS)
k_gather_rows(const uint8_t* __restrict__ in, const int64_t* __restrict__ in_off, uint32_t width,
              const uint32_t* __restrict__ idx, uint64_t n_out, const int64_t* __restrict__ out_off,
              uint8_t* __restrict__ out, uint32_t rows_per_tile) {
  extern __shared__ __align__(16) uint8_t smem[];
  uint8_t* stage = smem;                                        // GR_STAGE (+32 pad)
  int64_t* m_src = reinterpret_cast<int64_t*>(smem + GR_STAGE + 32);  // source byte offset
  int64_t* m_dst = m_src + GR_MAX_ROWS;                         // destination byte offset
  uint32_t* m_len = reinterpret_cast<uint32_t*>(m_dst + GR_MAX_ROWS);
  uint32_t* m_slot = m_len + GR_MAX_ROWS;                       // slot offset in stage
  uint32_t* s_misc = m_slot + GR_MAX_ROWS;                      // [0] round end, [1..8] warp sums
  const uint32_t stage_s = (uint32_t)__cvta_generic_to_shared(stage);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint64_t ntiles = (n_out + rows_per_tile - 1) / rows_per_tile;

  for (uint64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const uint64_t r0 = tile * rows_per_tile;
    const int T = (int)min((uint64_t)rows_per_tile, n_out - r0);

    // ---- phase 0: row metadata + slot offsets (exclusive scan of slot sizes) ----
    // slot = the row's bytes widened to the enclosing 16-byte-aligned source window
    uint32_t my_size[GR_MAX_ROWS / GR_THREADS];
    uint32_t my_sum = 0;
#pragma unroll
    for (int k = 0; k < GR_MAX_ROWS / GR_THREADS; ++k) {
      const int j = tid * (GR_MAX_ROWS / GR_THREADS) + k;  // blocked: scan order = row order
      uint32_t sz = 0;
      if (j < T) {
        const uint64_t r = r0 + j;
        int64_t s, d;
        uint32_t len;
        if constexpr (SPANS) {
          s = in_off[r];
          d = out_off[r];
          len = (uint32_t)(out_off[r + 1] - d);
        } else if constexpr (UNIFORM) {
          const uint32_t src_row = idx[r];
          s = (int64_t)src_row * width;
          d = (int64_t)r * width;
          len = width;
        } else {
          const uint32_t src_row = idx[r];
          s = in_off[src_row];
          len = (uint32_t)(in_off[src_row + 1] - s);
          d = out_off[r];
        }
        m_src[j] = s;
        m_dst[j] = d;
        m_len[j] = len;
        const uint32_t mis = (uint32_t)(((uintptr_t)in + s) & 15);
        sz = len ? ((mis + len + 15u) & ~15u) : 0u;
      }
      my_size[k] = sz;
      my_sum += sz;
    }
    {
      uint32_t incl = my_sum;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        uint32_t y = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += y;
      }
      if (lane == 31) s_misc[1 + warp] = incl;
      __syncthreads();
      uint32_t base = 0;
      for (int w = 0; w < warp; ++w) base += s_misc[1 + w];
      uint32_t run = base + incl - my_sum;
#pragma unroll
      for (int k = 0; k < GR_MAX_ROWS / GR_THREADS; ++k) {
        const int j = tid * (GR_MAX_ROWS / GR_THREADS) + k;
        if (j < T) m_slot[j] = run;
        run += my_size[k];
      }
    }
    __syncthreads();

    // ---- rounds: as many consecutive rows as fit in the staging buffer ----
    int round_start = 0;
    while (round_start < T) {
      const uint32_t base_slot = m_slot[round_start];
      if (tid == 0) s_misc[0] = (uint32_t)T;
      __syncthreads();
      for (int j = round_start + tid; j < T; j += GR_THREADS) {
        const uint32_t mis = (uint32_t)(((uintptr_t)in + m_src[j]) & 15);
        const uint32_t sz = m_len[j] ? ((mis + m_len[j] + 15u) & ~15u) : 0u;
        if (m_slot[j] - base_slot + sz > (uint32_t)GR_STAGE) atomicMin(&s_misc[0], (uint32_t)j);
      }
      __syncthreads();
      const int round_end = (int)s_misc[0];
      __syncthreads();
      if (round_end == round_start) {
        copy_big_row(in + m_src[round_start], out + m_dst[round_start], (int64_t)m_len[round_start]);
        round_start += 1;
        continue;
      }

      // ---- phase 1: global -> stage, 16 B cp.async ----
      // short rows (a slot of 2-3 windows: string keys): one lane per row -- with 8 lanes per
      // row a warp covered 4 rows per trip and most lanes idled (ncu: 36 % of the kernel's
      // instructions on 20-byte rows); longer rows: 8 lanes per row
      const int64_t round_out_bytes =
          (round_end < T ? m_dst[round_end] : m_dst[T - 1] + (int64_t)m_len[T - 1]) - m_dst[round_start];
      if (round_out_bytes <= (int64_t)40 * (round_end - round_start)) {
        for (int j = round_start + tid; j < round_end; j += GR_THREADS) {
          const uint32_t len = m_len[j];
          if (len == 0) continue;
          const uint8_t* g = in + m_src[j];
          const uint32_t mis = (uint32_t)((uintptr_t)g & 15);
          const uint8_t* g16 = g - mis;
          const uint32_t nchunks = (mis + len + 15u) >> 4;
          const uint32_t so = stage_s + (m_slot[j] - base_slot);
          for (uint32_t c = 0; c < nchunks; ++c) cp_async16(so + 16 * c, g16 + 16 * (size_t)c);
        }
        cp_async_wait_all();
      } else {
        const int sub = lane >> 3, sl = lane & 7;
        for (int j = round_start + warp * 4 + sub; j < round_end; j += GR_WARPS * 4) {
          const uint32_t len = m_len[j];
          if (len == 0) continue;
          const uint8_t* g = in + m_src[j];
          const uint32_t mis = (uint32_t)((uintptr_t)g & 15);
          const uint8_t* g16 = g - mis;
          const uint32_t nchunks = (mis + len + 15u) >> 4;
          const uint32_t so = stage_s + (m_slot[j] - base_slot);
          for (uint32_t c = sl; c < nchunks; c += 8) cp_async16(so + 16 * c, g16 + 16 * (size_t)c);
        }
        cp_async_wait_all();
      }
      __syncthreads();

      // ---- phase 2: stage -> global, FLAT over the round's output bytes ----
      // The rows of a round are consecutive output rows, so their bytes are ONE contiguous
      // range of `out`: thread q produces the q-th 16-byte-aligned unit of it (one 16-byte
      // store; a warp writes 512 dense bytes).  The unit's row is found by binary search over
      // the rows' output offsets; a unit inside one row is two unaligned 16-byte windows of the
      // stage joined by funnel shifts, a unit that spans rows (every unit of short rows) is
      // put together bytewise.  One warp per row cost ~150 thread instructions per 16 bytes on
      // 136-byte rows (the round-1 kernel was issue-bound at a quarter of the uniform gather's
      // rate); this mapping needs about a third of that.
      {
        const int64_t o0 = m_dst[round_start];
        const int64_t o1 = round_end < T ? m_dst[round_end] : m_dst[T - 1] + (int64_t)m_len[T - 1];
        uint8_t* ob = out + o0;
        const uint32_t nbytes = (uint32_t)(o1 - o0);
        const uint32_t head = min(nbytes, (uint32_t)((16 - ((uintptr_t)ob & 15)) & 15));
        const uint32_t nunits = (nbytes - head) >> 4;
        const uint32_t tail0 = head + (nunits << 4);
        auto row_of = [&](int64_t o) -> int {  // last row of the round that starts at or before byte o
          int lo = round_start, hi = round_end;
          while (hi - lo > 1) {
            const int mid = (lo + hi) >> 1;
            if (m_dst[mid] <= o) lo = mid;
            else hi = mid;
          }
          return lo;
        };
        auto stage_addr = [&](int j) -> uint32_t {  // stage byte address of row j's first byte
          return (m_slot[j] - base_slot) + (uint32_t)(((uintptr_t)in + m_src[j]) & 15);
        };
        uint4* ov = reinterpret_cast<uint4*>(ob + head);
        const uint32_t* sw = reinterpret_cast<const uint32_t*>(stage);
        // Short rows (string keys): nearly every unit spans rows and is assembled byte by byte
        // after a binary search (ncu: 45 % of the kernel's instructions on 20-byte rows).  When
        // the round's output also fits behind its slots in the staging buffer, the rows are
        // first packed there -- one thread per row, at the same 16-byte phase as `ob` -- and the
        // packed bytes leave as aligned 16-byte units.
        uint32_t slots_end;  // staging bytes the round's slots take
        {
          const int jl = round_end - 1;
          const uint32_t mis = (uint32_t)(((uintptr_t)in + m_src[jl]) & 15);
          slots_end = (m_slot[jl] - base_slot) + (m_len[jl] ? ((mis + m_len[jl] + 15u) & ~15u) : 0u);
        }
        const uint32_t pack0 = ((slots_end + 15u) & ~15u) + (uint32_t)((uintptr_t)ob & 15);
        if (nbytes <= 40u * (uint32_t)(round_end - round_start) && pack0 + nbytes + 16u <= (uint32_t)GR_STAGE) {
          uint8_t* pk = stage + pack0;
          for (int j = round_start + tid; j < round_end; j += GR_THREADS) {
            const uint32_t len = m_len[j];
            const uint8_t* sp = stage + stage_addr(j);
            uint8_t* dp = pk + (uint32_t)(m_dst[j] - o0);
            for (uint32_t b = 0; b < len; ++b) dp[b] = sp[b];
          }
          __syncthreads();
          const uint4* pv = reinterpret_cast<const uint4*>(pk + head);
          for (uint32_t q = tid; q < nunits; q += GR_THREADS) ov[q] = pv[q];
          if ((uint32_t)tid < head + (nbytes - tail0)) {
            const uint32_t t = (uint32_t)tid < head ? (uint32_t)tid : tail0 + ((uint32_t)tid - head);
            ob[t] = pk[t];
          }
          __syncthreads();
          round_start = round_end;
          continue;
        }
        for (uint32_t q = tid; q < nunits; q += GR_THREADS) {
          const int64_t o = o0 + head + ((int64_t)q << 4);
          int j = row_of(o);
          uint32_t r = (uint32_t)(o - m_dst[j]);
          uint32_t len = m_len[j];
          uint4 v;
          if (r + 16 <= len) {
            const uint32_t a = stage_addr(j) + r;
            const uint32_t* pa = sw + (a >> 2);
            const uint32_t sa = (a & 3) * 8;
            const uint32_t w0 = pa[0], w1 = pa[1], w2 = pa[2], w3 = pa[3], w4 = pa[4];
            v.x = __funnelshift_r(w0, w1, sa);
            v.y = __funnelshift_r(w1, w2, sa);
            v.z = __funnelshift_r(w2, w3, sa);
            v.w = __funnelshift_r(w3, w4, sa);
          } else {
            uint32_t x[4] = {0u, 0u, 0u, 0u};
            uint32_t a = stage_addr(j) + r;
#pragma unroll
            for (int bi = 0; bi < 16; ++bi) {
              while (r >= len) {  // next non-empty row (the unit lies inside the round's bytes)
                ++j;
                r = 0;
                len = m_len[j];
                a = stage_addr(j);
              }
              x[bi >> 2] |= (uint32_t)stage[a] << (8 * (bi & 3));
              ++a;
              ++r;
            }
            v = make_uint4(x[0], x[1], x[2], x[3]);
          }
          ov[q] = v;
        }
        // head and tail bytes of the round's range (at most 15 each)
        if ((uint32_t)tid < head + (nbytes - tail0)) {
          const uint32_t t = (uint32_t)tid < head ? (uint32_t)tid : tail0 + ((uint32_t)tid - head);
          const int64_t o = o0 + t;
          const int j = row_of(o);
          ob[t] = stage[stage_addr(j) + (uint32_t)(o - m_dst[j])];
        }
      }
      __syncthreads();
      round_start = round_end;
    }
    __syncthreads();
  }
}

int grid_for(uint64_t n, int threads, int per_sm) {
  const int cap = xchg_gather_ctas_per_sm();
  if (cap) per_sm = std::min(per_sm, cap);
  return (int)std::max<uint64_t>(1, std::min<uint64_t>((n + threads - 1) / threads,
                                                       (uint64_t)xchg_sm_count() * per_sm));
}

struct alignas(16) Bytes16 {
  uint32_t x, y, z, w;
};

}  // namespace

extern "C" int xchg_gather_rows(const uint8_t* in_data, uint32_t width, const uint32_t* idx,
                                uint64_t n_out, uint8_t* out_data, void* stream);
int xchg_internal_gather_rows_uniform(const uint8_t* in, uint32_t w, const uint32_t* idx,
                                      uint64_t n_out, uint8_t* out, cudaStream_t s);

extern "C" int xchg_gather_fixed(const void* in, uint32_t width, const uint32_t* idx, uint64_t n_out,
                                 void* out, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  if (n_out == 0) return XCHG_OK;
  XCHG_REQUIRE(in && idx && out, "NULL in / idx / out");
  XCHG_REQUIRE(width > 0, "width must be > 0");
  const int g = grid_for(n_out, 256, 8);
  const uintptr_t al = (uintptr_t)in | (uintptr_t)out;
  if (width == 1) {
    XCHG_LAUNCH(XK_GATHER_FIXED, s, k_gather_fixed<uint8_t><<<g, 256, 0, s>>>((const uint8_t*)in, idx, n_out, (uint8_t*)out));
  } else if (width == 2 && al % 2 == 0) {
    XCHG_LAUNCH(XK_GATHER_FIXED, s, k_gather_fixed<uint16_t><<<g, 256, 0, s>>>((const uint16_t*)in, idx, n_out, (uint16_t*)out));
  } else if (width == 4 && al % 4 == 0) {
    XCHG_LAUNCH(XK_GATHER_FIXED, s, k_gather_fixed<uint32_t><<<g, 256, 0, s>>>((const uint32_t*)in, idx, n_out, (uint32_t*)out));
  } else if (width == 8 && al % 8 == 0) {
    XCHG_LAUNCH(XK_GATHER_FIXED, s, k_gather_fixed<uint64_t><<<g, 256, 0, s>>>((const uint64_t*)in, idx, n_out, (uint64_t*)out));
  } else if (width == 16 && al % 16 == 0) {
    XCHG_LAUNCH(XK_GATHER_FIXED, s, k_gather_fixed<Bytes16><<<g, 256, 0, s>>>((const Bytes16*)in, idx, n_out, (Bytes16*)out));
  } else {
    return xchg_gather_rows((const uint8_t*)in, width, idx, n_out, (uint8_t*)out, stream);
  }
  XCHG_LAUNCH_CHECK();
  return XCHG_OK;
}

extern "C" int xchg_gather_bitmap(const uint8_t* in, int64_t in_bit_offset, const uint32_t* idx,
                                  uint64_t n_out, uint8_t* out, uint64_t* null_count_out,
                                  void* stream) {
  if (n_out == 0) return XCHG_OK;
  XCHG_REQUIRE(in && idx && out, "NULL in / idx / out");
  XCHG_REQUIRE(((uintptr_t)out & 3) == 0, "bitmap output must be 4-byte aligned (and padded to 4 bytes)");
  const int g = grid_for(n_out, 256, 8);
  XCHG_LAUNCH(XK_GATHER_BITMAP, (cudaStream_t)stream, k_gather_bitmap<<<g, 256, 0, (cudaStream_t)stream>>>(in, in_bit_offset, idx, n_out, (uint32_t*)out,
                                                        (unsigned long long*)null_count_out));
  XCHG_LAUNCH_CHECK();
  return XCHG_OK;
}

extern "C" int xchg_gather_offsets_workspace_bytes(uint64_t n_out, uint64_t* bytes_out) {
  XCHG_REQUIRE(bytes_out != nullptr, "bytes_out is NULL");
  *bytes_out = ((n_out + SC_TILE - 1) / SC_TILE + 1) * 8 + 256;
  return XCHG_OK;
}

extern "C" int xchg_gather_offsets(const int64_t* in_off, const uint32_t* idx, uint64_t n_out,
                                   int64_t* out_off, void* workspace, uint64_t workspace_bytes,
                                   void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  XCHG_REQUIRE(out_off != nullptr, "out_off is NULL");
  if (n_out == 0) {
    XCHG_CUDA_TRY(cudaMemsetAsync(out_off, 0, 8, s));
    return XCHG_OK;
  }
  XCHG_REQUIRE(in_off && idx && workspace, "NULL in_off / idx / workspace");
  const uint64_t ntiles = (n_out + SC_TILE - 1) / SC_TILE;
  int64_t* tiles = (int64_t*)xchg_align_up((uint64_t)(uintptr_t)workspace, 8);
  if ((uint64_t)((uint8_t*)tiles - (uint8_t*)workspace) + ntiles * 8 > workspace_bytes) {
    xchg_set_error("gather_offsets workspace too small");
    return XCHG_ERR_WORKSPACE;
  }
  XCHG_LAUNCH(XK_GATHER_OFFSETS, s, k_lengths_reduce<<<(unsigned)ntiles, SC_THREADS, 0, s>>>(in_off, idx, n_out, tiles));
  XCHG_LAUNCH_CHECK();
  XCHG_LAUNCH(XK_GATHER_OFFSETS, s, k_scan_tiles<<<1, SC_THREADS, 0, s>>>(tiles, ntiles));
  XCHG_LAUNCH_CHECK();
  XCHG_LAUNCH(XK_GATHER_OFFSETS, s, k_lengths_scan<<<(unsigned)ntiles, SC_THREADS, 0, s>>>(in_off, idx, n_out, tiles, out_off));
  XCHG_LAUNCH_CHECK();
  return XCHG_OK;
}

static int launch_gather_rows(bool uniform, const uint8_t* in, const int64_t* in_off, uint32_t width,
                              const uint32_t* idx, uint64_t n_out, const int64_t* out_off,
                              uint8_t* out, uint32_t rows_per_tile, cudaStream_t s, bool spans = false) {
  XCHG_ONCE_PER_DEVICE({
    XCHG_CUDA_TRY((cudaFuncSetAttribute(k_gather_rows<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        (int)GR_SMEM)));
    XCHG_CUDA_TRY(cudaFuncSetAttribute(k_gather_rows<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)GR_SMEM));
    XCHG_CUDA_TRY(cudaFuncSetAttribute(k_gather_rows<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)GR_SMEM));
  });
  const uint64_t ntiles = (n_out + rows_per_tile - 1) / rows_per_tile;
  const int cap = xchg_gather_ctas_per_sm();
  const int grid = (int)std::min<uint64_t>(ntiles, (uint64_t)xchg_sm_count() * (cap ? cap : 3 * 4));
  if (spans)
    XCHG_LAUNCH(XK_GATHER_ROWS, s, (k_gather_rows<false, true><<<grid, GR_THREADS, GR_SMEM, s>>>(in, in_off, width, idx, n_out, out_off,
                                                                                            out, rows_per_tile)));
  else if (uniform)
    XCHG_LAUNCH(XK_GATHER_ROWS, s, k_gather_rows<true><<<grid, GR_THREADS, GR_SMEM, s>>>(in, in_off, width, idx, n_out, out_off, out,
                                                          rows_per_tile));
  else
    XCHG_LAUNCH(XK_GATHER_ROWS, s, k_gather_rows<false><<<grid, GR_THREADS, GR_SMEM, s>>>(in, in_off, width, idx, n_out, out_off, out,
                                                           rows_per_tile));
  XCHG_LAUNCH_CHECK();
  return XCHG_OK;
}

extern "C" int xchg_gather_rows(const uint8_t* in_data, uint32_t width, const uint32_t* idx,
                                uint64_t n_out, uint8_t* out_data, void* stream) {
  if (n_out == 0 || width == 0) return XCHG_OK;
  XCHG_REQUIRE(in_data && idx && out_data, "NULL in / idx / out");
  {
    // flat word-per-thread kernel (gather_uniform.cu); rc 1 = row too wide for it
    const int rc = xchg_internal_gather_rows_uniform(in_data, width, idx, n_out, out_data,
                                                     (cudaStream_t)stream);
    if (rc != 1) return rc;
  }
  // rows per tile: fill the staging buffer (slot <= width + 30 bytes)
  const uint32_t slot = (width + 15u + 15u) & ~15u;
  uint32_t rpt = std::max<uint32_t>(1, std::min<uint32_t>(GR_MAX_ROWS, GR_STAGE / slot));
  return launch_gather_rows(true, in_data, nullptr, width, idx, n_out, nullptr, out_data, rpt,
                            (cudaStream_t)stream);
}

extern "C" int xchg_gather_bytes(const uint8_t* in_data, const int64_t* in_off, const uint32_t* idx,
                                 uint64_t n_out, const int64_t* out_off, uint8_t* out_data,
                                 void* stream) {
  if (n_out == 0) return XCHG_OK;
  XCHG_REQUIRE(in_off && idx && out_off, "NULL in_off / idx / out_off");
  XCHG_REQUIRE(in_data && out_data, "NULL data pointers");
  return launch_gather_rows(false, in_data, in_off, 0, idx, n_out, out_off, out_data, GR_MAX_ROWS,
                            (cudaStream_t)stream);
}

extern "C" int xchg_gather_spans(const int64_t* in_off, const uint32_t* idx, uint64_t n_out,
                                 int64_t* src_start_out, int64_t* out_off, void* workspace,
                                 uint64_t workspace_bytes, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  XCHG_REQUIRE(out_off != nullptr, "out_off is NULL");
  if (n_out == 0) {
    XCHG_CUDA_TRY(cudaMemsetAsync(out_off, 0, 8, s));
    return XCHG_OK;
  }
  XCHG_REQUIRE(in_off && idx && src_start_out && workspace, "NULL in_off / idx / src_start_out / workspace");
  const uint64_t ntiles = (n_out + SC_TILE - 1) / SC_TILE;
  int64_t* tiles = (int64_t*)xchg_align_up((uint64_t)(uintptr_t)workspace, 8);
  if ((uint64_t)((uint8_t*)tiles - (uint8_t*)workspace) + ntiles * 8 > workspace_bytes) {
    xchg_set_error("gather_spans workspace too small");
    return XCHG_ERR_WORKSPACE;
  }
  XCHG_LAUNCH(XK_GATHER_OFFSETS, s,
              k_spans_reduce<<<(unsigned)ntiles, SC_THREADS, 0, s>>>(in_off, idx, n_out, src_start_out, out_off, tiles));
  XCHG_LAUNCH_CHECK();
  XCHG_LAUNCH(XK_GATHER_OFFSETS, s, k_scan_tiles<<<1, SC_THREADS, 0, s>>>(tiles, ntiles));
  XCHG_LAUNCH_CHECK();
  XCHG_LAUNCH(XK_GATHER_OFFSETS, s, k_spans_scan<<<(unsigned)ntiles, SC_THREADS, 0, s>>>(n_out, tiles, out_off));
  XCHG_LAUNCH_CHECK();
  return XCHG_OK;
}

extern "C" int xchg_gather_bytes_spans(const uint8_t* in_data, const int64_t* src_start, uint64_t n_out,
                                       const int64_t* out_off, uint8_t* out_data, void* stream) {
  if (n_out == 0) return XCHG_OK;
  XCHG_REQUIRE(src_start && out_off, "NULL src_start / out_off");
  XCHG_REQUIRE(in_data && out_data, "NULL data pointers");
  return launch_gather_rows(false, in_data, src_start, 0, nullptr, n_out, out_off, out_data, GR_MAX_ROWS,
                            (cudaStream_t)stream, true);
}

extern "C" int xchg_offsets_widen(const int32_t* in, uint64_t count, int64_t base, int64_t* out,
                                  void* stream) {
  if (count == 0) return XCHG_OK;
  XCHG_REQUIRE(in && out, "NULL in / out");
  XCHG_LAUNCH(XK_OFFSETS_UTIL, (cudaStream_t)stream, k_offsets_widen<<<grid_for(count, 256, 8), 256, 0, (cudaStream_t)stream>>>(in, count, base, out));
  XCHG_LAUNCH_CHECK();
  return XCHG_OK;
}

extern "C" int xchg_offsets_narrow(const int64_t* in, uint64_t count, int32_t* out, void* stream) {
  if (count == 0) return XCHG_OK;
  XCHG_REQUIRE(in && out, "NULL in / out");
  XCHG_LAUNCH(XK_OFFSETS_UTIL, (cudaStream_t)stream, k_offsets_narrow<<<grid_for(count, 256, 8), 256, 0, (cudaStream_t)stream>>>(in, count, out));
  XCHG_LAUNCH_CHECK();
  return XCHG_OK;
}

extern "C" int xchg_offsets_uniform32(uint64_t count, uint32_t width, int32_t* out, void* stream) {
  if (count == 0) return XCHG_OK;
  XCHG_REQUIRE(out != nullptr, "NULL out");
  if ((count - 1) * (uint64_t)width > 0x7FFFFFFFull) {
    xchg_set_error("uniform int32 offsets overflow: %llu rows x %u bytes",
                   (unsigned long long)(count - 1), width);
    return XCHG_ERR_LIMIT;
  }
  XCHG_LAUNCH(XK_OFFSETS_UTIL, (cudaStream_t)stream, k_offsets_uniform32<<<grid_for(count, 256, 8), 256, 0, (cudaStream_t)stream>>>(count, width, out));
  XCHG_LAUNCH_CHECK();
  return XCHG_OK;
}

extern "C" int xchg_offsets_uniform64(uint64_t count, uint64_t width, int64_t base, int64_t* out,
                                      void* stream) {
  if (count == 0) return XCHG_OK;
  XCHG_REQUIRE(out != nullptr, "NULL out");
  XCHG_LAUNCH(XK_OFFSETS_UTIL, (cudaStream_t)stream, k_offsets_uniform64<<<grid_for(count, 256, 8), 256, 0, (cudaStream_t)stream>>>(count, width, base, out));
  XCHG_LAUNCH_CHECK();
  return XCHG_OK;
}

extern "C" int xchg_offsets_check_uniform(const int64_t* off, uint64_t n_rows, uint64_t width,
                                          uint32_t* flag_out, void* stream) {
  if (n_rows == 0) return XCHG_OK;
  XCHG_REQUIRE(off && flag_out, "NULL off / flag_out");
  XCHG_LAUNCH(XK_OFFSETS_UTIL, (cudaStream_t)stream, k_offsets_check_uniform<<<grid_for(n_rows, 256, 8), 256, 0, (cudaStream_t)stream>>>(off, n_rows, width,
                                                                                     flag_out));
  XCHG_LAUNCH_CHECK();
  return XCHG_OK;
}

// ------------------------------------------------- allocation-free variant ---
//
// xchg_gather_offsets_static: same result as xchg_gather_offsets, but the scan's
// scratch (one int64 partial sum per 2048-row tile) lives in a fixed array inside
// the library image instead of a caller-supplied workspace.
//
// Who needs it: hosts that receive their input and output device buffers from
// somebody else and own no device allocator themselves -- a plain ctypes/cffi
// embedding inside a worker whose GPU memory has been handed over wholesale to
// another component's pool (a framework caching allocator, an RMM-style arena), or
// a latency-critical stage that must not call cudaMalloc/cudaFree because those
// calls serialise with all work on the device.  Such callers can pass the buffers
// they were given, but cannot produce an extra workspace.
//
// Cost and limits: the array (XCHG_STATIC_GATHER_MAX_ROWS / 2048 x 8 bytes =
// 256 KiB) is part of every CUDA context that loads this library and is held
// until the process exits.  Requests above XCHG_STATIC_GATHER_MAX_ROWS are
// refused.  All calls on a device share the one array: two calls whose kernels
// may overlap on the GPU (different streams, or different host threads on
// different streams) race on it -- serialise them.  Nothing in the array has to
// start at zero: every slot a call reads was written earlier by that same call,
// so contents left behind by the previous call are harmless.
//
// Set-up: the array's name is only a host-side handle.  The first call on each
// device (with that device current) resolves its device address with
// cudaGetSymbolAddress and caches it; make that first call outside any stream
// capture.
constexpr uint64_t STATIC_MAX_TILES = XCHG_STATIC_GATHER_MAX_ROWS / SC_TILE;
static_assert(XCHG_STATIC_GATHER_MAX_ROWS % SC_TILE == 0, "limit must be a whole number of tiles");

static __device__ __align__(256) int64_t g_static_tile_sums[STATIC_MAX_TILES];

namespace {
constexpr int MAX_DEVICES = 64;
std::mutex g_static_mu;
int64_t* g_static_addr[MAX_DEVICES] = {nullptr};

// Device address of g_static_tile_sums on the current device (resolved once).
int static_scratch(int64_t** out) {
  int dev = -1;
  XCHG_CUDA_TRY(cudaGetDevice(&dev));
  XCHG_REQUIRE(dev >= 0 && dev < MAX_DEVICES, "device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lk(g_static_mu);
  if (g_static_addr[dev] == nullptr) {
    int64_t* ptr = nullptr;
    XCHG_CUDA_TRY(cudaGetSymbolAddress((void**)&ptr, g_static_tile_sums));
    XCHG_REQUIRE(ptr != nullptr, "cudaGetSymbolAddress returned NULL");
    g_static_addr[dev] = ptr;
  }
  *out = g_static_addr[dev];
  return XCHG_OK;
}
}  // namespace

extern "C" int xchg_gather_offsets_static(const int64_t* in_off, const uint32_t* idx,
                                          uint64_t n_out, int64_t* out_off, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  XCHG_REQUIRE(out_off != nullptr, "out_off is NULL");
  if (n_out > XCHG_STATIC_GATHER_MAX_ROWS) {
    xchg_set_error("xchg_gather_offsets_static: n_out=%llu exceeds the built-in scratch "
                   "(XCHG_STATIC_GATHER_MAX_ROWS=%llu rows); use xchg_gather_offsets with a "
                   "caller workspace instead",
                   (unsigned long long)n_out, (unsigned long long)XCHG_STATIC_GATHER_MAX_ROWS);
    return XCHG_ERR_LIMIT;
  }
  if (n_out == 0) {
    XCHG_CUDA_TRY(cudaMemsetAsync(out_off, 0, 8, s));
    return XCHG_OK;
  }
  XCHG_REQUIRE(in_off && idx, "NULL in_off / idx");
  int64_t* tiles = nullptr;
  int rc = static_scratch(&tiles);
  if (rc) return rc;
  const uint64_t ntiles = (n_out + SC_TILE - 1) / SC_TILE;
  XCHG_LAUNCH(XK_GATHER_OFFSETS, s, k_lengths_reduce<<<(unsigned)ntiles, SC_THREADS, 0, s>>>(in_off, idx, n_out, tiles));
  XCHG_LAUNCH_CHECK();
  XCHG_LAUNCH(XK_GATHER_OFFSETS, s, k_scan_tiles<<<1, SC_THREADS, 0, s>>>(tiles, ntiles));
  XCHG_LAUNCH_CHECK();
  XCHG_LAUNCH(XK_GATHER_OFFSETS, s, k_lengths_scan<<<(unsigned)ntiles, SC_THREADS, 0, s>>>(in_off, idx, n_out, tiles, out_off));
  XCHG_LAUNCH_CHECK();
  return XCHG_OK;
}

// ------------------------------------------------ exchange-side utilities ---
// Validity bitmaps cross the all-to-all as one byte per row (per-destination row
// ranges of a bit-packed buffer are not byte aligned); offsets cross as the
// sender's absolute values and are re-based on arrival.
namespace {

__global__ void __launch_bounds__(256)
k_bitmap_unpack(const uint8_t* __restrict__ in, int64_t bit0, uint64_t n, uint8_t* __restrict__ out) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    out[i] = bit_is_set(in, (int64_t)i + bit0) ? 1 : 0;
}

__global__ void __launch_bounds__(256)
k_bitmap_pack(const uint8_t* __restrict__ in, uint64_t n, uint32_t* __restrict__ out_words,
              unsigned long long* zero_count) {
  const uint64_t nwords = (n + 31) / 32;
  const int lane = threadIdx.x & 31;
  const uint64_t warp0 = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const uint64_t nwarps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
  unsigned long long zeros = 0;
  for (uint64_t w = warp0; w < nwords; w += nwarps) {
    const uint64_t i = w * 32 + lane;
    const bool bit = i < n ? in[i] != 0 : false;
    const unsigned word = __ballot_sync(0xffffffffu, bit);
    if (lane == 0) {
      out_words[w] = word;
      const unsigned live = (w * 32 + 32 <= n) ? 0xffffffffu : ((1u << (n - w * 32)) - 1u);
      zeros += __popc(~word & live);
    }
  }
  if (zero_count && lane == 0 && zeros) atomicAdd(zero_count, zeros);
}

__global__ void k_offsets_rebase(const int64_t* __restrict__ in, uint64_t count, int64_t base,
                                 int64_t* __restrict__ out) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  const int64_t first = in[0];
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride)
    out[i] = in[i] - first + base;
}

}  // namespace

namespace {
__global__ void __launch_bounds__(256)
k_gather_bitmap_bytes(const uint8_t* __restrict__ in, int64_t in_bit, const uint32_t* __restrict__ idx,
                      uint64_t n, uint8_t* __restrict__ out) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    out[i] = bit_is_set(in, (int64_t)idx[i] + in_bit) ? 1 : 0;
}
}  // namespace

extern "C" int xchg_gather_bitmap_bytes(const uint8_t* in, int64_t in_bit_offset, const uint32_t* idx,
                                        uint64_t n_out, uint8_t* out_bytes, void* stream) {
  if (n_out == 0) return XCHG_OK;
  XCHG_REQUIRE(in && idx && out_bytes, "NULL in / idx / out");
  XCHG_LAUNCH(XK_GATHER_BITMAP, (cudaStream_t)stream, k_gather_bitmap_bytes<<<grid_for(n_out, 256, 8), 256, 0, (cudaStream_t)stream>>>(in, in_bit_offset, idx, n_out, out_bytes));
  XCHG_LAUNCH_CHECK();
  return XCHG_OK;
}

extern "C" int xchg_bitmap_unpack(const uint8_t* in, int64_t in_bit_offset, uint64_t n, uint8_t* out,
                                  void* stream) {
  if (n == 0) return XCHG_OK;
  XCHG_REQUIRE(in && out, "NULL in / out");
  XCHG_LAUNCH(XK_GATHER_BITMAP, (cudaStream_t)stream, k_bitmap_unpack<<<grid_for(n, 256, 8), 256, 0, (cudaStream_t)stream>>>(in, in_bit_offset, n, out));
  XCHG_LAUNCH_CHECK();
  return XCHG_OK;
}

extern "C" int xchg_bitmap_pack(const uint8_t* in, uint64_t n, uint8_t* out, uint64_t* zero_count_out,
                                void* stream) {
  if (n == 0) return XCHG_OK;
  XCHG_REQUIRE(in && out, "NULL in / out");
  XCHG_REQUIRE(((uintptr_t)out & 3) == 0, "bitmap output must be 4-byte aligned (and padded to 4 bytes)");
  XCHG_LAUNCH(XK_GATHER_BITMAP, (cudaStream_t)stream, k_bitmap_pack<<<grid_for(n, 256, 8), 256, 0, (cudaStream_t)stream>>>(in, n, (uint32_t*)out, (unsigned long long*)zero_count_out));
  XCHG_LAUNCH_CHECK();
  return XCHG_OK;
}

extern "C" int xchg_offsets_rebase(const int64_t* in, uint64_t count, int64_t base, int64_t* out,
                                   void* stream) {
  if (count == 0) return XCHG_OK;
  XCHG_REQUIRE(in && out, "NULL in / out");
  XCHG_LAUNCH(XK_OFFSETS_UTIL, (cudaStream_t)stream, k_offsets_rebase<<<grid_for(count, 256, 8), 256, 0, (cudaStream_t)stream>>>(in, count, base, out));
  XCHG_LAUNCH_CHECK();
  return XCHG_OK;
}
