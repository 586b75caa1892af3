// Kernels of the multi-GPU exchange's control path, for sm_100a: K1 boundary sampling,
// boundary selection from the all-gathered samples, the per-destination row histogram,
// contiguous segment copies into peer memory and the arrival flags that order the
// chunks of an exchange between GPUs.
//
// In the reference these are: one Ray task per block drawing `random.sample` rows
// (python/ray/data/_internal/planner/exchange/sort_task_spec.py:229 ->
// _internal/arrow_block.py:341-344), a driver-side Python sort of the samples with the
// quantile pick (sort_task_spec.py:200-226), and the object store's transfers plus task
// dependencies between map, merge and reduce rounds
// (push_based_shuffle_task_scheduler.py:548-561).  Here nothing leaves the GPUs: samples
// cross in one all-gather, every rank picks the same boundaries with one small kernel and
// keeps them in HBM for the partition kernels, and a chunk's arrival is a 4-byte flag in
// the receiver's memory instead of a task dependency.
#include <algorithm>
#include <cstdlib>

#include "common.cuh"
#include "feistel.cuh"
#include "keynorm.cuh"

using namespace xchg_feistel;

namespace {

__host__ __device__ __forceinline__ uint64_t splitmix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}

// ------------------------------------------------------------------ K1 -----

constexpr int SAMPLE_BLOCKS_PER_LAUNCH = 32;
struct SampleBlocks {
  uint64_t start[SAMPLE_BLOCKS_PER_LAUNCH];  // first row of the block in the column
  uint64_t rows[SAMPLE_BLOCKS_PER_LAUNCH];
  uint64_t seed[SAMPLE_BLOCKS_PER_LAUNCH];
  uint32_t pos[SAMPLE_BLOCKS_PER_LAUNCH];    // first output slot of the block
  uint32_t k[SAMPLE_BLOCKS_PER_LAUNCH];      // samples to draw: min(per_block, rows)
};

// One CTA per block.  k distinct rows: the first k values of a seeded bijection of
// [0, rows) (every row when k == rows).
template <typename T>
__global__ void __launch_bounds__(128)
k_sample_keys(const T* __restrict__ keys, const uint8_t* __restrict__ validity, int64_t vbit, int desc,
              SampleBlocks sb, uint64_t* __restrict__ out) {
  const int b = blockIdx.x;
  const uint64_t rows = sb.rows[b];
  const uint32_t k = sb.k[b];
  Feistel f;
  make_feistel(sb.seed[b], rows, &f);
  for (uint32_t j = threadIdx.x; j < k; j += blockDim.x) {
    const uint64_t r = (k == rows) ? (uint64_t)j : feistel_walk(f, j, false);
    const uint64_t row = sb.start[b] + r;
    const bool valid = validity ? bit_is_set(validity, (int64_t)row + vbit) : true;
    out[sb.pos[b] + j] = valid ? norm_value<T>(keys[row], desc != 0) : null_key<T>();
  }
}

// -------------------------------------------------- boundary selection -----

constexpr uint32_t SEL_CAP = 16384;  // samples one CTA sorts in shared memory (128 KB)
constexpr int SEL_THREADS = 1024;

// gathered = nmsgs messages of msg_words u64 each; word count_word of a message = the
// number of samples it carries, the samples start at word hdr_words.  Every rank runs this
// on the same gathered bytes and obtains the same boundaries.
__global__ void __launch_bounds__(SEL_THREADS)
k_select_boundaries(const uint64_t* __restrict__ gathered, int nmsgs, uint64_t msg_words,
                    uint32_t count_word, uint32_t hdr_words, uint32_t R, uint64_t* __restrict__ bounds,
                    uint64_t* __restrict__ info) {
  extern __shared__ __align__(16) uint8_t smem_raw[];
  uint64_t* s = reinterpret_cast<uint64_t*>(smem_raw);
  __shared__ uint64_t prefix[65];
  const int tid = threadIdx.x;
  if (tid == 0) {
    uint64_t acc = 0;
    for (int g = 0; g < nmsgs; ++g) {
      prefix[g] = acc;
      acc += gathered[(uint64_t)g * msg_words + count_word];
    }
    prefix[nmsgs] = acc;
  }
  __syncthreads();
  const uint64_t total = prefix[nmsgs];
  const uint64_t stride = total > SEL_CAP ? (total + SEL_CAP - 1) / SEL_CAP : 1;  // thin out evenly
  const uint32_t m = (uint32_t)((total + stride - 1) / stride);
  uint32_t P = 1;
  while (P < m) P <<= 1;
  for (uint32_t i = tid; i < P; i += SEL_THREADS) {
    uint64_t v = 0xFFFFFFFFFFFFFFFFull;
    if (i < m) {
      const uint64_t gidx = (uint64_t)i * stride;
      int g = 0;
      while (g + 1 < nmsgs && prefix[g + 1] <= gidx) ++g;
      v = gathered[(uint64_t)g * msg_words + hdr_words + (gidx - prefix[g])];
    }
    s[i] = v;
  }
  __syncthreads();
  // bitonic sort of P values
  for (uint32_t kk = 2; kk <= P; kk <<= 1) {
    for (uint32_t j = kk >> 1; j > 0; j >>= 1) {
      for (uint32_t i = tid; i < P; i += SEL_THREADS) {
        const uint32_t ixj = i ^ j;
        if (ixj > i) {
          const uint64_t a = s[i], b = s[ixj];
          const bool up = (i & kk) == 0;
          if ((a > b) == up) {
            s[i] = b;
            s[ixj] = a;
          }
        }
      }
      __syncthreads();
    }
  }
  // the reference's pick (sort_task_spec.py:221-226): sample int(q * (len - 1)) for
  // q in numpy.linspace(0, 1, R + 1)[1:-1], in float64 like numpy
  if (m > 0) {
    const double step = 1.0 / (double)R;
    for (uint32_t i = tid + 1; i < R; i += SEL_THREADS) {
      const double q = (double)i * step;
      const uint64_t idx = (uint64_t)(q * (double)(m - 1));
      bounds[i - 1] = s[idx < m ? idx : m - 1];
    }
  }
  if (tid == 0 && info) {
    info[0] = total;
    info[1] = m;
  }
}

// ------------------------------------------------ destination histogram -----

constexpr int PH_THREADS = 256;

template <typename T>
__global__ void __launch_bounds__(PH_THREADS)
k_partition_hist(const T* __restrict__ keys, uint64_t n, int desc, const uint64_t* __restrict__ bounds,
                 DestRule rule, unsigned long long* __restrict__ counts) {
  __shared__ uint64_t thr[XCHG_MAX_THRESHOLDS + 1];
  __shared__ uint8_t strict[XCHG_MAX_THRESHOLDS + 1];
  __shared__ uint32_t whist[PH_THREADS / 32][XCHG_MAX_THRESHOLDS + 1];
  const int tid = threadIdx.x, warp = tid >> 5;
  for (int j = tid; j < rule.nthr; j += PH_THREADS) {
    thr[j] = bounds[rule.pick[j]];
    strict[j] = rule.strict[j];
  }
  for (int j = tid; j < (PH_THREADS / 32) * (XCHG_MAX_THRESHOLDS + 1); j += PH_THREADS)
    (&whist[0][0])[j] = 0;
  __syncthreads();
  const uint64_t stride = (uint64_t)gridDim.x * PH_THREADS;
  for (uint64_t i = (uint64_t)blockIdx.x * PH_THREADS + tid; i < n; i += stride) {
    const uint64_t k = norm_value<T>(keys[i], desc != 0);
    atomicAdd(&whist[warp][dest_of(k, thr, strict, rule.nthr)], 1u);
  }
  __syncthreads();
  for (int d = tid; d <= rule.nthr; d += PH_THREADS) {
    uint32_t c = 0;
#pragma unroll
    for (int w = 0; w < PH_THREADS / 32; ++w) c += whist[w][d];
    if (c) atomicAdd(&counts[d], (unsigned long long)c);
  }
}

// ------------------------------------------------------- segment copies -----

constexpr int CS_MAX_SEGS = 16;
constexpr uint32_t CS_TILE = 32768;  // bytes per CTA step
struct CopySegs {
  const uint8_t* src[CS_MAX_SEGS];
  uint8_t* dst[CS_MAX_SEGS];
  uint64_t bytes[CS_MAX_SEGS];
  int nseg;
};

// Contiguous copies of several segments in one launch, their tiles interleaved (tile r of
// every segment, then tile r + 1, ...): the exchange's key runs reach every peer at the
// same pace.  16-byte units when source, destination and length allow, bytes otherwise.
__global__ void __launch_bounds__(256)
k_copy_segments(CopySegs cs, uint64_t max_tiles) {
  const uint64_t nwork = max_tiles * (uint64_t)cs.nseg;
  for (uint64_t work = blockIdx.x; work < nwork; work += gridDim.x) {
    const int sg = (int)(work % (uint64_t)cs.nseg);
    const uint64_t off = (work / (uint64_t)cs.nseg) * CS_TILE;
    if (off >= cs.bytes[sg]) continue;
    const uint32_t len = (uint32_t)min((uint64_t)CS_TILE, cs.bytes[sg] - off);
    const uint8_t* src = cs.src[sg] + off;
    uint8_t* dst = cs.dst[sg] + off;
    if ((((uintptr_t)src | (uintptr_t)dst) & 15) == 0) {
      const uint4* s4 = reinterpret_cast<const uint4*>(src);
      uint4* d4 = reinterpret_cast<uint4*>(dst);
      const uint32_t nv = len >> 4;
      for (uint32_t q = threadIdx.x; q < nv; q += 256) st_stream_u4(d4 + q, ld_stream_u4(s4 + q));
      for (uint32_t q = (nv << 4) + threadIdx.x; q < len; q += 256) dst[q] = src[q];
    } else if ((((uintptr_t)src | (uintptr_t)dst) & 7) == 0) {
      const uint64_t* s8 = reinterpret_cast<const uint64_t*>(src);
      uint64_t* d8 = reinterpret_cast<uint64_t*>(dst);
      const uint32_t nv = len >> 3;
      for (uint32_t q = threadIdx.x; q < nv; q += 256) d8[q] = s8[q];
      for (uint32_t q = (nv << 3) + threadIdx.x; q < len; q += 256) dst[q] = src[q];
    } else {
      for (uint32_t q = threadIdx.x; q < len; q += 256) dst[q] = src[q];
    }
  }
}

// The same copy by the TMA unit: one warp per CTA, lane 0 moves 32 KB stages global ->
// shared (cp.async.bulk + mbarrier) -> global (cp.async.bulk, bulk group), two stages so
// that the store of one stage drains while the next one loads.  No thread touches the data;
// towards a peer GPU the stores leave the SM as the TMA unit's large bursts instead of
// per-thread 16-byte stores.  Needs 16-byte aligned segments; tails (< 16 bytes) go bytewise.
constexpr uint32_t CB_STAGE = 32768;
constexpr int CB_STAGES = 2;

__device__ __forceinline__ void mbar_init(uint32_t mbar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(mbar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t mbar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mbar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t mbar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@!p bra WAIT_%=;\n\t}"
      ::"r"(mbar), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_load(uint32_t smem_dst, const void* gsrc, uint32_t bytes, uint32_t mbar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_dst), "l"(gsrc), "r"(bytes), "r"(mbar) : "memory");
}
__device__ __forceinline__ void bulk_store(void* gdst, uint32_t smem_src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
               ::"l"(gdst), "r"(smem_src), "r"(bytes) : "memory");
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}

__global__ void __launch_bounds__(32)
k_copy_segments_bulk(CopySegs cs, uint64_t max_tiles) {
  extern __shared__ __align__(128) uint8_t cb_smem[];
  __shared__ __align__(8) uint64_t mbar_mem[CB_STAGES];
  const uint32_t stage_s = (uint32_t)__cvta_generic_to_shared(cb_smem);
  const uint32_t mbar_s = (uint32_t)__cvta_generic_to_shared(mbar_mem);
  const int lane = threadIdx.x;
  if (lane == 0) {
    for (int i = 0; i < CB_STAGES; ++i) mbar_init(mbar_s + 8 * i, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncwarp();
  const uint64_t nwork = max_tiles * (uint64_t)cs.nseg;
  uint32_t it = 0;
  for (uint64_t work = blockIdx.x; work < nwork; work += gridDim.x) {
    const int sg = (int)(work % (uint64_t)cs.nseg);
    const uint64_t off = (work / (uint64_t)cs.nseg) * CB_STAGE;
    if (off >= cs.bytes[sg]) continue;
    const uint32_t len = (uint32_t)min((uint64_t)CB_STAGE, cs.bytes[sg] - off);
    const uint32_t body = len & ~15u;
    const uint8_t* src = cs.src[sg] + off;
    uint8_t* dst = cs.dst[sg] + off;
    if (lane == 0 && body) {
      const uint32_t st = it % CB_STAGES;
      // the store issued from this stage two items ago must have read its bytes
      asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(CB_STAGES - 1) : "memory");
      mbar_expect_tx(mbar_s + 8 * st, body);
      bulk_load(stage_s + st * CB_STAGE, src, body, mbar_s + 8 * st);
      mbar_wait(mbar_s + 8 * st, (it / CB_STAGES) & 1u);
      bulk_store(dst, stage_s + st * CB_STAGE, body);
      ++it;
    }
    for (uint32_t q = body + lane; q < len; q += 32) dst[q] = src[q];
  }
  if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

// -------------------------------------------------------- arrival flags -----

constexpr int FLAG_MAX = 64;
struct FlagPtrs {
  uint32_t* p[FLAG_MAX];
  int n;
};

// Runs after the stores it announces (stream order): every thread publishes one flag with
// a system-scope release, so a peer that observes the value also observes those stores.
__global__ void k_signal_flags(FlagPtrs fp, uint32_t value) {
  const int t = threadIdx.x;
  if (t >= fp.n) return;
  __threadfence_system();
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(fp.p[t]), "r"(value) : "memory");
}

__device__ __forceinline__ uint64_t global_timer_ns() {
  uint64_t t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// One thread per flag spins (system-scope acquire loads) until the flag has reached
// `value`; kernels behind this one on the stream then see the announced data.  A flag
// that does not arrive within timeout_ns sets *error and ends the wait, so a lost peer
// cannot hang the GPU.
__global__ void k_wait_flags(const uint32_t* flags, int n, uint32_t value, uint64_t timeout_ns,
                             uint32_t* error) {
  const int t = threadIdx.x;
  if (t >= n) return;
  const uint64_t t0 = global_timer_ns();
  uint32_t backoff = 32;
  while (true) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(flags + t) : "memory");
    if ((int32_t)(v - value) >= 0) break;
    if (global_timer_ns() - t0 > timeout_ns) {
      atomicExch(error, 1u + (uint32_t)t);
      break;
    }
    __nanosleep(backoff);
    if (backoff < 2048) backoff <<= 1;
  }
}

}  // namespace

extern "C" int xchg_sample_keys(const void* keys, int key_type, int descending, const uint8_t* validity,
                                int64_t validity_bit_offset, const uint64_t* block_start_host,
                                const uint64_t* block_rows_host, int nblocks, uint32_t per_block,
                                uint64_t seed, uint64_t first_block_index, uint64_t* samples_out,
                                uint64_t capacity, uint64_t* count_out_host, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  XCHG_REQUIRE(count_out_host != nullptr, "count_out_host is NULL");
  XCHG_REQUIRE(nblocks >= 0 && (nblocks == 0 || (block_start_host && block_rows_host)), "NULL block arrays");
  uint64_t total = 0;
  for (int b = 0; b < nblocks; ++b) total += std::min<uint64_t>(per_block, block_rows_host[b]);
  *count_out_host = total;
  if (total == 0) return XCHG_OK;
  XCHG_REQUIRE(keys && samples_out, "NULL keys / samples_out");
  if (total > capacity) {
    xchg_set_error("xchg_sample_keys: %llu samples exceed the capacity of %llu; lower per_block",
                   (unsigned long long)total, (unsigned long long)capacity);
    return XCHG_ERR_LIMIT;
  }
  uint64_t pos = 0;
  int b = 0;
  while (b < nblocks) {
    SampleBlocks sb;
    int cnt = 0;
    for (; b < nblocks && cnt < SAMPLE_BLOCKS_PER_LAUNCH; ++b) {
      const uint64_t rows = block_rows_host[b];
      const uint32_t k = (uint32_t)std::min<uint64_t>(per_block, rows);
      if (k == 0) continue;
      XCHG_REQUIRE(rows <= XCHG_MAX_ROWS, "block %d exceeds XCHG_MAX_ROWS", b);
      sb.start[cnt] = block_start_host[b];
      sb.rows[cnt] = rows;
      sb.seed[cnt] = splitmix64(seed ^ splitmix64(first_block_index + (uint64_t)b));
      sb.pos[cnt] = (uint32_t)pos;
      sb.k[cnt] = k;
      pos += k;
      ++cnt;
    }
    if (cnt == 0) continue;
    const bool ok = xchg_dispatch_key_type(key_type, [&](auto tag) {
      using T = typename decltype(tag)::type;
      XCHG_LAUNCH(XK_PARTITION, s,
                  k_sample_keys<T><<<cnt, 128, 0, s>>>((const T*)keys, validity, validity_bit_offset,
                                                       descending, sb, samples_out));
    });
    XCHG_REQUIRE(ok, "unknown key_type %d", key_type);
    XCHG_LAUNCH_CHECK();
  }
  return XCHG_OK;
}

extern "C" int xchg_select_boundaries(const uint64_t* gathered, int nmsgs, uint64_t msg_words,
                                      uint32_t count_word, uint32_t hdr_words, uint32_t num_partitions,
                                      uint64_t* bounds_out, uint64_t* info_out, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  XCHG_REQUIRE(gathered && nmsgs >= 1 && nmsgs <= 64, "gathered NULL or nmsgs outside [1, 64]");
  XCHG_REQUIRE(count_word < hdr_words && hdr_words <= msg_words, "bad message layout");
  XCHG_REQUIRE(num_partitions >= 1, "num_partitions must be >= 1");
  XCHG_REQUIRE(num_partitions == 1 || bounds_out, "NULL bounds_out");
  XCHG_ONCE_PER_DEVICE(XCHG_CUDA_TRY(cudaFuncSetAttribute(
      k_select_boundaries, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(SEL_CAP * 8))));
  XCHG_LAUNCH(XK_PARTITION, s,
              k_select_boundaries<<<1, SEL_THREADS, SEL_CAP * 8, s>>>(gathered, nmsgs, msg_words, count_word,
                                                                     hdr_words, num_partitions, bounds_out,
                                                                     info_out));
  XCHG_LAUNCH_CHECK();
  return XCHG_OK;
}

static int fill_rule(DestRule* rule, const uint32_t* pick_host, const uint8_t* strict_host, int nthr) {
  XCHG_REQUIRE(nthr >= 0 && nthr <= XCHG_MAX_THRESHOLDS, "nthr must be in [0, %d]", XCHG_MAX_THRESHOLDS);
  XCHG_REQUIRE(nthr == 0 || pick_host, "NULL pick_host");
  rule->nthr = nthr;
  for (int i = 0; i < nthr; ++i) {
    rule->pick[i] = pick_host[i];
    rule->strict[i] = strict_host ? strict_host[i] : 0;
  }
  return XCHG_OK;
}

int xchg_internal_fill_rule(DestRule* rule, const uint32_t* pick_host, const uint8_t* strict_host, int nthr) {
  return fill_rule(rule, pick_host, strict_host, nthr);
}

extern "C" int xchg_partition_hist(const void* keys, int key_type, int descending, uint64_t n,
                                   const uint64_t* bounds, const uint32_t* pick_host,
                                   const uint8_t* strict_host, int nthr, uint64_t* counts_out,
                                   void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  XCHG_REQUIRE(counts_out != nullptr, "NULL counts_out");
  DestRule rule;
  int rc = fill_rule(&rule, pick_host, strict_host, nthr);
  if (rc) return rc;
  XCHG_REQUIRE(nthr == 0 || bounds, "NULL bounds");
  XCHG_CUDA_TRY(cudaMemsetAsync(counts_out, 0, (size_t)(nthr + 1) * 8, s));
  if (n == 0) return XCHG_OK;
  XCHG_REQUIRE(keys != nullptr, "NULL keys");
  if (n > XCHG_MAX_ROWS) {
    xchg_set_error("n=%llu exceeds XCHG_MAX_ROWS", (unsigned long long)n);
    return XCHG_ERR_LIMIT;
  }
  const int blocks = (int)std::min<uint64_t>((n + PH_THREADS - 1) / PH_THREADS, (uint64_t)xchg_sm_count() * 8);
  const bool ok = xchg_dispatch_key_type(key_type, [&](auto tag) {
    using T = typename decltype(tag)::type;
    XCHG_LAUNCH(XK_PARTITION, s,
                k_partition_hist<T><<<blocks, PH_THREADS, 0, s>>>((const T*)keys, n, descending, bounds, rule,
                                                                 (unsigned long long*)counts_out));
  });
  XCHG_REQUIRE(ok, "unknown key_type %d", key_type);
  XCHG_LAUNCH_CHECK();
  return XCHG_OK;
}

extern "C" int xchg_copy_segments(int nseg, const void* const* src_host, void* const* dst_host,
                                  const uint64_t* bytes_host, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  if (nseg == 0) return XCHG_OK;
  XCHG_REQUIRE(nseg > 0 && src_host && dst_host && bytes_host, "bad segment arrays");
  for (int base = 0; base < nseg; base += CS_MAX_SEGS) {
    CopySegs cs;
    cs.nseg = 0;
    uint64_t max_tiles = 0;
    for (int i = base; i < nseg && cs.nseg < CS_MAX_SEGS; ++i) {
      if (bytes_host[i] == 0) continue;
      XCHG_REQUIRE(src_host[i] && dst_host[i], "NULL pointer in segment %d", i);
      cs.src[cs.nseg] = (const uint8_t*)src_host[i];
      cs.dst[cs.nseg] = (uint8_t*)dst_host[i];
      cs.bytes[cs.nseg] = bytes_host[i];
      max_tiles = std::max<uint64_t>(max_tiles, (bytes_host[i] + CS_TILE - 1) / CS_TILE);
      ++cs.nseg;
    }
    if (cs.nseg == 0) continue;
    // XCHG_COPY_BULK=1: the TMA variant (16-byte aligned segments only)
    const char* bulk_env = getenv("XCHG_COPY_BULK");  // read per call: a profiling switch
    const int bulk_mode = bulk_env ? atoi(bulk_env) : 0;
    bool aligned = true;
    for (int i = 0; i < cs.nseg; ++i) aligned = aligned && ((((uintptr_t)cs.src[i] | (uintptr_t)cs.dst[i]) & 15) == 0);
    if (bulk_mode && aligned) {
      constexpr int smem = CB_STAGE * CB_STAGES;
      XCHG_ONCE_PER_DEVICE(XCHG_CUDA_TRY(cudaFuncSetAttribute(k_copy_segments_bulk,
                                                              cudaFuncAttributeMaxDynamicSharedMemorySize, smem)));
      const uint64_t tiles = ((max_tiles * CS_TILE + CB_STAGE - 1) / CB_STAGE);
      const int grid = (int)std::min<uint64_t>(tiles * cs.nseg, (uint64_t)xchg_sm_count() * 3);
      XCHG_LAUNCH(XK_MISC, s, k_copy_segments_bulk<<<grid, 32, smem, s>>>(cs, tiles));
    } else {
      const int grid = (int)std::min<uint64_t>(max_tiles * cs.nseg, (uint64_t)xchg_sm_count() * 8);
      XCHG_LAUNCH(XK_MISC, s, k_copy_segments<<<grid, 256, 0, s>>>(cs, max_tiles));
    }
    XCHG_LAUNCH_CHECK();
  }
  return XCHG_OK;
}

extern "C" int xchg_signal_flags(uint32_t* const* flag_ptrs_host, int n, uint32_t value, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  if (n == 0) return XCHG_OK;
  XCHG_REQUIRE(n > 0 && flag_ptrs_host, "bad flag array");
  for (int base = 0; base < n; base += FLAG_MAX) {
    FlagPtrs fp;
    fp.n = std::min(FLAG_MAX, n - base);
    for (int i = 0; i < fp.n; ++i) {
      XCHG_REQUIRE(flag_ptrs_host[base + i] != nullptr, "NULL flag pointer %d", base + i);
      fp.p[i] = flag_ptrs_host[base + i];
    }
    XCHG_LAUNCH(XK_MISC, s, k_signal_flags<<<1, FLAG_MAX, 0, s>>>(fp, value));
    XCHG_LAUNCH_CHECK();
  }
  return XCHG_OK;
}

extern "C" int xchg_wait_flags(const uint32_t* flags, int n, uint32_t value, uint64_t timeout_ns,
                               uint32_t* error_flag, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  if (n == 0) return XCHG_OK;
  XCHG_REQUIRE(n > 0 && n <= 1024 && flags && error_flag, "bad arguments");
  XCHG_LAUNCH(XK_MISC, s, k_wait_flags<<<1, ((n + 31) / 32) * 32, 0, s>>>(flags, n, value, timeout_ns, error_flag));
  XCHG_LAUNCH_CHECK();
  return XCHG_OK;
}
