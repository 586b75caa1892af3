// K5, uniform-width rows (GraySort 90-byte payloads, 1 KiB tensor rows): the
// payload hot loop of Dataset.sort / random_shuffle.  Replaces Table.take on the
// payload column (python/ray/data/_internal/arrow_ops/transform_pyarrow.py:45-71).
//
// out row i = in row idx[i], every row `w` bytes.  Per CTA and tile of T rows:
//   phase 0  idx -> source byte offsets in shared memory (coalesced);
//   phase 1  every 16-byte window of every source row -> stage via cp.async.cg, the
//            row keeping its source alignment inside a fixed slot (no registers
//            held while the random reads are in flight: T*ceil16(w+15) bytes of
//            loads outstanding per CTA, 3 CTAs per SM);
//   phase 2  FLAT mapping: thread q produces the q-th 4-byte-aligned word of the
//            tile's contiguous output range; row = word_byte / w by multiply-high,
//            two aligned LDS + one funnel shift re-align source to destination;
//            words that straddle two rows are assembled bytewise.  Stores are dense
//            128-byte warp writes.
// The first round-1 version used one warp per row for phase 2 (105 warp
// instructions per 90-byte row, issue-bound at 1.3 TB/s, profiles/r01_summary.md);
// this mapping needs ~10.
// Algorithmic bytes: 2*w + 4 per row.  Rows whose width and both base pointers are
// multiples of 16 take the vector variant (16-byte words, no re-alignment).
#include <algorithm>
#include <cstdlib>

#include "common.cuh"

namespace {

constexpr int GU_THREADS = 256;
constexpr int GU_STAGE_DEFAULT_KB = 56;
constexpr int GU_STAGE_MAX = 56 * 1024;
constexpr int GU_MAX_ROWS = 1024;
// int64 src[rows_per_tile] | uint32 a0[rows_per_tile + 1] | stage | pad: the row tables are sized
// by the tile's rows, so a small stage makes a small CTA (the peer-store gathers run with a
// few KB per CTA beside kernels that need the SM's shared memory)
__host__ __device__ constexpr size_t gu_meta(uint32_t rows_per_tile) { return (((size_t)rows_per_tile * 12 + 4) + 15) & ~(size_t)15; }
constexpr size_t gu_smem(int stage_bytes, uint32_t rows_per_tile) {
  return gu_meta(rows_per_tile) + (size_t)stage_bytes + 32;
}

// Bytes of row staging per CTA.  Smaller stages let more CTAs share an SM, so that more
// of them are in their (latency-bound) load phase while others write out: XCHG_GATHER_STAGE_KB
// (16..56) overrides the default.
static int gu_stage_bytes() {
  const int over = xchg_gather_stage_kb();  // per-launch override (xchg_set_gather_stage_kb)
  if (over >= 4 && over <= 56) return over * 1024;
  static int v = 0;
  if (v == 0) {
    const char* e = getenv("XCHG_GATHER_STAGE_KB");
    int kb = e ? atoi(e) : 0;
    if (kb < 4 || kb > 56) kb = GU_STAGE_DEFAULT_KB;
    v = kb * 1024;
  }
  return v;
}

// .L2::64B: a miss fetches the 64 bytes around the access from HBM instead of the whole
// 128-byte line (measured on B200, profiles/r01_fetch_granularity.md: 127 B of DRAM
// reads per random 32-byte access by default, 64 B with this qualifier;
// cudaLimitMaxL2FetchGranularity has no effect).  A 90-byte row then costs ~153 B of
// reads instead of ~219 B.
__device__ __forceinline__ void cp_async16_cg(uint32_t smem_addr, const void* g) {
  asm volatile("cp.async.cg.shared.global.L2::64B [%0], [%1], 16;" ::"r"(smem_addr), "l"(g) : "memory");
}
__device__ __forceinline__ void cp_async_drain() {
  asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory");
}

// x / d for a precomputed magic = ceil(2^32 / d); magic == 0 encodes d == 1
__device__ __forceinline__ uint32_t udiv_magic(uint32_t x, uint32_t magic) {
  return magic ? __umulhi(x, magic) : x;
}

// Several output segments in ONE launch (the multi-GPU exchange: segment d = the rows this
// rank sends to destination d, out[d] = a pointer into that peer's arena).  The tiles of
// the segments are interleaved (tile r of every segment, then tile r+1, ...), so a rank
// stores to all its peers at once and every GPU's NVLink ingress is shared evenly by all
// senders for the whole kernel, whatever the ranks' relative progress -- one launch per
// destination leaves that to chance.
constexpr int GU_MAX_SEGS = 16;
struct GatherSegs {
  uint64_t row0[GU_MAX_SEGS];  // first entry of the segment in idx
  uint64_t n[GU_MAX_SEGS];     // rows
  uint8_t* out[GU_MAX_SEGS];
  int nseg;
};

template <bool VEC16, bool MULTI>
__global__ void __launch_bounds__(GU_THREADS)
k_gather_rows_uniform(const uint8_t* __restrict__ in, uint32_t w, const uint32_t* __restrict__ idx_,
                      uint64_t n_out_, uint8_t* __restrict__ out_, uint32_t rows_per_tile,
                      uint32_t slot, uint32_t magic_w, uint32_t magic_cpr, GatherSegs segs,
                      uint64_t max_tiles) {
  extern __shared__ __align__(16) uint8_t smem[];
  int64_t* m_src = reinterpret_cast<int64_t*>(smem);
  uint32_t* m_a0 = reinterpret_cast<uint32_t*>(m_src + rows_per_tile);
  uint8_t* stage = smem + gu_meta(rows_per_tile);
  const uint32_t stage_s = (uint32_t)__cvta_generic_to_shared(stage);
  const int tid = threadIdx.x;
  const uint32_t cpr = slot >> 4;  // 16-byte chunks per slot
  const uint64_t nwork = MULTI ? max_tiles * (uint64_t)segs.nseg
                               : (n_out_ + rows_per_tile - 1) / rows_per_tile;
  for (uint64_t work = blockIdx.x; work < nwork; work += gridDim.x) {
    const uint32_t* idx = idx_;
    uint64_t n_out = n_out_;
    uint8_t* out = out_;
    uint64_t tile = work;
    if constexpr (MULTI) {
      const int sg = (int)(work % (uint64_t)segs.nseg);
      tile = work / (uint64_t)segs.nseg;
      idx = idx_ + segs.row0[sg];
      n_out = segs.n[sg];
      out = segs.out[sg];
      if (tile * rows_per_tile >= n_out) continue;  // uniform per CTA: no barrier is skipped
    }
    const uint64_t r0 = tile * rows_per_tile;
    const uint32_t T = (uint32_t)min((uint64_t)rows_per_tile, n_out - r0);

    // ---- phase 0 ----
    for (uint32_t j = tid; j < T; j += GU_THREADS) {
      const int64_t s = (int64_t)idx[r0 + j] * w;
      m_src[j] = s;
      m_a0[j] = j * slot + (VEC16 ? 0u : (uint32_t)(((uintptr_t)in + s) & 15));
    }
    __syncthreads();

    // ---- phase 1: global -> stage ----
    const uint32_t nchunks = T * cpr;
    for (uint32_t c = tid; c < nchunks; c += GU_THREADS) {
      const uint32_t j = udiv_magic(c, magic_cpr);
      const uint32_t k = c - j * cpr;
      const uint8_t* g = in + m_src[j];
      const uint32_t mis = VEC16 ? 0u : (uint32_t)((uintptr_t)g & 15);
      if (16 * k < mis + w) cp_async16_cg(stage_s + j * slot + 16 * k, g - mis + 16 * (size_t)k);
    }
    cp_async_drain();
    __syncthreads();

    // ---- phase 2: stage -> global ----
    uint8_t* o = out + r0 * w;  // the tile's output range is contiguous
    const uint32_t nbytes = T * w;
    if constexpr (VEC16) {
      // slot == w, everything 16-byte aligned: the stage IS the output layout
      const uint4* s4 = reinterpret_cast<const uint4*>(stage);
      uint4* o4 = reinterpret_cast<uint4*>(o);
      for (uint32_t q = tid; q < (nbytes >> 4); q += GU_THREADS) o4[q] = s4[q];
    } else if (w >= 16) {
      // 16 output bytes per thread.  Unit at tile byte t lies in row j = t / w from its
      // byte r; if the row ends inside the unit (k = w - r < 16) the rest comes from the
      // start of row j+1.  Both 16-byte windows are read unaligned from the stage (five
      // aligned words + four funnel shifts each; the slots' slack keeps the over-reads in
      // bounds) and merged bytewise with masks: no divergence, one 16-byte store.
      const uint32_t head = min(nbytes, (uint32_t)((16 - ((uintptr_t)o & 15)) & 15));
      const uint32_t nunits = (nbytes - head) >> 4;
      const uint32_t tail0 = head + (nunits << 4);
      uint4* ov = reinterpret_cast<uint4*>(o + head);
      const uint32_t* sw = reinterpret_cast<const uint32_t*>(stage);
      for (uint32_t q = tid; q < nunits; q += GU_THREADS) {
        const uint32_t t = head + (q << 4);
        const uint32_t j = udiv_magic(t, magic_w);
        const uint32_t r = t - j * w;
        const uint32_t k = w - r;  // bytes of row j available from r (>= 1)
        const uint32_t a = m_a0[j] + r;
        const uint32_t* pa = sw + (a >> 2);
        const uint32_t sa = (a & 3) * 8;
        const uint32_t a0w = pa[0], a1w = pa[1], a2w = pa[2], a3w = pa[3], a4w = pa[4];
        uint4 v;
        v.x = __funnelshift_r(a0w, a1w, sa);
        v.y = __funnelshift_r(a1w, a2w, sa);
        v.z = __funnelshift_r(a2w, a3w, sa);
        v.w = __funnelshift_r(a3w, a4w, sa);
        if (k < 16) {
          // window that has row j+1's first bytes at unit offsets k..15
          const uint32_t b = m_a0[j + 1] - k;
          const uint32_t* pb = sw + (b >> 2);
          const uint32_t sb = (b & 3) * 8;
          const uint32_t b0w = pb[0], b1w = pb[1], b2w = pb[2], b3w = pb[3], b4w = pb[4];
          uint32_t u[4] = {__funnelshift_r(b0w, b1w, sb), __funnelshift_r(b1w, b2w, sb),
                           __funnelshift_r(b2w, b3w, sb), __funnelshift_r(b3w, b4w, sb)};
          uint32_t x[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            // bytes of word i that still belong to row j: clamp(k - 4i, 0, 4)
            const int keep = min(4, max(0, (int)k - 4 * i));
            const uint32_t mask = keep >= 4 ? 0xFFFFFFFFu : ((1u << (8 * keep)) - 1u);
            x[i] = (x[i] & mask) | (u[i] & ~mask);
          }
          v = make_uint4(x[0], x[1], x[2], x[3]);
        }
        ov[q] = v;
      }
      // head and tail bytes of the tile (at most 15 each)
      if ((uint32_t)tid < head + (nbytes - tail0)) {
        const uint32_t t = (uint32_t)tid < head ? (uint32_t)tid : tail0 + ((uint32_t)tid - head);
        const uint32_t j = udiv_magic(t, magic_w);
        o[t] = stage[m_a0[j] + (t - j * w)];
      }
    } else {
      const uint32_t head = min(nbytes, (uint32_t)((4 - ((uintptr_t)o & 3)) & 3));
      const uint32_t nwords = (nbytes - head) >> 2;
      const uint32_t tail0 = head + (nwords << 2);
      uint32_t* ow = reinterpret_cast<uint32_t*>(o + head);
      const uint32_t* sw = reinterpret_cast<const uint32_t*>(stage);
      for (uint32_t q = tid; q < nwords; q += GU_THREADS) {
        const uint32_t t = head + (q << 2);  // tile-relative byte of this word
        const uint32_t j = udiv_magic(t, magic_w);
        const uint32_t r = t - j * w;
        uint32_t v;
        if (r + 4 <= w) {
          const uint32_t a = m_a0[j] + r;
          v = __funnelshift_r(sw[a >> 2], sw[(a >> 2) + 1], (a & 3) * 8);
        } else {
          v = 0;
#pragma unroll
          for (uint32_t b = 0; b < 4; ++b) {
            uint32_t jb = j, rb = r + b;
            while (rb >= w) {  // w < 4 can put several rows into one word
              rb -= w;
              ++jb;
            }
            v |= (uint32_t)stage[m_a0[jb] + rb] << (8 * b);
          }
        }
        ow[q] = v;
      }
      // head and tail bytes of the tile (at most 3 each)
      if ((uint32_t)tid < head + (nbytes - tail0)) {
        const uint32_t t = (uint32_t)tid < head ? (uint32_t)tid : tail0 + ((uint32_t)tid - head);
        const uint32_t j = udiv_magic(t, magic_w);
        o[t] = stage[m_a0[j] + (t - j * w)];
      }
    }
    __syncthreads();
  }
}

// ---- TMA variant for rows whose width and bases are multiples of 16 bytes ----
// (1 KiB tensor rows of random_shuffle).  One warp per CTA, no thread touches the data:
// the lanes issue one cp.async.bulk per row (global -> shared, completion counted in bytes
// on an mbarrier), lane 0 then stores the tile -- contiguous in the output -- with ONE bulk
// store; two stages, so a tile's store drains while the next tile's rows arrive.
constexpr int GB_STAGES = 2;
constexpr uint32_t GB_STAGE_BYTES = 40 * 1024;

__global__ void __launch_bounds__(32)
k_gather_rows_bulk(const uint8_t* __restrict__ in, uint32_t w, const uint32_t* __restrict__ idx, uint64_t n_out,
                   uint8_t* __restrict__ out, uint32_t rows_per_tile) {
  extern __shared__ __align__(128) uint8_t gb_smem[];
  __shared__ __align__(8) uint64_t mbar_mem[GB_STAGES];
  const uint32_t stage_s = (uint32_t)__cvta_generic_to_shared(gb_smem);
  const uint32_t mbar_s = (uint32_t)__cvta_generic_to_shared(mbar_mem);
  const int lane = threadIdx.x;
  if (lane == 0) {
    for (int i = 0; i < GB_STAGES; ++i)
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(mbar_s + 8 * i) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncwarp();
  const uint64_t ntiles = (n_out + rows_per_tile - 1) / rows_per_tile;
  uint32_t it = 0;
  for (uint64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
    const uint64_t r0 = tile * rows_per_tile;
    const uint32_t T = (uint32_t)min((uint64_t)rows_per_tile, n_out - r0);
    const uint32_t st = it % GB_STAGES;
    const uint32_t sbase = stage_s + st * GB_STAGE_BYTES;
    const uint32_t mb = mbar_s + 8 * st;
    if (lane == 0) {
      // the bulk store issued from this stage two tiles ago must have read its bytes
      asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(GB_STAGES - 1) : "memory");
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mb), "r"(T * w) : "memory");
    }
    __syncwarp();
    for (uint32_t j = lane; j < T; j += 32) {
      const uint8_t* g = in + (uint64_t)idx[r0 + j] * w;
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                   ::"r"(sbase + j * w), "l"(g), "r"(w), "r"(mb) : "memory");
    }
    if (lane == 0) {
      const uint32_t parity = (it / GB_STAGES) & 1u;
      asm volatile(
          "{\n\t.reg .pred p;\n\t"
          "WAITB_%=:\n\t"
          "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
          "@!p bra WAITB_%=;\n\t}"
          ::"r"(mb), "r"(parity) : "memory");
      asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                   ::"l"(out + r0 * w), "r"(sbase), "r"(T * w) : "memory");
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    }
    __syncwarp();
  }
  if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

}  // namespace

static GatherSegs no_segs() {
  GatherSegs g;
  g.nseg = 0;
  return g;
}

static int gather_uniform_launch(const uint8_t* in, uint32_t w, const uint32_t* idx, uint64_t n_out,
                                 uint8_t* out, const GatherSegs* multi, cudaStream_t s) {
  XCHG_ONCE_PER_DEVICE({
    XCHG_CUDA_TRY(cudaFuncSetAttribute(k_gather_rows_uniform<false, false>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gu_smem(GU_STAGE_MAX, GU_MAX_ROWS)));
    XCHG_CUDA_TRY(cudaFuncSetAttribute(k_gather_rows_uniform<true, false>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gu_smem(GU_STAGE_MAX, GU_MAX_ROWS)));
    XCHG_CUDA_TRY(cudaFuncSetAttribute(k_gather_rows_uniform<false, true>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gu_smem(GU_STAGE_MAX, GU_MAX_ROWS)));
    XCHG_CUDA_TRY(cudaFuncSetAttribute(k_gather_rows_uniform<true, true>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gu_smem(GU_STAGE_MAX, GU_MAX_ROWS)));
  });
  uintptr_t align_bits = (uintptr_t)in | (uintptr_t)out;
  if (multi)
    for (int d = 0; d < multi->nseg; ++d) align_bits |= (uintptr_t)multi->out[d];
  const bool vec16 = (w % 16 == 0) && (align_bits % 16 == 0);
  // XCHG_GATHER_BULK=1: aligned rows through the TMA unit (A/B switch, read per call)
  const char* bulk_env = getenv("XCHG_GATHER_BULK");
  if (bulk_env && atoi(bulk_env) == 1 && vec16 && !multi && w <= GB_STAGE_BYTES) {
    constexpr int smem = GB_STAGES * GB_STAGE_BYTES;
    XCHG_ONCE_PER_DEVICE(XCHG_CUDA_TRY(cudaFuncSetAttribute(k_gather_rows_bulk,
                                                            cudaFuncAttributeMaxDynamicSharedMemorySize, smem)));
    const uint32_t rpt = GB_STAGE_BYTES / w;
    const uint64_t ntiles = (n_out + rpt - 1) / rpt;
    const int cap = xchg_gather_ctas_per_sm();
    const int grid = (int)std::min<uint64_t>(ntiles, (uint64_t)xchg_sm_count() * (cap ? cap : 2 * 8));
    XCHG_LAUNCH(XK_GATHER_ROWS, s, k_gather_rows_bulk<<<grid, 32, smem, s>>>(in, w, idx, n_out, out, rpt));
    XCHG_LAUNCH_CHECK();
    return XCHG_OK;
  }
  const uint32_t slot = vec16 ? w : ((w + 15u + 15u) & ~15u);
  int stage_bytes = gu_stage_bytes();
  if (slot > (uint32_t)stage_bytes) stage_bytes = GU_STAGE_MAX;  // wide rows: the largest stage
  const uint32_t rpt = std::min<uint32_t>(GU_MAX_ROWS, (uint32_t)stage_bytes / slot);
  if (rpt == 0) return 1;  // row wider than the stage: caller takes the generic path
  const size_t smem_bytes = gu_smem(stage_bytes, rpt);
  // CTAs of this shape an SM holds (227 KB of shared memory, 1 KB reserved per CTA)
  const int resident = std::max(1, std::min(8, (int)((227 * 1024) / (smem_bytes + 1024))));
  const uint32_t cpr = slot >> 4;
  // exact for every dividend below 2^32 / divisor, which bounds all uses (tile bytes, chunks)
  const uint32_t magic_w = w == 1 ? 0u : (uint32_t)((0x100000000ull + w - 1) / w);
  const uint32_t magic_cpr = cpr == 1 ? 0u : (uint32_t)((0x100000000ull + cpr - 1) / cpr);
  uint64_t ntiles = (n_out + rpt - 1) / rpt;
  uint64_t max_tiles = 0;
  if (multi) {
    for (int d = 0; d < multi->nseg; ++d) max_tiles = std::max<uint64_t>(max_tiles, (multi->n[d] + rpt - 1) / rpt);
    ntiles = max_tiles * (uint64_t)multi->nseg;
    if (ntiles == 0) return XCHG_OK;
  }
  // persistent CTAs (3 resident per SM, 8 waves of tiles each by default); a cap leaves
  // SMs' shared memory to kernels running beside a link-bound peer-store gather
  const int cap = xchg_gather_ctas_per_sm();
  const int grid = (int)std::min<uint64_t>(ntiles, (uint64_t)xchg_sm_count() * (cap ? cap : resident * 8));
  const GatherSegs sg = multi ? *multi : no_segs();
#define XCHG_GU_LAUNCH(VEC, MULTI)                                                                   \
  XCHG_LAUNCH(XK_GATHER_ROWS, s,                                                                     \
              k_gather_rows_uniform<VEC, MULTI><<<grid, GU_THREADS, smem_bytes, s>>>(                    \
                  in, w, idx, n_out, out, rpt, slot, magic_w, magic_cpr, sg, max_tiles))
  if (multi) {
    if (vec16) XCHG_GU_LAUNCH(true, true);
    else XCHG_GU_LAUNCH(false, true);
  } else {
    if (vec16) XCHG_GU_LAUNCH(true, false);
    else XCHG_GU_LAUNCH(false, false);
  }
#undef XCHG_GU_LAUNCH
  XCHG_LAUNCH_CHECK();
  return XCHG_OK;
}

int xchg_internal_gather_rows_uniform(const uint8_t* in, uint32_t w, const uint32_t* idx,
                                      uint64_t n_out, uint8_t* out, cudaStream_t s) {
  return gather_uniform_launch(in, w, idx, n_out, out, nullptr, s);
}

extern "C" int xchg_gather_rows_multi(const void* in, uint32_t row_bytes, const uint32_t* idx, int nseg,
                                      const uint64_t* seg_row0_host, const uint64_t* seg_rows_host,
                                      void* const* seg_out_host, void* stream) {
  if (nseg == 0) return XCHG_OK;
  XCHG_REQUIRE(nseg > 0 && nseg <= GU_MAX_SEGS, "nseg must be in [1, %d]", GU_MAX_SEGS);
  XCHG_REQUIRE(in && idx && seg_row0_host && seg_rows_host && seg_out_host, "NULL argument");
  XCHG_REQUIRE(row_bytes > 0, "row_bytes must be positive");
  GatherSegs g;
  g.nseg = nseg;
  for (int d = 0; d < nseg; ++d) {
    XCHG_REQUIRE(seg_rows_host[d] == 0 || seg_out_host[d], "NULL output of segment %d", d);
    XCHG_REQUIRE(seg_rows_host[d] <= XCHG_MAX_ROWS, "segment %d exceeds XCHG_MAX_ROWS", d);
    g.row0[d] = seg_row0_host[d];
    g.n[d] = seg_rows_host[d];
    g.out[d] = (uint8_t*)seg_out_host[d];
  }
  const int rc = gather_uniform_launch((const uint8_t*)in, row_bytes, idx, 0, nullptr, &g, (cudaStream_t)stream);
  if (rc == 1) {
    xchg_set_error("rows of %u bytes exceed the staging tile of xchg_gather_rows_multi; gather the "
                   "segments one by one with xchg_gather_rows", row_bytes);
    return XCHG_ERR_LIMIT;
  }
  return rc;
}
