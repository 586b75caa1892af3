// K7: seeded random permutation for Dataset.random_shuffle, for sm_100a.
//
// Replaces numpy.random.RandomState(seed).permutation(n) as the reference calls it
// from python/ray/data/_internal/arrow_block.py:225-231 (ArrowBlockAccessor.
// random_shuffle; map side shuffle_task_spec.py:81-84, reduce side :120-123).  The
// reference's generator is MT19937 driving a sequential Fisher-Yates; its VALUES are
// not pinned by any reference test (only same-seed-equal,
// different-seed-different and multiset preservation are), so the device uses a
// counter-based construction instead that needs no sequential state and no sort:
//
//   a 4-round balanced Feistel network over 2^b (b = even bit width covering n) whose
//   round function is Philox-2x32-10 keyed by the seed, restricted to [0, n) by cycle
//   walking (apply again while the value is >= n; the domain is < 4n, so < 4
//   applications on average; lanes that finish early take the next index, see the kernel).  It is a bijection of [0, n) for every seed, both
//   directions are computable per element (forward = output position of a source row,
//   inverse = source row of an output position), and any sub-range can be evaluated
//   independently -- which is what lets every GPU compute the destination of its own
//   rows of ONE global permutation without communication.
//
// Pure ALU work (~100 integer multiplies per row), 4 bytes written per row: the
// permutation itself is far from any roofline; the payload gather that consumes it
// (K5) is the HBM-bound part.
#include <algorithm>

#include "common.cuh"
#include "feistel.cuh"

namespace {

using namespace xchg_feistel;

// Cycle walking takes a geometrically distributed number of applications per element (the
// domain is up to 4n), so with one element per lane a warp runs as long as its unluckiest lane:
// ~8 applications where the average is < 3 (1.36 ms for 25 M rows).  Here a warp owns a
// contiguous slice of the indices and a lane that lands inside [0, n) takes the next unassigned
// index of the slice at once: the warp's work is the SUM of the walks, not 32 x the longest.
// Same function, same values.
constexpr uint32_t SH_WARP_SLICE = 2048;

__global__ void __launch_bounds__(256)
k_shuffle_positions(Feistel f, uint64_t first, uint64_t count, int inverse, uint32_t* __restrict__ out) {
  const unsigned lane = threadIdx.x & 31;
  const unsigned lt_mask = (1u << lane) - 1u;
  const uint64_t warps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
  const uint64_t nslices = (count + SH_WARP_SLICE - 1) / SH_WARP_SLICE;
  const bool inv = inverse != 0;
  for (uint64_t sl = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; sl < nslices; sl += warps) {
    const uint64_t w0 = sl * SH_WARP_SLICE, w1 = min(w0 + (uint64_t)SH_WARP_SLICE, count);
    uint64_t idx = w0 + lane;      // the element this lane is walking
    uint64_t pending = w0 + 32;    // first index nobody has taken yet (warp-uniform)
    bool have = idx < w1;
    uint64_t x = first + idx;
    while (__any_sync(0xffffffffu, have)) {
      if (have) x = feistel_once(f, x, inv);
      const bool done = have && x < f.n;
      if (done) out[idx] = (uint32_t)x;
      const unsigned dm = __ballot_sync(0xffffffffu, done);
      if (done) {
        idx = pending + (uint64_t)__popc(dm & lt_mask);
        have = idx < w1;
        x = first + idx;
      }
      pending += (uint64_t)__popc(dm);
    }
  }
}

__global__ void __launch_bounds__(256)
k_invert_permutation(const uint32_t* __restrict__ pos, uint64_t n, uint64_t base,
                     uint32_t* __restrict__ inv) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    inv[(uint64_t)pos[i] - base] = (uint32_t)i;
}

// destination partition of a position under an even split of n_total into parts:
// partition p owns [p*n/parts, (p+1)*n/parts) (floor), i.e. p = the largest p with
// floor(p*n/parts) <= pos
__global__ void __launch_bounds__(256)
k_even_split_dest(const uint32_t* __restrict__ pos, uint64_t n, uint64_t n_total, uint32_t parts,
                  uint32_t* __restrict__ dest, unsigned long long* counts) {
  extern __shared__ uint32_t s_cnt[];
  for (uint32_t j = threadIdx.x; j < parts; j += blockDim.x) s_cnt[j] = 0;
  __syncthreads();
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const uint64_t p = pos[i];
    // first guess, then fix up the floor effects
    uint64_t d = (p * parts + parts - 1) / n_total;
    if (d >= parts) d = parts - 1;
    while (d > 0 && (d * n_total) / parts > p) --d;
    while (d + 1 < parts && ((d + 1) * n_total) / parts <= p) ++d;
    dest[i] = (uint32_t)d;
    atomicAdd(&s_cnt[d], 1u);
  }
  __syncthreads();
  for (uint32_t j = threadIdx.x; j < parts; j += blockDim.x)
    if (s_cnt[j]) atomicAdd(&counts[j], (unsigned long long)s_cnt[j]);
}

int blocks_for(uint64_t n) {
  return (int)std::max<uint64_t>(1, std::min<uint64_t>((n + 255) / 256, (uint64_t)xchg_sm_count() * 8));
}

}  // namespace

extern "C" int xchg_shuffle_positions(uint64_t seed, uint64_t n_total, uint64_t first, uint64_t count,
                                      int inverse, uint32_t* out, void* stream) {
  if (count == 0) return XCHG_OK;
  XCHG_REQUIRE(out != nullptr, "NULL out");
  XCHG_REQUIRE(n_total >= 1 && first + count <= n_total, "range [first, first+count) exceeds n_total");
  if (n_total > 0xFFFFFFFFull) {
    xchg_set_error("xchg_shuffle_positions: n_total=%llu exceeds 2^32-1 (positions are uint32)",
                   (unsigned long long)n_total);
    return XCHG_ERR_LIMIT;
  }
  Feistel f;
  make_feistel(seed, n_total, &f);
  XCHG_LAUNCH(XK_SHUFFLE, (cudaStream_t)stream,
              k_shuffle_positions<<<blocks_for(count), 256, 0, (cudaStream_t)stream>>>(f, first, count,
                                                                                     inverse, out));
  XCHG_LAUNCH_CHECK();
  return XCHG_OK;
}

extern "C" int xchg_invert_permutation(const uint32_t* pos, uint64_t n, uint64_t base, uint32_t* inv_out,
                                       void* stream) {
  if (n == 0) return XCHG_OK;
  XCHG_REQUIRE(pos && inv_out, "NULL pos / inv_out");
  XCHG_LAUNCH(XK_SHUFFLE, (cudaStream_t)stream,
              k_invert_permutation<<<blocks_for(n), 256, 0, (cudaStream_t)stream>>>(pos, n, base, inv_out));
  XCHG_LAUNCH_CHECK();
  return XCHG_OK;
}

extern "C" int xchg_even_split_dest(const uint32_t* pos, uint64_t n, uint64_t n_total, uint32_t parts,
                                    uint32_t* dest_out, uint64_t* counts_out, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  XCHG_REQUIRE(parts >= 1 && parts <= 8192, "parts must be in [1, 8192]");
  XCHG_REQUIRE(counts_out != nullptr, "NULL counts_out");
  XCHG_CUDA_TRY(cudaMemsetAsync(counts_out, 0, (size_t)parts * 8, s));
  if (n == 0) return XCHG_OK;
  XCHG_REQUIRE(pos && dest_out, "NULL pos / dest_out");
  XCHG_REQUIRE(n_total >= 1, "n_total must be >= 1");
  XCHG_LAUNCH(XK_SHUFFLE, s,
              k_even_split_dest<<<blocks_for(n), 256, (size_t)parts * 4, s>>>(
                  pos, n, n_total, parts, dest_out, (unsigned long long*)counts_out));
  XCHG_LAUNCH_CHECK();
  return XCHG_OK;
}
