// K6: stable k-way merge of sorted runs of (u64 radix key, u32 row id), for sm_100a.
//
// Replaces the reduce side of the reference's sort, which concatenates the M sorted
// partitions a reducer receives and sorts them again from scratch
// (python/ray/data/_internal/arrow_block.py:529-541 -> _internal/arrow_ops/
// transform_pyarrow.py:540-547: concat + pac.sort_indices + take).  Its result is the
// stable sort of the concatenation, i.e. ties keep (run index, position in run) order --
// exactly what a stable merge of the runs produces, without looking at a key more than
// ceil(log2 K) times.
//
// Rounds of pairwise merges (runs 0+1, 2+3, ...; an odd run out is carried over), each
// round two launches over all pairs.  k_merge_splits: one thread per 4096-element output
// tile does the merge-path search in global memory (~log2(n) dependent loads) that bounds
// the tile's slices of the left and right run; doing this inside the merge CTAs put ~20 us
// of load latency in front of every tile (measured: 1.5 TB/s of algorithmic bytes).
// k_merge_pairs: a CTA stages both slices of its tile in shared memory, every thread finds
// its own diagonal there, merges 16 elements sequentially (left run wins ties: stable; the
// two head keys live in registers, one shared-memory load per step) and the tile leaves
// through shared memory as coalesced stores.
// Algorithmic bytes: 24 B per element and round (12 read + 12 written).
#include <algorithm>

#include "common.cuh"

namespace {

constexpr int MG_THREADS = 256;
constexpr int MG_ITEMS = 16;
constexpr int MG_TILE = MG_THREADS * MG_ITEMS;
constexpr int MG_MAX_PAIRS = 64;
// Staging slot of tile element e.  A thread owns 16 consecutive elements; unpadded, the 32
// lanes of a warp would hit one bank (stride 16 x 8 B = 128 B): ncu showed 71 % of the
// shared-memory wavefronts as bank conflicts and the L1 pipe at 97 %.  One spare slot per 16
// makes the stride 17: conflict-free for the register -> staging writes and for runs that
// come mostly from one side.
__host__ __device__ constexpr uint32_t mg_slot(uint32_t e) { return e + (e >> 4); }
constexpr uint32_t MG_SLOTS = mg_slot(MG_TILE);
constexpr size_t MG_SMEM = (size_t)MG_SLOTS * 12;

struct MergePairs {
  // pair j merges A = [a0, a0 + na) and B = [a0 + na, a0 + na + nb) of the input into the
  // same positions [a0, a0 + na + nb) of the output (nb == 0: carried over)
  uint64_t a0[MG_MAX_PAIRS];
  uint32_t na[MG_MAX_PAIRS], nb[MG_MAX_PAIRS];
  uint32_t tile0[MG_MAX_PAIRS + 1];  // first tile of every pair
  int npairs;
};

// number of elements taken from A among the first `diag` outputs of merge(A, B); ties -> A
__device__ __forceinline__ uint32_t merge_path(const uint64_t* __restrict__ a, uint32_t na,
                                               const uint64_t* __restrict__ b, uint32_t nb, uint32_t diag) {
  uint32_t lo = diag > nb ? diag - nb : 0, hi = min(diag, na);
  while (lo < hi) {
    const uint32_t mid = (lo + hi) >> 1;
    // A[mid] goes before B[diag - 1 - mid] iff A[mid] <= B[...]
    if (a[mid] <= b[diag - 1 - mid]) lo = mid + 1;
    else hi = mid;
  }
  return lo;
}

// the same search on the two slices staged in shared memory (A at elements [0, la), B behind it)
__device__ __forceinline__ uint32_t merge_path_staged(const uint64_t* sk, uint32_t la, uint32_t lb, uint32_t diag) {
  uint32_t lo = diag > lb ? diag - lb : 0, hi = min(diag, la);
  while (lo < hi) {
    const uint32_t mid = (lo + hi) >> 1;
    if (sk[mg_slot(mid)] <= sk[mg_slot(la + diag - 1 - mid)]) lo = mid + 1;
    else hi = mid;
  }
  return lo;
}

__device__ __forceinline__ int pair_of_tile(const MergePairs& mp, uint32_t tile) {
  int j = 0;
  while (j + 1 < mp.npairs && mp.tile0[j + 1] <= tile) ++j;  // <= 64 pairs: a short walk
  return j;
}

// splits[t] = elements of the left run among the outputs before tile t of its pair
__global__ void __launch_bounds__(128)
k_merge_splits(const uint64_t* __restrict__ kin, MergePairs mp, uint32_t tiles, uint32_t* __restrict__ splits) {
  const uint32_t tile = blockIdx.x * blockDim.x + threadIdx.x;
  if (tile >= tiles) return;
  const int j = pair_of_tile(mp, tile);
  const uint64_t* A = kin + mp.a0[j];
  const uint32_t na = mp.na[j], nb = mp.nb[j];
  splits[tile] = merge_path(A, na, A + na, nb, (tile - mp.tile0[j]) * (uint32_t)MG_TILE);
}

// CHECK (first round, on request): *flag |= 1 when a run is not ascending.  Every element of
// both runs is staged by exactly one tile, so comparing neighbours in the staging area (and a
// slice's first element with its predecessor in global memory) sees every adjacent pair.
template <bool PRESPLIT, bool CHECK>
__global__ void __launch_bounds__(MG_THREADS, 4)
k_merge_pairs(const uint64_t* __restrict__ kin, const uint32_t* __restrict__ vin,
              uint64_t* __restrict__ kout, uint32_t* __restrict__ vout, MergePairs mp,
              const uint32_t* __restrict__ splits, uint32_t* __restrict__ flag) {
  extern __shared__ __align__(16) uint8_t mg_smem[];
  uint64_t* sk = reinterpret_cast<uint64_t*>(mg_smem);
  uint32_t* sv = reinterpret_cast<uint32_t*>(mg_smem + (size_t)MG_SLOTS * 8);
  __shared__ uint32_t s_split[2];
  const int tid = threadIdx.x;
  const uint32_t tile = blockIdx.x;
  const int j = pair_of_tile(mp, tile);
  const uint64_t a0 = mp.a0[j];
  const uint32_t na = mp.na[j], nb = mp.nb[j];
  const uint64_t* A = kin + a0;
  const uint64_t* B = A + na;
  const uint32_t d0 = (tile - mp.tile0[j]) * (uint32_t)MG_TILE;
  const uint32_t d1 = min(d0 + (uint32_t)MG_TILE, na + nb);
  uint32_t as, ae;
  if (PRESPLIT) {
    as = splits[tile];
    ae = (tile + 1 < mp.tile0[j + 1]) ? splits[tile + 1] : na;  // the pair's last tile ends its left run
  } else {
    if (tid < 2) s_split[tid] = merge_path(A, na, B, nb, tid == 0 ? d0 : d1);
    __syncthreads();
    as = s_split[0];
    ae = s_split[1];
  }
  // merge-path splits of sorted runs never decrease and never advance by more than a tile; runs
  // that are NOT sorted can break both: keep the slices inside the staging area and say so
  bool bad = false;
  if (ae < as || ae - as > d1 - d0) {
    bad = true;
    ae = ae < as ? as : as + (d1 - d0);
  }
  const uint32_t bs = d0 - as, be = d1 - ae;
  const uint32_t la = ae - as, lb = be - bs;  // la + lb = d1 - d0 <= MG_TILE
  // stage: A slice at [0, la), B slice at [la, la + lb).  A fixed trip count and cp.async keep
  // all 16 loads of a thread in flight at once (a loop over the two slices with a load and a
  // shared-memory store per step went to memory 16 times in a row: ~29 us per tile)
  const uint32_t total = la + lb;
  const uint32_t sk_s = (uint32_t)__cvta_generic_to_shared(sk);
  const uint32_t sv_s = (uint32_t)__cvta_generic_to_shared(sv);
#pragma unroll
  for (int r = 0; r < MG_ITEMS; ++r) {
    const uint32_t i = (uint32_t)tid + (uint32_t)r * MG_THREADS;
    if (i < total) {
      // position of staged element i in the round's input
      const uint64_t g = i < la ? a0 + as + i : a0 + na + bs + (i - la);
      const uint32_t p = mg_slot(i);
      asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(sk_s + p * 8u), "l"(kin + g) : "memory");
      if (vin) asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(sv_s + p * 4u), "l"(vin + g) : "memory");
      else sv[p] = (uint32_t)g;
    }
  }
  asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory");
  __syncthreads();
  if (CHECK) {
#pragma unroll
    for (int r = 0; r < MG_ITEMS; ++r) {
      const uint32_t i = (uint32_t)tid + (uint32_t)r * MG_THREADS;
      if (i < total && i != 0 && i != la) bad |= sk[mg_slot(i - 1)] > sk[mg_slot(i)];
    }
    if (tid == 0 && la > 0 && as > 0) bad |= A[as - 1] > sk[0];
    if (tid == 1 && lb > 0 && bs > 0) bad |= B[bs - 1] > sk[mg_slot(la)];
    if (__syncthreads_or(bad) && tid == 0) atomicOr(flag, 1u);
  }
  const uint32_t my0 = min((uint32_t)tid * MG_ITEMS, total);
  const uint32_t my1 = min(my0 + MG_ITEMS, total);
  uint32_t ia = merge_path_staged(sk, la, lb, my0);
  uint32_t ib = my0 - ia;
  uint64_t rk[MG_ITEMS];
  uint32_t rv[MG_ITEMS];
  uint32_t rs[MG_ITEMS];
  // head keys of both slices in registers; a read past a slice's end stays inside the
  // staging area (key slots, then the values) and is never used
  uint64_t ka = sk[mg_slot(ia)], kb = sk[mg_slot(la + ib)];
#pragma unroll
  for (int i = 0; i < MG_ITEMS; ++i) {
    const bool take_a = ib >= lb || (ia < la && ka <= kb);
    rk[i] = take_a ? ka : kb;
    rs[i] = take_a ? ia : la + ib;
    if (take_a) ka = sk[mg_slot(++ia)];
    else kb = sk[mg_slot(la + (++ib))];
  }
#pragma unroll
  for (int i = 0; i < MG_ITEMS; ++i) rv[i] = sv[mg_slot(min(rs[i], (uint32_t)MG_TILE - 1))];
  __syncthreads();
#pragma unroll
  for (int i = 0; i < MG_ITEMS; ++i) {
    if (my0 + i < my1) {
      sk[mg_slot(my0 + i)] = rk[i];
      sv[mg_slot(my0 + i)] = rv[i];
    }
  }
  __syncthreads();
#pragma unroll
  for (int r = 0; r < MG_ITEMS; ++r) {
    const uint32_t i = (uint32_t)tid + (uint32_t)r * MG_THREADS;
    if (i < total) {
      kout[a0 + d0 + i] = sk[mg_slot(i)];
      vout[a0 + d0 + i] = sv[mg_slot(i)];
    }
  }
}

// *flag |= 1 when some run is not sorted ascending (callers then fall back to a full sort).
// Only for a single run, which is copied; merges check while they stage (k_merge_pairs<., true>).
__global__ void __launch_bounds__(256)
k_check_runs(const uint64_t* __restrict__ keys, uint64_t n, const uint64_t* __restrict__ run_start, int nruns,
             uint32_t* __restrict__ flag) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  bool bad = false;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x + 1; i < n; i += stride) {
    if (keys[i] < keys[i - 1]) {
      // an inversion is fine only across a run boundary
      int lo = 0, hi = nruns;  // first run whose start is > i
      while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (run_start[mid] <= i) lo = mid + 1;
        else hi = mid;
      }
      if (run_start[lo - 1] != i) bad = true;
    }
  }
  if (__syncthreads_or(bad) && threadIdx.x == 0) atomicOr(flag, 1u);
}

// tile splits of one round: at most n / MG_TILE + one partial tile per pair
uint64_t merge_split_bytes(uint64_t n) { return xchg_align_up((n / MG_TILE + MG_MAX_PAIRS + 2) * 4, 256); }

}  // namespace

extern "C" int xchg_merge_workspace_bytes(uint64_t n, uint64_t* bytes_out) {
  XCHG_REQUIRE(bytes_out != nullptr, "bytes_out is NULL");
  if (n > XCHG_MAX_ROWS) {
    xchg_set_error("n=%llu exceeds XCHG_MAX_ROWS", (unsigned long long)n);
    return XCHG_ERR_LIMIT;
  }
  *bytes_out = xchg_align_up(n * 8, 256) + xchg_align_up(n * 4, 256) + 256 + (2 * MG_MAX_PAIRS + 2) * 8 +
               merge_split_bytes(n);
  return XCHG_OK;
}

extern "C" int xchg_merge_sorted_runs(const uint64_t* keys_in, const uint32_t* vals_in,
                                      const uint64_t* run_start_host, int nruns, uint64_t* keys_out,
                                      uint32_t* vals_out, void* workspace, uint64_t workspace_bytes,
                                      uint32_t* unsorted_flag_out, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  XCHG_REQUIRE(nruns >= 0 && (nruns == 0 || run_start_host), "bad runs");
  if (unsorted_flag_out) XCHG_CUDA_TRY(cudaMemsetAsync(unsorted_flag_out, 0, 4, s));
  if (nruns == 0) return XCHG_OK;
  const uint64_t n = run_start_host[nruns];
  if (n == 0) return XCHG_OK;
  XCHG_REQUIRE(keys_in && keys_out && vals_out, "NULL keys_in / keys_out / vals_out");
  XCHG_REQUIRE(run_start_host[0] == 0, "run_start_host[0] must be 0");
  if (n > XCHG_MAX_ROWS) {
    xchg_set_error("n=%llu exceeds XCHG_MAX_ROWS", (unsigned long long)n);
    return XCHG_ERR_LIMIT;
  }
  XCHG_REQUIRE(nruns <= 2 * MG_MAX_PAIRS, "at most %d runs per call", 2 * MG_MAX_PAIRS);
  for (int r = 0; r < nruns; ++r)
    XCHG_REQUIRE(run_start_host[r] <= run_start_host[r + 1], "run_start_host must not decrease");

  XCHG_ONCE_PER_DEVICE(
      XCHG_CUDA_TRY(cudaFuncSetAttribute(k_merge_pairs<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)MG_SMEM));
      XCHG_CUDA_TRY(cudaFuncSetAttribute(k_merge_pairs<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)MG_SMEM));
      XCHG_CUDA_TRY(cudaFuncSetAttribute(k_merge_pairs<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)MG_SMEM));
      XCHG_CUDA_TRY(cudaFuncSetAttribute(k_merge_pairs<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)MG_SMEM)));
  // runs of this round (empty runs dropped: they merge to nothing)
  uint64_t starts[2 * MG_MAX_PAIRS + 1];
  int k = 0;
  for (int r = 0; r < nruns; ++r)
    if (run_start_host[r + 1] > run_start_host[r]) starts[k++] = run_start_host[r];
  starts[k] = n;

  uint8_t* base = (uint8_t*)xchg_align_up((uint64_t)(uintptr_t)workspace, 256);
  uint64_t* ws_keys = (uint64_t*)base;
  uint32_t* ws_vals = (uint32_t*)(base + xchg_align_up(n * 8, 256));
  int rounds = 0;
  for (int kk = k; kk > 1; kk = (kk + 1) / 2) ++rounds;
  // tile splits live behind the run starts; without a workspace that holds them (a caller
  // merging two runs may pass none) the merge CTAs search for themselves
  const uint64_t fixed = xchg_align_up(n * 8, 256) + xchg_align_up(n * 4, 256) + 256 + (2 * MG_MAX_PAIRS + 2) * 8;
  uint32_t* d_splits = nullptr;
  if (workspace && (uint64_t)(base - (uint8_t*)workspace) + fixed + merge_split_bytes(n) <= workspace_bytes)
    d_splits = (uint32_t*)(base + fixed);

  if (unsorted_flag_out && rounds == 0) {
    // a single run is copied, not merged: its order is checked by a pass of its own.
    // run starts to the device: the tail of the workspace
    XCHG_REQUIRE(workspace != nullptr, "the sortedness check needs the workspace");
    uint64_t need = xchg_align_up(n * 8, 256) + xchg_align_up(n * 4, 256);
    XCHG_REQUIRE((uint64_t)(base - (uint8_t*)workspace) + need + (uint64_t)(k + 1) * 8 <= workspace_bytes,
                 "merge workspace too small for the sortedness check");
    uint64_t* d_starts = (uint64_t*)(base + need);
    XCHG_CUDA_TRY(cudaMemcpyAsync(d_starts, starts, (size_t)(k + 1) * 8, cudaMemcpyHostToDevice, s));
    const int blocks = (int)std::min<uint64_t>((n + 255) / 256, (uint64_t)xchg_sm_count() * 8);
    XCHG_LAUNCH(XK_MERGE, s, k_check_runs<<<blocks, 256, 0, s>>>(keys_in, n, d_starts, k, unsorted_flag_out));
    XCHG_LAUNCH_CHECK();
  }

  if (rounds == 0) {  // one run: the merge is a copy
    XCHG_CUDA_TRY(cudaMemcpyAsync(keys_out, keys_in, n * 8, cudaMemcpyDeviceToDevice, s));
    if (vals_in) {
      XCHG_CUDA_TRY(cudaMemcpyAsync(vals_out, vals_in, n * 4, cudaMemcpyDeviceToDevice, s));
    } else {
      MergePairs mp;
      mp.npairs = 1;
      mp.a0[0] = 0;
      mp.na[0] = (uint32_t)n;
      mp.nb[0] = 0;
      mp.tile0[0] = 0;
      mp.tile0[1] = (uint32_t)((n + MG_TILE - 1) / MG_TILE);
      XCHG_LAUNCH(XK_MERGE, s,
                  (k_merge_pairs<false, false><<<mp.tile0[1], MG_THREADS, MG_SMEM, s>>>(keys_in, nullptr, keys_out,
                                                                                        vals_out, mp, nullptr, nullptr)));
      XCHG_LAUNCH_CHECK();
    }
    return XCHG_OK;
  }
  if (rounds > 1) {
    XCHG_REQUIRE(workspace != nullptr, "NULL workspace");
    uint64_t need = xchg_align_up(n * 8, 256) + xchg_align_up(n * 4, 256);
    if ((uint64_t)(base - (uint8_t*)workspace) + need > workspace_bytes) {
      xchg_set_error("merge workspace too small: have %llu, need %llu", (unsigned long long)workspace_bytes,
                     (unsigned long long)(need + 512));
      return XCHG_ERR_WORKSPACE;
    }
  }
  const uint64_t* kin = keys_in;
  const uint32_t* vin = vals_in;
  for (int round = 1; round <= rounds; ++round) {
    // the last round writes the output; the rounds before it alternate towards it
    const bool to_out = ((rounds - round) % 2) == 0;
    uint64_t* ko = to_out ? keys_out : ws_keys;
    uint32_t* vo = to_out ? vals_out : ws_vals;
    MergePairs mp;
    mp.npairs = 0;
    uint32_t tiles = 0;
    uint64_t next[2 * MG_MAX_PAIRS + 1];
    int nk = 0;
    for (int r = 0; r < k; r += 2) {
      const bool has_b = r + 1 < k;  // an odd run out is carried over (merged with nothing)
      const uint64_t a0 = starts[r];
      const uint64_t mid = starts[r + 1];
      const uint64_t end = has_b ? starts[r + 2] : mid;
      const int j = mp.npairs++;
      mp.a0[j] = a0;
      mp.na[j] = (uint32_t)(mid - a0);
      mp.nb[j] = (uint32_t)(end - mid);
      mp.tile0[j] = tiles;
      tiles += (uint32_t)((mp.na[j] + (uint64_t)mp.nb[j] + MG_TILE - 1) / MG_TILE);
      next[nk++] = a0;
    }
    mp.tile0[mp.npairs] = tiles;
    next[nk] = n;
    const bool check = round == 1 && unsorted_flag_out != nullptr;  // every run takes part in round 1
    if (d_splits) {
      XCHG_LAUNCH(XK_MERGE, s, k_merge_splits<<<(tiles + 127) / 128, 128, 0, s>>>(kin, mp, tiles, d_splits));
      XCHG_LAUNCH_CHECK();
      if (check) {
        XCHG_LAUNCH(XK_MERGE, s, (k_merge_pairs<true, true><<<tiles, MG_THREADS, MG_SMEM, s>>>(kin, vin, ko, vo, mp, d_splits,
                                                                                             unsorted_flag_out)));
      } else {
        XCHG_LAUNCH(XK_MERGE, s, (k_merge_pairs<true, false><<<tiles, MG_THREADS, MG_SMEM, s>>>(kin, vin, ko, vo, mp, d_splits,
                                                                                              nullptr)));
      }
    } else if (check) {
      XCHG_LAUNCH(XK_MERGE, s, (k_merge_pairs<false, true><<<tiles, MG_THREADS, MG_SMEM, s>>>(kin, vin, ko, vo, mp, nullptr,
                                                                                            unsorted_flag_out)));
    } else {
      XCHG_LAUNCH(XK_MERGE, s, (k_merge_pairs<false, false><<<tiles, MG_THREADS, MG_SMEM, s>>>(kin, vin, ko, vo, mp, nullptr,
                                                                                             nullptr)));
    }
    XCHG_LAUNCH_CHECK();
    for (int i = 0; i <= nk; ++i) starts[i] = next[i];
    k = nk;
    kin = ko;
    vin = vo;
  }
  return XCHG_OK;
}
