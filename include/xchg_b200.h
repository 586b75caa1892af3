/*
 * xchg_b200.h -- C-ABI of the B200-native exchange engine (libxchg_b200.so).
 *
 * Drop-in boundary for ONE hot path of the reference: Ray Data's all-to-all
 * exchange (Dataset.sort / random_shuffle / groupby().aggregate on Arrow
 * blocks).  The reference has no FFI for this path -- its boundary is the Python
 * class surface BlockAccessor / ExchangeTaskSpec (python/ray/data/block.py:
 * 256-512, python/ray/data/_internal/planner/exchange/interfaces.py:14-102) and
 * its arithmetic lives in pyarrow / numpy calls.  Each entry point below names
 * the reference call site it replaces.  The Python host layer
 * (cluster_anywhere_b200/) binds these with ctypes and mirrors the reference's
 * class surface on top; INTEGRATION.md shows the binding a reference
 * maintainer would add.
 *
 * Conventions
 *  - All pointers are DEVICE pointers unless the name ends in _host.
 *  - Nothing here allocates: the caller passes workspaces (sizes from the
 *    *_workspace_bytes queries) and a CUDA stream (cudaStream_t as void*, NULL =
 *    legacy default stream).  Calls are asynchronous on that stream unless
 *    stated otherwise.
 *  - Every function returns 0 on success, a negative xchg_status otherwise;
 *    xchg_last_error() returns a thread-local message.  There is no CPU
 *    fallback: without a CUDA device every compute entry point fails.
 *  - Row counts per call are limited to XCHG_MAX_ROWS (2^30 - 1).
 *  - Validity bitmaps are Arrow LSB-first bitmaps; NULL means "no nulls".
 */
#ifndef XCHG_B200_H_
#define XCHG_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define XCHG_ABI_VERSION 1
#define XCHG_MAX_ROWS ((uint64_t)((1u << 30) - 1))
/* widest key range xchg_dense_groups addresses directly */
#define XCHG_DENSE_MAX_RANGE ((uint64_t)1 << 28)

typedef enum {
  XCHG_OK = 0,
  XCHG_ERR_INVALID = -1,   /* bad argument */
  XCHG_ERR_CUDA = -2,      /* CUDA runtime error (message has the detail) */
  XCHG_ERR_WORKSPACE = -3, /* workspace too small */
  XCHG_ERR_LIMIT = -4      /* size beyond a documented maximum */
} xchg_status;

/* Physical key types (Arrow primitive layouts). */
typedef enum {
  XCHG_I8 = 0, XCHG_I16 = 1, XCHG_I32 = 2, XCHG_I64 = 3,
  XCHG_U8 = 4, XCHG_U16 = 5, XCHG_U32 = 6, XCHG_U64 = 7,
  XCHG_F32 = 8, XCHG_F64 = 9
} xchg_key_type;

int xchg_abi_version(void);
const char* xchg_last_error(void);
/* Number of visible CUDA devices (0 if none; never fails). */
int xchg_device_count(void);

/* ------------------------------------------------------- measurement ---- */
/* No counterpart in the reference (it reports wall-clock task stats,
 * python/ray/data/block.py:141-207).  Kernels launched by this library since
 * load; and optional per-kernel device timing: while enabled, every launch is
 * bracketed by CUDA events on its stream; xchg_profile_read synchronises the
 * recorded events and returns the accumulated milliseconds / launch count of one
 * kernel class (ids: 0 normalize+hist, 1 scan_hist, 2 onesweep pass, 3 row
 * gather, 4 fixed gather, 5 bitmap gather, 6 offsets scan, 7 partition bounds,
 * 8 offsets utilities, 9 misc, 10 merge, 11 shuffle, 12 aggregate, 13 local
 * (shared-memory) sort, 14 MSD level bookkeeping: segment histograms, scans, lists). */
uint64_t xchg_launch_count(void);
int xchg_profile_enable(int on);
int xchg_profile_read(int kernel_id, double* total_ms, uint64_t* launches);

/* Device-wide hint (cudaLimitMaxL2FetchGranularity) for the CURRENT device: how
 * many bytes the L2 fetches from HBM per miss.  The gathers of this engine read
 * random 8..128-byte rows; with the 128-byte default a 90-byte row costs ~219 B of
 * DRAM reads (profiles/r01_summary.md), with 32 it costs its sectors.  No
 * counterpart in the reference.  bytes = 32, 64 or 128. */
int xchg_set_l2_fetch_granularity(int bytes);

/* Cap on the CTAs per SM the gather kernels launch (0 = each kernel's default).  The
 * multi-GPU exchange sets a small cap while its peer-store gathers -- NVLink-bound, a
 * few CTAs per SM saturate the links -- run on a side stream beside the receive-side
 * key sort, so that they do not take the SMs' shared memory away from it.  Process
 * wide; no counterpart in the reference. */
int xchg_set_gather_ctas_per_sm(int ctas_per_sm);
/* Shared-memory staging per CTA (KB, 4..56; 0 = the default 56) of the uniform-row gather
 * for the launches that follow: a gather that stores into peer memory is link-bound and runs
 * with a few KB per CTA so that the receive-side sort and gathers keep the SM's shared
 * memory.  Process wide; no counterpart in the reference. */
int xchg_set_gather_stage_kb(int kb);

/* ---------------------------------------------------------------- sort ---- */

/* Which passes of xchg_sort_column go through HBM.  0 (default): the hybrid -- passes
 * from the top digit place down, every segment of equal higher digits finished inside
 * shared memory as soon as it has <= 4096 rows (random 64-bit keys: 2 global passes + 1
 * local sort instead of 8 global passes).  1: every pass global (LSD), kept for A/B
 * measurements; the environment variable XCHG_SORT_ALGO=lsd selects it at load.  Both
 * are stable and produce identical output.  Process wide; no counterpart in the
 * reference (pyarrow picks its own algorithm inside sort_indices). */
int xchg_set_sort_algorithm(int algorithm);

/* Workspace bytes xchg_sort_column needs for n rows. */
int xchg_sort_workspace_bytes(uint64_t n, uint64_t* bytes_out);

/*
 * Stable sort of one key column -> row permutation.
 * Replaces pyarrow.compute.sort_indices as called from
 * python/ray/data/_internal/arrow_ops/transform_pyarrow.py:41 (sort) and :546
 * (concat_and_sort), one key column per call; multi-key sorts chain calls from
 * the LAST key to the first, feeding perm_out back as perm_in (LSD over
 * columns; every call is stable).
 *
 * Order (Arrow's, both directions): finite numbers in the requested direction,
 * then NaN, then null; ties keep input order; -0.0 == +0.0.
 *
 *  keys        n values of key_type (indexed by ORIGINAL row id)
 *  validity    Arrow validity bitmap of the column or NULL; bit i+validity_bit_offset
 *  perm_in     NULL = identity; else the current row order (n row ids)
 *  perm_out    n row ids: perm_out[j] = original row id at sorted position j
 *  norm_keys_out  optional (may be NULL): n order-preserving 64-bit radix keys in
 *              sorted order (ascending as unsigned whatever the direction);
 *              input to xchg_partition_bounds
 *  Synchronises the stream internally: once for the 9 pass-activity flags, and in the
 *  hybrid once per global level after the first (segment counts; two for random keys).
 */
int xchg_sort_column(const void* keys, int key_type, int descending,
                     const uint8_t* validity, int64_t validity_bit_offset,
                     const uint32_t* perm_in, uint64_t n,
                     uint32_t* perm_out, uint64_t* norm_keys_out,
                     void* workspace, uint64_t workspace_bytes, void* stream);

/*
 * Key normalisation only (no sort): norm_out[i] = radix key of
 * keys[perm ? perm[i] : i].  Used to build the secondary-key arrays
 * xchg_partition_bounds searches.  For 64-bit integer columns nulls cannot be
 * folded into 64 bits; they map to UINT64_MAX (callers treat that value as
 * "null" only when the column has nulls).
 */
int xchg_normalize_keys(const void* keys, int key_type, int descending,
                        const uint8_t* validity, int64_t validity_bit_offset,
                        const uint32_t* perm, uint64_t n, uint64_t* norm_out,
                        void* stream);

/*
 * Inverse of the key normalisation for INTEGER key types: sorted radix keys ->
 * sorted raw key values, out[i] = key whose radix key is norm[i].  Produces the
 * sorted key column sequentially instead of gathering it through the
 * permutation (Table.take on the key column,
 * python/ray/data/_internal/arrow_ops/transform_pyarrow.py:70).  Slots of null
 * rows receive an unspecified value (they are masked by the validity bitmap).
 */
int xchg_denormalize_keys(const uint64_t* norm, int key_type, int descending, uint64_t n,
                          void* out, void* stream);

/*
 * Boundary search on sorted normalised keys -> partition cut indices.
 * Replaces find_partition_index / find_partitions
 * (python/ray/data/_internal/util.py:728-828): for boundary b, walk the key
 * columns c = 0..ncols-1 narrowing [lo,hi) to the run whose normalised key lies
 * in [lower[b*ncols+c], upper[b*ncols+c]]; stop early when lo == hi; result is
 * hi if first_descending else lo.  flags per (boundary, column): bit0 = the equal
 * range is empty (hi = lo = lower_bound), bit1 = boundary tuple ends before this
 * column, bit2 = None boundary component (sorts after every row of the run on an
 * ascending column: lo = hi = run end; before every row on a descending one:
 * lo = hi = run start -- util.py:765-766 with :775-799),
 * bit3 = boundary is NaN.
 * Descending columns reproduce numpy.searchsorted over the reference's reversed
 * view probe by probe (util.py:775-799), so that cut points stay identical even
 * where NaN / null keys make that view unsorted for numpy; keys >=
 * last_threshold_host[c] are the NaN / null class when has_last_host[c] != 0.
 *
 *  norm_cols_host   HOST array of ncols device pointers, each n keys in final
 *                   sorted order; col_desc_host / last_threshold_host /
 *                   has_last_host: HOST arrays of ncols entries
 *  lower/upper/flags  nb*ncols entries, device
 *  cuts_out    nb uint64 row indices, device
 */
int xchg_partition_bounds(const uint64_t* const* norm_cols_host, const uint8_t* col_desc_host,
                          const uint64_t* last_threshold_host, const uint8_t* has_last_host,
                          int ncols, uint64_t n,
                          const uint64_t* lower, const uint64_t* upper,
                          const uint8_t* flags, int nb, int first_descending,
                          uint64_t* cuts_out, void* stream);

/*
 * The same boundary rule stated per ROW, for rows that are not sorted yet:
 * dest_out[i] = index of the range partition row i belongs to (0..nb), i.e. the
 * partition find_partitions (python/ray/data/_internal/util.py:806-828) would
 * slice it into after the sort: the number of boundaries whose cut index is <= the
 * row's sorted position (cut = lo -> boundaries the row is not strictly before;
 * cut = hi when first_descending -> boundaries the row is strictly after,
 * util.py:803).  Used by the multi-GPU exchange to split a rank's rows by
 * destination before the sort (the "map" side of SortTaskSpec,
 * python/ray/data/_internal/planner/exchange/sort_task_spec.py:127-138, without
 * sorting first).  Boundaries must be in sort order.
 *
 *  norm_cols_host  HOST array of ncols device pointers: normalised keys of every
 *                  key column in ROW order (xchg_normalize_keys with perm = NULL);
 *                  col_desc_host: HOST array of ncols direction flags
 *  lower/upper/flags  as xchg_partition_bounds (nb*ncols entries, device)
 *  dest_out     n uint32, device
 *  counts_out   optional: nb+1 uint64 rows per partition, device (zeroed here)
 */
int xchg_assign_partitions(const uint64_t* const* norm_cols_host, const uint8_t* col_desc_host,
                           int ncols, uint64_t n,
                           const uint64_t* lower, const uint64_t* upper, const uint8_t* flags,
                           int nb, int first_descending, uint32_t* dest_out,
                           uint64_t* counts_out, void* stream);

/* ------------------------------------- multi-GPU exchange: control path ---- */
/*
 * K1, boundary sampling.  Replaces SortTaskSpec._sample_block -> ArrowBlockAccessor._sample
 * (python/ray/data/_internal/planner/exchange/sort_task_spec.py:229,
 * _internal/arrow_block.py:341-344: `random.sample` of min(n_samples, rows) rows of every
 * block and a take of the key column).  Block b = rows [block_start_host[b],
 * +block_rows_host[b]) of the key column; min(per_block, rows) DISTINCT rows of it are
 * drawn (the first values of a Philox-keyed bijection of [0, rows), keyed by seed and the
 * block's global index first_block_index + b; the reference's draw is unseeded) and their
 * radix keys (xchg_normalize_keys' transform) are written back to back to samples_out.
 * *count_out_host = number of samples (known without a read-back).  XCHG_ERR_LIMIT if
 * they exceed `capacity`.
 */
int xchg_sample_keys(const void* keys, int key_type, int descending, const uint8_t* validity,
                     int64_t validity_bit_offset, const uint64_t* block_start_host,
                     const uint64_t* block_rows_host, int nblocks, uint32_t per_block,
                     uint64_t seed, uint64_t first_block_index, uint64_t* samples_out,
                     uint64_t capacity, uint64_t* count_out_host, void* stream);
/*
 * The driver half of SortTaskSpec.sample_boundaries (sort_task_spec.py:200-226) on the
 * device: sort all samples, pick the num_partitions - 1 boundaries at
 * int(q * (len - 1)), q in linspace(0, 1, num_partitions + 1)[1:-1] (float64, as numpy).
 * `gathered` = nmsgs (<= 64) messages of msg_words uint64 each, as one all-gather of the
 * ranks' sample buffers delivers them: word count_word of a message = its number of
 * samples, samples from word hdr_words on.  More than 16384 samples in total are thinned
 * out evenly first.  bounds_out: num_partitions - 1 radix keys in sort order (untouched
 * when there is no sample); info_out (may be NULL): {samples received, samples used}.
 * One CTA; every rank computes identical boundaries from identical input.
 */
int xchg_select_boundaries(const uint64_t* gathered, int nmsgs, uint64_t msg_words,
                           uint32_t count_word, uint32_t hdr_words, uint32_t num_partitions,
                           uint64_t* bounds_out, uint64_t* info_out, void* stream);
/*
 * Map side of the exchange for ONE fixed-width key column without nulls, boundaries
 * resident on the device: destination of every row = the range partition
 * find_partitions (python/ray/data/_internal/util.py:806-828) slices it into, for the
 * nthr (<= 255) thresholds bounds[pick_host[i]] (in sort order).  strict_host[i] != 0: rows
 * equal to threshold i stay left of it (descending keys, util.py:803); NULL = none strict.
 * Same rule as xchg_assign_partitions, without materialising radix keys or a destination
 * array.
 *
 * xchg_partition_hist   counts_out[d] (nthr + 1 uint64, device, zeroed here) = rows of
 *                       destination d.
 * xchg_partition_rows   stable partition: keys_out / rows_out (n entries each) = radix key
 *                       and row id of every row, ordered by destination, input order kept
 *                       inside a destination (SortTaskSpec.map without the sort,
 *                       sort_task_spec.py:127-138).  counts = xchg_partition_hist's output
 *                       for the same arguments.  One RAW onesweep pass: 8 B read + 12 B
 *                       written per row.
 */
int xchg_partition_hist(const void* keys, int key_type, int descending, uint64_t n,
                        const uint64_t* bounds, const uint32_t* pick_host, const uint8_t* strict_host,
                        int nthr, uint64_t* counts_out, void* stream);
int xchg_partition_workspace_bytes(uint64_t n, uint64_t* bytes_out);
int xchg_partition_rows(const void* keys, int key_type, int descending, uint64_t n,
                        const uint64_t* bounds, const uint32_t* pick_host, const uint8_t* strict_host,
                        int nthr, const uint64_t* counts, uint64_t* keys_out, uint32_t* rows_out,
                        void* workspace, uint64_t workspace_bytes, void* stream);
/*
 * dst_host[i][0 .. bytes_host[i]) = src_host[i][...] for nseg segments in one launch, the
 * segments' tiles interleaved; destinations may be peer memory (the key runs of the
 * exchange).  Replaces the object-store transfer of a map task's outputs
 * (pull_based_shuffle_task_scheduler.py:84-120) for contiguous runs.
 */
int xchg_copy_segments(int nseg, const void* const* src_host, void* const* dst_host,
                       const uint64_t* bytes_host, void* stream);
/*
 * Arrival flags between GPUs.  The reference orders map, merge and reduce rounds by task
 * dependencies on object references (push_based_shuffle_task_scheduler.py:548-561,
 * :696-730); here a sender announces that a chunk has been stored by writing `value` to a
 * 4-byte flag in the receiver's memory, and the receiver's stream waits for it.
 * xchg_signal_flags  every flag_ptrs_host[i] (device pointers, peer memory allowed) := value,
 *                    released at system scope after all earlier work of the stream.
 * xchg_wait_flags    blocks the stream until flags[0..n) (local memory) have all reached
 *                    `value` (wrap-around compare).  After timeout_ns without it,
 *                    *error_flag (device uint32) := 1 + index of a missing flag and the
 *                    stream continues: check it before trusting what follows.
 */
int xchg_signal_flags(uint32_t* const* flag_ptrs_host, int n, uint32_t value, void* stream);
int xchg_wait_flags(const uint32_t* flags, int n, uint32_t value, uint64_t timeout_ns,
                    uint32_t* error_flag, void* stream);

/* ---------------------------------------------------------------- merge ---- */
/*
 * K6: stable k-way merge of sorted runs.  Replaces the reduce side of the reference's
 * sort -- ArrowBlockAccessor.merge_sorted_blocks -> concat_and_sort
 * (python/ray/data/_internal/arrow_block.py:529-541, _internal/arrow_ops/
 * transform_pyarrow.py:540-547: concatenate the runs, sort everything again) -- by a merge:
 * the result is identical (ties keep (run index, position in run) order), every key is
 * read ceil(log2(nruns)) times instead of being re-sorted.
 *
 *  keys_in    radix keys (xchg_normalize_keys) of all runs back to back; run r =
 *             [run_start_host[r], run_start_host[r+1]), ascending inside every run
 *  vals_in    row ids travelling with the keys, or NULL = positions 0..n-1
 *  run_start_host  HOST array, nruns + 1 entries, run_start_host[0] = 0 (<= 128 runs)
 *  keys_out / vals_out  n merged entries each (must not alias the inputs)
 *  workspace  xchg_merge_workspace_bytes(n) bytes
 *  unsorted_flag_out  optional device uint32: set to 1 when some run is NOT ascending (the
 *             output is then meaningless; callers fall back to xchg_sort_column)
 */
int xchg_merge_workspace_bytes(uint64_t n, uint64_t* bytes_out);
int xchg_merge_sorted_runs(const uint64_t* keys_in, const uint32_t* vals_in,
                           const uint64_t* run_start_host, int nruns, uint64_t* keys_out,
                           uint32_t* vals_out, void* workspace, uint64_t workspace_bytes,
                           uint32_t* unsorted_flag_out, void* stream);

/* ------------------------------------------------- string / binary keys ---- */
/*
 * Sort keys of variable-width columns (Arrow binary / string, int64 byte offsets as this
 * engine stages them).  Arrow's order -- what pyarrow.compute.sort_indices gives the
 * reference at python/ray/data/_internal/arrow_ops/transform_pyarrow.py:41 -- is bytewise
 * (unsigned), a proper prefix before the longer string.  That order equals the
 * lexicographic order of (chunk 0, chunk 1, ..., length) where chunk c = bytes
 * [8c, 8c+8) as a big-endian u64, zero padded; callers sort LSD over those columns with
 * xchg_sort_column (length as XCHG_U32 first, then the chunks as XCHG_U64 from the last
 * to the first, the column's validity on the final call), chaining perm_out -> perm_in.
 *
 * xchg_string_lengths   len_out[i] = bytes of row i; *max_len_out (device uint64) = the
 *                       longest row (tells the caller how many chunks exist)
 * xchg_string_chunk_keys  keys_out[i] = chunk `chunk` of row i
 */
int xchg_string_lengths(const int64_t* offsets, uint64_t n, uint32_t* len_out, uint64_t* max_len_out,
                        void* stream);
int xchg_string_chunk_keys(const int64_t* offsets, const uint8_t* data, uint64_t n, uint32_t chunk,
                           uint64_t* keys_out, void* stream);
/*
 * find_partition_index (python/ray/data/_internal/util.py:728-803) for ONE string key
 * column: cut index of every boundary on the SORTED column (nulls last; the reference
 * replaces nulls by NULL_SENTINEL, :765-772, so a None boundary equals the null rows).
 * Boundaries are packed like an Arrow binary array (bound_offsets nb+1 entries,
 * bound_data); bound_flags[b] bit2 = None boundary.  Descending reproduces
 * numpy.searchsorted over the reversed view probe by probe, as xchg_partition_bounds does.
 */
int xchg_partition_bounds_bytes(const int64_t* sorted_offsets, const uint8_t* sorted_data,
                                const uint8_t* sorted_validity, uint64_t n,
                                const int64_t* bound_offsets, const uint8_t* bound_data,
                                const uint8_t* bound_flags, int nb, int descending,
                                uint64_t* cuts_out, void* stream);
/*
 * find_partition_index for key TUPLES that mix numeric and string / binary columns (the
 * reference narrows [left, right) column by column whatever their types, util.py:757-802).
 * cols_host: HOST array of ncols descriptors.  kind 0 = numeric: `norm` holds the column's n
 * radix keys in sorted order, lower / upper / flags entries [b * ncols + c] as for
 * xchg_partition_bounds; kind 1 = string: the SORTED column (offsets, data, validity) and
 * the boundary strings of that column packed like an Arrow binary array (bound_offsets nb + 1
 * entries, bound_data); of flags only bit1 (tuple ended) and bit2 (None component: equals the
 * null rows, which hold NULL_SENTINEL) are read.  Descending columns replay numpy's probes
 * over the reversed view.  cuts_out: nb uint64, device.
 */
typedef struct {
  int32_t kind;             /* 0 numeric, 1 string / binary */
  int32_t descending;
  const uint64_t* norm;     /* kind 0 */
  uint64_t last_threshold;  /* kind 0: radix keys >= this are the NaN / null class ... */
  int32_t has_last;         /* ... when has_last != 0 */
  int32_t reserved;
  const int64_t* offsets;   /* kind 1: n + 1 byte offsets */
  const uint8_t* data;
  const uint8_t* validity;  /* may be NULL */
  const int64_t* bound_offsets;
  const uint8_t* bound_data;
} xchg_mixed_col;
int xchg_partition_bounds_mixed(const xchg_mixed_col* cols_host, int ncols, uint64_t n,
                                const uint64_t* lower, const uint64_t* upper, const uint8_t* flags,
                                int nb, int first_descending, uint64_t* cuts_out, void* stream);
/* xchg_assign_partitions for such tuples: per ROW of the UNSORTED columns (kind 0: `norm` holds
 * the radix keys in row order; kind 1: the column as stored).  counts_out may be NULL. */
int xchg_assign_partitions_mixed(const xchg_mixed_col* cols_host, int ncols, uint64_t n,
                                 const uint64_t* lower, const uint64_t* upper, const uint8_t* flags,
                                 int nb, int first_descending, uint32_t* dest_out, uint64_t* counts_out,
                                 void* stream);
/* The same rule per row of an UNSORTED string column (see xchg_assign_partitions). */
int xchg_assign_partitions_bytes(const int64_t* offsets, const uint8_t* data, const uint8_t* validity,
                                 uint64_t n, const int64_t* bound_offsets, const uint8_t* bound_data,
                                 const uint8_t* bound_flags, int nb, int descending,
                                 uint32_t* dest_out, uint64_t* counts_out, void* stream);

/* head_out[j] = 1 unless row j of the level-0 order is known to repeat the string of row j-1 in
 * the FINAL order (fflag[j] == 0 and its tie group is closed; active may be NULL = no open
 * group), or j is a multiple of `period` (0 = never).  Runs of repeats let a caller copy a sorted
 * key column from one source row per run instead of fetching every row's own copy; the period
 * keeps a heavy hitter from becoming ONE hot line that every SM reads. */
int xchg_string_run_heads(const uint8_t* fflag, const uint32_t* active, uint64_t m, uint32_t period,
                          uint8_t* head_out, void* stream);
/*
 * MSD refinement of a string sort.  After a stable sort by chunk 0 only the rows that tie
 * there can be out of order, and of those only the ones whose tie group holds at least two
 * DIFFERENT strings (a heavy hitter's run of identical keys is final).  The entry points
 * below work on a LIST of m rows in their current order -- level 0: every row, in the order
 * xchg_sort_column produced; later levels: the rows still unresolved -- and let the caller
 * re-sort only those by the next chunk (xchg_sort_column on the list's chunk keys, then on
 * the list's group ids, both stable).  Same result as the LSD chain over every chunk: the
 * order pyarrow.compute.sort_indices gives the reference (transform_pyarrow.py:41).
 *
 * xchg_string_tie_flags     gflag[j] = list entry j starts a new tie group (previous-level
 *                           group id, this level's key or null-ness differs from entry j-1);
 *                           fflag[j] = gflag[j], or the two rows' strings differ anywhere from
 *                           byte from_byte on (lengths included).  rows = the list's row ids,
 *                           keys = its keys at this level, gid_prev may be NULL (one group).
 * xchg_string_active_rows   active_out[j] = 1 when entry j's group (gid from a scan of gflag,
 *                           num_groups groups) contains two different strings, else 0.
 *                           unresolved_scratch: num_groups bytes.
 * xchg_scatter_flagged      list_out[excl_scan[j]] = j for every j with flags[j] != 0.
 * xchg_string_chunk_keys_rows  xchg_string_chunk_keys for listed rows; chunk = UINT32_MAX
 *                           gives the rows' lengths (the last level: a proper prefix first).
 * xchg_scatter_u32          out[idx[j]] = vals[j].
 */
int xchg_string_tie_flags(const int64_t* offsets, const uint8_t* data, const uint8_t* validity,
                          const uint32_t* rows, const uint64_t* keys, const uint32_t* gid_prev,
                          uint64_t m, uint64_t from_byte, uint8_t* gflag_out, uint8_t* fflag_out,
                          void* stream);
int xchg_string_active_rows(const uint8_t* gflag, const uint8_t* fflag, const uint32_t* gid, uint64_t m,
                            uint64_t num_groups, uint8_t* unresolved_scratch, uint32_t* active_out,
                            void* stream);
int xchg_scatter_flagged(const uint32_t* flags, const int64_t* excl_scan, uint64_t m, uint32_t* list_out,
                         void* stream);
int xchg_string_chunk_keys_rows(const int64_t* offsets, const uint8_t* data, const uint32_t* rows,
                                uint64_t m, uint32_t chunk, uint64_t* keys_out, void* stream);
int xchg_scatter_u32(const uint32_t* idx, const uint32_t* vals, uint64_t m, uint32_t* out, void* stream);

/* -------------------------------------------------------------- gather ---- */

/*
 * out[i] = in[idx[i]] for fixed-width elements of `width` bytes (1,2,4,8,16 are
 * vectorised; any other width takes the byte-row path).  Replaces
 * pyarrow.Table.take for primitive columns
 * (python/ray/data/_internal/arrow_ops/transform_pyarrow.py:45-71).
 */
int xchg_gather_fixed(const void* in, uint32_t width, const uint32_t* idx, uint64_t n_out,
                      void* out, void* stream);

/* out bit i = in bit (idx[i] + in_bit_offset); also counts nulls (zero bits)
 * into *null_count_out (device uint64, may be NULL; must be zeroed by caller). */
int xchg_gather_bitmap(const uint8_t* in, int64_t in_bit_offset, const uint32_t* idx,
                       uint64_t n_out, uint8_t* out, uint64_t* null_count_out, void* stream);

/*
 * Variable-width take (Arrow binary/string/large_* /list child bytes).
 * Replaces Table.take on var-width columns (same call site).
 *  step 1  xchg_gather_offsets: out_off[0]=0, out_off[i+1]=out_off[i]+len(idx[i])
 *          (int64, n_out+1 entries); in_off is int64 (n_in+1 entries, absolute).
 *  step 2  xchg_gather_bytes: copies the row bytes.
 */
int xchg_gather_offsets_workspace_bytes(uint64_t n_out, uint64_t* bytes_out);
int xchg_gather_offsets(const int64_t* in_off, const uint32_t* idx, uint64_t n_out,
                        int64_t* out_off, void* workspace, uint64_t workspace_bytes,
                        void* stream);
/*
 * Allocation-free form of xchg_gather_offsets (no counterpart in the reference).
 * For hosts that are handed their device buffers by another component and have
 * no way to obtain scratch memory themselves (all device memory belongs to a
 * foreign pool, or cudaMalloc's device-wide serialisation is unacceptable in
 * their loop): the scan's scratch is a fixed 256 KiB array inside the library
 * image, resident in every context that loads the library until exit.
 *  - n_out above XCHG_STATIC_GATHER_MAX_ROWS is refused with XCHG_ERR_LIMIT.
 *  - All calls on a device share that array; calls whose kernels may overlap on
 *    the GPU race on it, so serialise them.  Left-over contents of an earlier
 *    call are harmless (each call overwrites every slot before reading it).
 *  - The first call on each device resolves the array's device address; make it
 *    outside any stream capture, with the target device current.  The address is
 *    cached per device ordinal for the life of the process: after cudaDeviceReset
 *    (or any teardown of the device's primary context) it is stale and this entry
 *    point must not be used again in that process.
 * Output is identical to xchg_gather_offsets.
 */
#define XCHG_STATIC_GATHER_MAX_ROWS ((uint64_t)(64u << 20))
int xchg_gather_offsets_static(const int64_t* in_off, const uint32_t* idx, uint64_t n_out,
                               int64_t* out_off, void* stream);
int xchg_gather_bytes(const uint8_t* in_data, const int64_t* in_off, const uint32_t* idx,
                      uint64_t n_out, const int64_t* out_off, uint8_t* out_data,
                      void* stream);
/*
 * The same take with the rows' source spans kept between the two steps: xchg_gather_spans is
 * xchg_gather_offsets that also stores src_start_out[i] = in_off[idx[i]] (int64, n_out entries);
 * xchg_gather_bytes_spans copies the bytes from those starts.  in_off -- one random 16-byte read
 * per row, a 64-128-byte DRAM fill -- is then read once per take instead of three times.  Same
 * output; workspace as xchg_gather_offsets_workspace_bytes.
 */
int xchg_gather_spans(const int64_t* in_off, const uint32_t* idx, uint64_t n_out,
                      int64_t* src_start_out, int64_t* out_off, void* workspace,
                      uint64_t workspace_bytes, void* stream);
int xchg_gather_bytes_spans(const uint8_t* in_data, const int64_t* src_start, uint64_t n_out,
                            const int64_t* out_off, uint8_t* out_data, void* stream);

/* Uniform-width rows (every row `width` bytes; GraySort payload, tensor rows):
 * out row i = in row idx[i]; no offsets are read. */
/* xchg_gather_rows_multi: the same for up to 16 output segments in ONE launch: segment d
 * takes idx[seg_row0[d] .. seg_row0[d] + seg_rows[d]) and writes seg_out[d] (HOST arrays of
 * nseg entries; seg_out[d] may point into a peer GPU's memory).  The segments' tiles are
 * interleaved so that all outputs are written at the same pace: the multi-GPU exchange's
 * stores reach every peer at once instead of one destination after the other.  Replaces the
 * per-reducer transfers of pull_based_shuffle_task_scheduler.py:84-120 like xchg_gather_rows
 * into a peer arena does.  XCHG_ERR_LIMIT when a row does not fit the staging tile (~56 KB). */
int xchg_gather_rows_multi(const void* in_data, uint32_t width, const uint32_t* idx, int nseg,
                           const uint64_t* seg_row0_host, const uint64_t* seg_rows_host,
                           void* const* seg_out_host, void* stream);
int xchg_gather_rows(const uint8_t* in_data, uint32_t width, const uint32_t* idx,
                     uint64_t n_out, uint8_t* out_data, void* stream);

/* Offsets utilities for Arrow staging: widen int32 offsets to absolute int64
 * (out[i] = in[i] - in[0] + base) and narrow back (out[i] = in[i] - in[0]). */
int xchg_offsets_widen(const int32_t* in, uint64_t count, int64_t base, int64_t* out, void* stream);
int xchg_offsets_narrow(const int64_t* in, uint64_t count, int32_t* out, void* stream);
/* out[i] = base + i*width, count entries (offsets of uniform rows). */
int xchg_offsets_uniform32(uint64_t count, uint32_t width, int32_t* out, void* stream);
int xchg_offsets_uniform64(uint64_t count, uint64_t width, int64_t base, int64_t* out, void* stream);
/* *flag_out (device uint32, caller zeroes) becomes nonzero unless every row has
 * `width` bytes. */
int xchg_offsets_check_uniform(const int64_t* off, uint64_t n_rows, uint64_t width,
                               uint32_t* flag_out, void* stream);

/* ------------------------------------------------------------- shuffle ---- */

/*
 * Seeded random permutation of [0, n_total).  Replaces
 * numpy.random.RandomState(seed).permutation(n) in ArrowBlockAccessor.random_shuffle
 * (python/ray/data/_internal/arrow_block.py:225-231; called from
 * ShuffleTaskSpec.map/reduce, planner/exchange/shuffle_task_spec.py:81-84, :120-123).
 * The reference's MT19937 values are not pinned by its tests; this permutation is a
 * Philox-2x32-10-keyed 4-round Feistel network with cycle walking: a bijection of
 * [0, n_total) for every seed, evaluable per element and per sub-range.
 *   inverse = 0:  out[i] = P(first + i)       output position of source row first+i
 *   inverse = 1:  out[i] = P^-1(first + i)    source row of output position first+i
 * so `take(P^-1(0..n))` is the shuffled table.  n_total <= 2^32 - 1.
 */
int xchg_shuffle_positions(uint64_t seed, uint64_t n_total, uint64_t first, uint64_t count,
                           int inverse, uint32_t* out, void* stream);
/* inv_out[pos[i] - base] = i  (pos holds n distinct values in [base, base+n)):
 * turns "row i goes to position pos[i]" into the gather index of that range. */
int xchg_invert_permutation(const uint32_t* pos, uint64_t n, uint64_t base, uint32_t* inv_out,
                            void* stream);
/* Owner of each position under an even split of [0, n_total) into `parts`
 * contiguous ranges (range p = [floor(p*n_total/parts), floor((p+1)*n_total/parts)) ):
 * dest_out[i] = p, counts_out[p] = rows (uint64[parts], zeroed here).  The shuffle's
 * analogue of the reference's R equal slices per mapper (shuffle_task_spec.py:90-93). */
int xchg_even_split_dest(const uint32_t* pos, uint64_t n, uint64_t n_total, uint32_t parts,
                         uint32_t* dest_out, uint64_t* counts_out, void* stream);

/* ----------------------------------------------------------- aggregate ---- */

/*
 * Hash group-by aggregate (one key column).  Replaces the arithmetic of
 * ArrowBlockAccessor.combine (python/ray/data/_internal/arrow_block.py:452-526: a
 * Python loop over groups, one pyarrow.compute call per group and aggregate,
 * :346-423) and of aggregate_combined_blocks (:544-663: heapq.merge over rows), for
 * Count / Sum / Min / Max / Mean / Std (python/ray/data/aggregate.py:97-326) with the
 * null policy of _internal/null_aggregate.py:11-25.
 *
 * Call sequence
 *   xchg_normalize_keys(key column)                      -> radix keys (row order)
 *   xchg_hash_groups(...)                                -> slot per row, group list
 *   xchg_sort_column(group keys as XCHG_U64)             -> groups in output order
 *                                                           (numbers < NaN < null)
 *   xchg_group_ids(...)                                  -> dense group id per row
 *   xchg_agg_stats_init / xchg_agg_count_rows / xchg_agg_accumulate[_m2]   (raw rows)
 *     or xchg_agg_merge_rows / xchg_agg_merge_states                      (partials)
 *   xchg_agg_finalize (per aggregate)  or  xchg_agg_pack_states (emit partials)
 *   xchg_group_keys_out                                  -> key column of the result
 *
 * Per value column the statistics live in `stats`: 5 arrays of num_groups 8-byte
 * words [cnt | sum | min | max | nan] (sum is int64 for integer columns, double
 * otherwise; min/max are order-preserving encodings) plus an optional double m2[]
 * for Std.  A PARTIAL row (map-side output, the reference's accumulator columns
 * _internal/null_aggregate.py:28-45) is 6 words: cnt, sum, min, max, nan, M2.
 */
/* Bytes of table workspace for `capacity` slots (a power of two, >= 2 x distinct keys). */
int xchg_hash_table_bytes(uint64_t capacity, uint64_t* bytes_out);
/*
 * Build the table.  slot_out[i] = slot of row i; group_keys_out / group_slots_out
 * (capacity + 2 entries each, caller-allocated) receive the num_groups distinct
 * (radix key, slot) pairs in table order, null group last.  Null rows (validity)
 * form their own group.  Synchronises the stream (reads the group count);
 * XCHG_ERR_LIMIT if the table overflowed.
 */
int xchg_hash_groups(const uint64_t* norm_keys, const uint8_t* validity, int64_t validity_bit_offset,
                     uint64_t n, uint64_t capacity, void* table_ws, uint64_t table_ws_bytes,
                     uint32_t* slot_out, uint64_t* group_keys_out, uint32_t* group_slots_out,
                     uint64_t* num_groups_out, void* stream);
/*
 * Direct-address alternative to xchg_hash_groups + xchg_sort_column + xchg_group_ids for
 * keys whose radix keys span a small range (integer keys such as ids, categories, dates:
 * BASELINE config 4 groups 2 B rows by an int32 key with 10 M distinct values).  Same
 * reference arithmetic as above (the grouping half of ArrowBlockAccessor.combine,
 * arrow_block.py:452-526); the reference has no such special case -- it always sorts.
 *
 * xchg_key_minmax: out3_host = {min, max, number of valid rows} of the radix keys of the
 *   valid rows ({UINT64_MAX, 0, 0} when there is none).  scratch32: 32 device bytes.
 *   Synchronises the stream.
 * xchg_dense_groups: presence byte per value in [key_min, key_min + range), scan, lookup:
 *   row_gid_out[i] = group id of row i in OUTPUT order (key ascending, null group last);
 *   group_keys_out (range + 1 entries at most; min(range, n) + 1 suffice) receives the
 *   groups' radix keys in that order, UINT64_MAX for the null group.  group_rows_out (may
 *   be NULL; same length): the number of rows of every group -- the marking pass then counts
 *   rows per value instead of setting a presence byte, and xchg_agg_count_rows is not needed.
 *   range <= XCHG_DENSE_MAX_RANGE.  Synchronises the stream (reads the group count).
 */
int xchg_key_minmax(const uint64_t* norm_keys, const uint8_t* validity, int64_t validity_bit_offset,
                    uint64_t n, void* scratch32, uint64_t* out3_host, void* stream);
int xchg_dense_groups_bytes(uint64_t range, uint64_t* bytes_out);
int xchg_dense_groups(const uint64_t* norm_keys, const uint8_t* validity, int64_t validity_bit_offset,
                      uint64_t n, uint64_t key_min, uint64_t range, void* workspace,
                      uint64_t workspace_bytes, uint32_t* row_gid_out, uint64_t* group_keys_out,
                      uint64_t* group_rows_out, uint64_t* num_groups_out, void* stream);
/*
 * Grouping of SORTED rows, for any key: several key columns, string / binary keys.
 * Replaces the run detection of ArrowBlockAccessor.combine (arrow_block.py:478-497: walk
 * the sorted block, keys_equal with NaN == NaN and null as its own group) and of
 * _get_block_boundaries (python/ray/data/block.py:515-592).  perm = the stable sort of the
 * rows by the key columns (xchg_sort_column chained over the columns, string keys through
 * xchg_string_* keys), or NULL when the rows are stored in key order already (the block of
 * a sorted dataset: GroupedData.map_groups, grouped_data.py:196-213); flags = n bytes,
 * zeroed by the caller.
 *
 * xchg_group_flags_fixed: flags[i] |= sorted row i differs from sorted row i-1 in this
 *   column; norm_sorted = the column's radix keys in SORTED order (xchg_normalize_keys
 *   with perm), validity (may be NULL) indexed by original row id.
 * xchg_group_flags_bytes: the same for a variable-width column (int64 offsets + bytes).
 * xchg_groups_from_flags: scan of the flags -> row_gid_out[original row] = group id in
 *   key order, group_first_out[g] = original row id of group g's first row (the take index
 *   of the output key columns).  perm NULL = identity (rows already sorted).  Synchronises
 *   the stream (reads the group count).
 */
int xchg_group_flags_fixed(const uint64_t* norm_sorted, const uint8_t* validity,
                           int64_t validity_bit_offset, const uint32_t* perm, uint64_t n,
                           uint8_t* flags, void* stream);
int xchg_group_flags_bytes(const int64_t* offsets, const uint8_t* data, const uint8_t* validity,
                           int64_t validity_bit_offset, const uint32_t* perm, uint64_t n,
                           uint8_t* flags, void* stream);
int xchg_groups_from_flags_bytes(uint64_t n, uint64_t* bytes_out);
int xchg_groups_from_flags(const uint8_t* flags, const uint32_t* perm, uint64_t n, void* workspace,
                           uint64_t workspace_bytes, uint32_t* row_gid_out, uint32_t* group_first_out,
                           uint64_t* num_groups_out, void* stream);
/* first_out[g] = smallest row index whose group id is g (row_gid from any of the grouping
 * front ends): the row to take a group's key value from.  Dataset.unique
 * (python/ray/data/dataset.py:2182-2221) is this + a take of the column. */
int xchg_group_first_rows(const uint32_t* row_gid, uint64_t n, uint64_t num_groups, uint32_t* first_out,
                          void* stream);
/* order = permutation that sorts the group list (from xchg_sort_column);
 * row_gid_out[i] = position of row i's group in that order. */
int xchg_group_ids(const uint32_t* order, const uint32_t* group_slots, uint64_t num_groups,
                   const uint32_t* slot, uint64_t n, uint64_t capacity, void* table_ws,
                   uint32_t* row_gid_out, void* stream);
/* stats: 5 * num_groups words, m2 (may be NULL): num_groups doubles. */
int xchg_agg_stats_init(uint64_t* stats, double* m2, uint64_t num_groups, void* stream);
/* rows[g] += number of rows of group g (Count, and the ignore_nulls=False test);
 * row_gid NULL = one global group (groupby key None, aggregate.py:61-63).  With more
 * than 65 536 groups the kernels issue one atomic per row; below that, lanes of a warp
 * that share a group are folded first (contended addresses). */
int xchg_agg_count_rows(const uint32_t* row_gid, uint64_t n, uint64_t num_groups, uint64_t* rows,
                        void* stream);
/* Accumulate one value column (value_type = xchg_key_type of its storage).
 * want_minmax = 0 skips the min / max / NaN-count statistics (Sum, Mean, Std and Count
 * need none of them).  rows (may be NULL): the per-group row counts of the SAME rows
 * (xchg_agg_count_rows); for a column without validity they are its value counts and are
 * added instead of being counted again. */
int xchg_agg_accumulate(const uint32_t* row_gid, const void* values, int value_type,
                        const uint8_t* validity, int64_t validity_bit_offset, uint64_t n,
                        uint64_t num_groups, uint64_t* stats, int want_minmax,
                        const uint64_t* rows, void* stream);
/* Second pass for Std: m2[g] += (x - sum[g]/cnt[g])^2 (after xchg_agg_accumulate). */
int xchg_agg_accumulate_m2(const uint32_t* row_gid, const void* values, int value_type,
                           const uint8_t* validity, int64_t validity_bit_offset, uint64_t n,
                           uint64_t num_groups, const uint64_t* stats, double* m2, void* stream);
/* K9: the same over partial rows (partial_states: n x 6 words). */
int xchg_agg_merge_rows(const uint32_t* row_gid, const int64_t* partial_rows, uint64_t n,
                        uint64_t* rows, void* stream);
int xchg_agg_merge_states(const uint32_t* row_gid, const uint64_t* partial_states, uint64_t n,
                          int sum_is_int, uint64_t num_groups, uint64_t* stats, double* m2,
                          void* stream);
int xchg_agg_pack_states(uint64_t num_groups, const uint64_t* stats, const double* m2,
                         uint64_t* partial_states_out, void* stream);
/* kind: 0 count, 1 sum, 2 min, 3 max, 4 mean, 5 std(ddof), 6 abs_max (AbsMax,
 * python/ray/data/aggregate.py:330-354: the larger magnitude of the group's min and max;
 * NaN counts as a null there, as the reference's row-wise null test does).  out: num_groups
 * 8-byte values (int64 for count and for sum/min/max/abs_max of integer columns, double
 * otherwise); valid_out: one byte per group (0 = null result: no non-null value, or a null
 * with ignore_nulls = 0). */
int xchg_agg_finalize(int kind, int ddof, int ignore_nulls, int value_type, uint64_t num_groups,
                      const uint64_t* rows, const uint64_t* stats, const double* m2, void* out,
                      uint8_t* valid_out, void* stream);
/* Sorted group radix keys -> key column of the result: int64 for integer keys, double
 * for float keys (the reference rebuilds rows from Python scalars). */
int xchg_group_keys_out(const uint64_t* sorted_norm_keys, int key_type, uint64_t num_groups,
                        void* out, void* stream);

/*
 * Exchange-side utilities.  The reference moves map outputs between tasks as
 * Arrow IPC objects through the plasma store
 * (python/ray/_private/arrow_serialization.py:103,
 * pull_based_shuffle_task_scheduler.py:84-120); here the column buffers cross an
 * all-to-all directly, and these three kernels convert what does not split at
 * byte granularity: validity bitmaps travel as one byte per row, offsets travel
 * as the sender's absolute values and are re-based on arrival
 * (out[i] = in[i] - in[0] + base).
 * xchg_bitmap_pack also counts zero bits into *zero_count_out (device uint64,
 * caller zeroes; may be NULL); out must be 4-byte aligned and padded to 4 bytes.
 */
int xchg_bitmap_unpack(const uint8_t* in, int64_t in_bit_offset, uint64_t n, uint8_t* out_bytes,
                       void* stream);
/* out_bytes[i] = bit (idx[i] + in_bit_offset) of `in`: take of a validity / boolean
 * buffer straight into the one-byte-per-row wire form (out_bytes may be peer memory). */
int xchg_gather_bitmap_bytes(const uint8_t* in, int64_t in_bit_offset, const uint32_t* idx,
                             uint64_t n_out, uint8_t* out_bytes, void* stream);
int xchg_bitmap_pack(const uint8_t* in_bytes, uint64_t n, uint8_t* out, uint64_t* zero_count_out,
                     void* stream);
int xchg_offsets_rebase(const int64_t* in, uint64_t count, int64_t base, int64_t* out, void* stream);

/*
 * Caller-owned host memory (the wire bytes of a block as they sit in the object store,
 * python/ray/_private/arrow_serialization.py:785-809): page-lock [ptr, ptr + bytes) so that
 * host->device copies out of it are asynchronous DMA instead of the driver's staged copy of
 * pageable memory.  Pass the buffer's exact range (do not round it out to pages: a neighbour
 * sharing its first or last page must not end up half inside a registered range, copies into
 * it would be refused).  A range that is already registered is
 * not an error (*already_out = 1; the caller must then NOT unregister it).  Read-only mappings
 * are registered read-only.
 */
int xchg_host_register(const void* ptr, uint64_t bytes, int* already_out);
int xchg_host_unregister(const void* ptr);

#ifdef __cplusplus
}
#endif
#endif /* XCHG_B200_H_ */
