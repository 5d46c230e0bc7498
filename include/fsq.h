/*
 * fsq.h — C ABI of libfsq_b200.so, the B200-native PBWT + segmentation path of founder-sequences.
 *
 * The reference (tsnorri/founder-sequences) has no plugin / FFI interface; the entry points below replace
 * the narrowest in-process C++ seams on the hot path.  Each cites the reference interface
 * it stands in for.  Plain pointers and sizes only; no C++ or torch types; never throws, never aborts:
 * every call returns an fsq_status and fsq_last_error() holds the message (the host wrapper turns a
 * non-zero status into libbio_fail(), lib/libbio/src/assert.cc:11-20).
 *
 * All arithmetic is integer and results are bit-identical to the reference's.  There is no CPU
 * fallback: without a CUDA device every compute call fails with FSQ_ERR_CUDA.
 *
 * Threading: a context may be used from any one host thread at a time; it owns its CUDA stream.
 */
#ifndef FSQ_H
#define FSQ_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum fsq_status {
	FSQ_OK = 0,
	FSQ_ERR_ARG = 1,          /* bad argument (null pointer, m == 0, L == 0, more than 15 distinct characters, ...) */
	FSQ_ERR_CUDA = 2,         /* CUDA runtime error, or no device */
	FSQ_ERR_SHORT = 3,        /* n < 2L: the long path does not apply; use fsq_distinct_rows (generate_context.cc:379-390) */
	FSQ_ERR_NOT_REDUCIBLE = 4,/* K == m: the reference exits with failure here (generate_context.cc:192-200) */
	FSQ_ERR_NOMEM = 5,
	FSQ_ERR_UNSUPPORTED = 6,  /* shape outside what the kernels are built for (m > 65535 - block columns, n >= 2^31) */
	FSQ_ERR_INTERNAL = 7
} fsq_status;

typedef struct fsq_ctx fsq_ctx;                     /* one per device / host thread */
typedef struct fsq_msa fsq_msa;                     /* column-major, 4-bit packed alignment resident in HBM */
typedef struct fsq_segmentation fsq_segmentation;   /* result of the long path, device + host copies */
typedef struct fsq_shard fsq_shard;                 /* one rank's part of a column-sharded run of the long path */

/* ---- context ------------------------------------------------------------------------------------------ */

int fsq_create(int device, fsq_ctx **out);
void fsq_destroy(fsq_ctx *ctx);
const char *fsq_last_error(const fsq_ctx *ctx);     /* ctx may be NULL: returns the message of a failed fsq_create */
const char *fsq_version(void);

/* Tuning knobs (all have defaults; unknown keys are an error): "block_columns" (columns per scan block), "cap"
 * (initial bound on the per-column top list; grown automatically when the DP proves it too small), "keep_dp"
 * (retain the full DP table for fsq_segmentation_dp_table), "store_states" (-1 auto, 0 re-scan for the boundary
 * states, 1 keep every column's state in pass B), "dp_chunks", "dp_max_rounds", "dp_guess_rounds" (DP chaining),
 * "kernel", "chunk", "warps", "stitch_sort", "stitch_group" (kernel selection, used by the parity tests). */
int fsq_set_option(fsq_ctx *ctx, const char *key, int64_t value);
int fsq_get_stat(const fsq_ctx *ctx, const char *key, double *value);   /* e.g. "ms_pack", "ms_pass_a", "ms_dp", "kernel_launches" */

/* ---- alignment in HBM --------------------------------------------------------------------------------- */

/* Replaces the row access `inputs[vec_idx][column_idx]` + `alphabet.char_to_comp` of
 * libbio::pbwt::build_prefix_and_divergence_arrays (lib/libbio/include/libbio/pbwt.hh:93-95,119-121) and the
 * sequence_vector of spans (include/founder_sequences/founder_sequences.hh:38-40).
 * rows: m host pointers to n bytes each.  comp: the reference alphabet's 256-entry char→comp table
 * (consecutive_alphabet.hh:259-264), or NULL to have it built here the same way (NUL = 0, bytes present get
 * 1.. in byte order).  Copies host→device and transposes/packs on the device. */
int fsq_msa_from_rows(fsq_ctx *ctx, const uint8_t *const *rows, uint32_t m, uint64_t n, const uint8_t comp[256], fsq_msa **out);

/* Same, rows already on the device as a row-major m × n byte matrix with the given row stride; comp must be given. */
int fsq_msa_from_device_rows(fsq_ctx *ctx, const uint8_t *d_rows, uint64_t row_stride, uint32_t m, uint64_t n, const uint8_t comp[256], fsq_msa **out);

/* Founder-mosaic synthetic alignment generated on the device, columns [col0, col0 + n) of the matrix defined by
 * (seed, founders, gaps) — see founder-sequences_b200/synth.py for the definition.  Used by bench.py and the tests. */
int fsq_msa_synth(fsq_ctx *ctx, uint32_t m, uint64_t n, uint64_t col0, uint64_t seed, uint32_t founders, int gaps, fsq_msa **out);
/* The same cells as ASCII rows (row-major m × n) into a device buffer, for end-to-end runs that start from host rows. */
int fsq_synth_rows_device(fsq_ctx *ctx, uint8_t *d_rows, uint64_t row_stride, uint32_t m, uint64_t n, uint64_t col0, uint64_t seed, uint32_t founders, int gaps);

/* Must run before fsq_destroy() of the context that made the alignment: packed alignments of up to 4 GB return to the
 * context's buffer cache (reused by the next upload on that context, released by fsq_destroy or when memory runs short). */
void fsq_msa_free(fsq_msa *msa);
uint32_t fsq_msa_rows(const fsq_msa *msa);
uint64_t fsq_msa_columns(const fsq_msa *msa);

/* ---- PBWT state tap (a1–a3) ---------------------------------------------------------------------------- */

/* (a_c, d_c) — the input permutation / divergence *before* column c, i.e. after columns [0, c) — for each of
 * the `count` ascending positions in `columns` (1 <= c <= n).  Replaces running
 * libbio::pbwt::pbwt_context::process() to column c and reading input_permutation()/input_divergence()
 * (lib/libbio/include/libbio/pbwt_context.hh:184-243, 439-473).  a, d: count × m. */
int fsq_pbwt_states(fsq_ctx *ctx, const fsq_msa *msa, const uint64_t *columns, uint32_t count, uint32_t *a, uint32_t *d);

/* ---- segmentation, long path (seam S1) ----------------------------------------------------------------- */

/* Replaces segmentation_lp_context::generate_traceback → update_samples_to_traceback_positions →
 * find_segments_greedy (founder-sequences/segmentation_lp_context.cc:26-390), whose result the reference hands
 * over as a segmentation_container (include/founder_sequences/segmentation_container.hh:15-20).
 * Blocking.  Returns FSQ_ERR_SHORT when n < 2L and FSQ_ERR_NOT_REDUCIBLE when K == m. */
int fsq_segment(fsq_ctx *ctx, const fsq_msa *msa, uint64_t segment_length_bound, fsq_segmentation **out);
void fsq_segmentation_free(fsq_segmentation *seg);

/* container.reduced_traceback.size() and container.max_segment_size */
uint32_t fsq_segmentation_count(const fsq_segmentation *seg);
uint32_t fsq_segmentation_max_segment_size(const fsq_segmentation *seg);
/* reduced_traceback[i].{lb, rb, segment_size}; arrays of fsq_segmentation_count() entries */
int fsq_segmentation_bounds(const fsq_segmentation *seg, uint64_t *lb, uint64_t *rb, uint32_t *size);
/* reduced_pbwt_samples[i].input_permutation() / input_divergence(); count × m each */
int fsq_segmentation_states(const fsq_segmentation *seg, uint32_t *a, uint32_t *d);
/* The same without the extra copy: pointers into pinned host memory owned by the handle (valid until it is freed).
 * The device->host copy happens on the first call of either function. */
int fsq_segmentation_states_view(const fsq_segmentation *seg, const uint32_t **a, const uint32_t **d);

/* Test taps.  The traceback before reduction (segmentation_lp_context.cc:191-224): count, then lb/rb/max/size. */
uint32_t fsq_segmentation_traceback_count(const fsq_segmentation *seg);
int fsq_segmentation_traceback(const fsq_segmentation *seg, uint64_t *lb, uint64_t *rb, uint32_t *max_size, uint32_t *size);
/* The whole DP table, n - L + 1 entries (needs option keep_dp = 1; entries the reference never writes hold its
 * defaults {0, 0, UINT32_MAX, UINT32_MAX}). */
int fsq_segmentation_dp_table(const fsq_segmentation *seg, uint64_t *lb, uint64_t *rb, uint32_t *max_size, uint32_t *size);

/* ---- short path (seam S1') ------------------------------------------------------------------------------ */

/* Replaces segmentation_sp_context::process (founder-sequences/segmentation_sp_context.cc:21-28):
 * distinct rows in PBWT order after all n columns, as (first row of the run, run length).  Arrays of m entries. */
int fsq_distinct_rows(fsq_ctx *ctx, const fsq_msa *msa, uint32_t *row_of_run, uint32_t *run_len, uint32_t *count);

/* ---- segment texts (seam S2) and edge weights (seam S3) ---------------------------------------------------- */

/* Replaces pbwt_context::unique_substring_count_idxs_lhs (lib/libbio/include/libbio/pbwt.hh:253-280, called at
 * founder-sequences/join_context.cc:73) and the gather + per-set sort of create_segment_texts_task::execute
 * (founder-sequences/create_segment_texts_task.cc:15-35) for reduced segment `segment`:
 * runs in PBWT order as CSR — run_off[n_runs + 1] offsets into ids[m], row ids ascending inside a run —
 * plus class_of_row[m] (the run each row belongs to).  The size-sort / padding to K slots (:38-82, std::sort and
 * double ceil) stays on the host. */
int fsq_segment_sets(fsq_ctx *ctx, const fsq_segmentation *seg, uint32_t segment, uint32_t *run_off, uint32_t *ids, uint32_t *class_of_row, uint32_t *n_runs);

/* Replaces the K² set_intersection_size calls of merge_segments_task::execute
 * (founder-sequences/merge_segments_task.cc:133-176; lib/libbio/include/libbio/algorithm/set_count.hh:61-89) for the
 * adjacent reduced segments (segment, segment + 1): W[i * runs_rhs + j] = |S_i ∩ T_j| over the runs in PBWT order. */
int fsq_pair_weights(fsq_ctx *ctx, const fsq_segmentation *seg, uint32_t segment, int32_t *W, uint32_t *runs_lhs, uint32_t *runs_rhs);

/* ---- column-sharded operation (one process per GPU; the exchange itself is done by the caller over NCCL) ----------
 *
 * The reference runs the path in one process (no counterpart); these are the stages of fsq_segment() cut
 * where column shards exchange data.  Rank r holds the columns [c0_r, c1_r) of ONE alignment of n_total columns as an
 * fsq_msa whose first column is c0_r (fsq_msa_synth with col0, or fsq_msa_set_first_column).  Order of calls, with the
 * caller's collectives in brackets (founder-sequences_b200/sharded.py is the reference driver):
 *
 *   fsq_shard_local_summary      local scan from the identity state + prefix composition -> this rank's (pi, delta)
 *   [all-gather of the summaries]
 *   fsq_shard_set_start          true state at c0_r = composition of the summaries of the ranks before; block starts
 *   fsq_shard_scan               exact scan: top lists (and kept states) for the own columns; resets the DP table
 *   repeat: fsq_shard_dp_round   own DP entries against the table as it stands
 *           [broadcast of every rank's slice of tb_max; max-reduce of changed / reach]
 *           fsq_shard_dp_sync    level guess (optional) + masks and rmq tables of the whole table
 *   until no rank changed anything
 *   fsq_shard_dp_final           last entry (rank with the last column); [broadcast of the tb_max/tb_lb/tb_size slices;
 *                                max-reduce of the overflow flag -> fsq_shard_grow_cap and back to fsq_shard_scan]
 *   fsq_shard_traceback          every rank, from the complete table
 *   fsq_shard_reduce_lists       rows of the reduction's count lists for the own traceback positions
 *   [broadcast of every rank's rows]
 *   fsq_shard_reduce             greedy reduction (every rank) + the states of the own reduced segments
 * Results are independent of the number of ranks and of the block size (tests/test_sharded*.py). */
int fsq_msa_set_first_column(fsq_msa *msa, uint64_t first_column);
int fsq_shard_create(fsq_ctx *ctx, const fsq_msa *msa, uint64_t n_total, uint64_t segment_length_bound, fsq_shard **out);
void fsq_shard_free(fsq_shard *s);
/* Device pointers (m uint32 each, owned by the shard): the state after this rank's last column had the scan started
 * from the identity state at its first column. */
int fsq_shard_local_summary(fsq_shard *s, const uint32_t **d_pi, const uint32_t **d_delta);
/* d_pis / d_deltas: device arrays [count][m] with the summaries of the ranks before this one, in rank order;
 * first_columns[count]: their first columns (host).  count == 0 for the rank that starts at column 0. */
int fsq_shard_set_start(fsq_shard *s, const uint32_t *d_pis, const uint32_t *d_deltas, const uint64_t *first_columns, uint32_t count);
int fsq_shard_scan(fsq_shard *s);
/* The DP table (device, `entries` uint32 each; entry t describes right bound t + L) and the range of entries this rank
 * computes, i.e. the slice it must send to the others. */
int fsq_shard_dp_table(fsq_shard *s, uint32_t **d_max, uint32_t **d_lb, uint32_t **d_size, uint64_t *entries, uint64_t *own_first, uint64_t *own_last);
int fsq_shard_dp_round(fsq_shard *s, uint32_t *changed, uint32_t *reach);
int fsq_shard_dp_sync(fsq_shard *s, int guess, uint32_t reach);
int fsq_shard_dp_single_chain(fsq_shard *s);       /* fallback: the next rounds run the own entries as one chain */
int fsq_shard_dp_final(fsq_shard *s, uint32_t *overflow);
int fsq_shard_grow_cap(fsq_shard *s);
int fsq_shard_traceback(fsq_shard *s, uint32_t *count, uint32_t *max_segment_size);
int fsq_shard_get_traceback(const fsq_shard *s, uint64_t *lb, uint64_t *rb, uint32_t *max_size, uint32_t *size);
/* The traceback positions whose right bound lies in (c0, c1]: the rows of the reduction lists that the rank holding
 * columns [c0, c1) fills.  The traceback is replicated, so every rank can answer this for every other rank. */
int fsq_shard_position_range(const fsq_shard *s, uint64_t c0, uint64_t c1, uint32_t *first, uint32_t *last);
/* Device array [count][width] of (v, #{d > v}) pairs (2 uint32 each) and the rows [own_first, own_last) this rank filled. */
int fsq_shard_reduce_lists(fsq_shard *s, uint32_t **d_lists, uint32_t *width, uint32_t *own_first, uint32_t *own_last);
/* Reduced segments: their number and the range [own_first, own_last) whose states this rank holds. */
int fsq_shard_reduce(fsq_shard *s, uint32_t *count, uint32_t *own_first, uint32_t *own_last);
int fsq_shard_bounds(const fsq_shard *s, uint64_t *lb, uint64_t *rb, uint32_t *size);
int fsq_shard_states(fsq_shard *s, uint32_t *a, uint32_t *d);      /* (own_last - own_first) x m each, to host memory */

#ifdef __cplusplus
}
#endif
#endif /* FSQ_H */
