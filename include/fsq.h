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

/* ---- context ------------------------------------------------------------------------------------------ */

int fsq_create(int device, fsq_ctx **out);
void fsq_destroy(fsq_ctx *ctx);
const char *fsq_last_error(const fsq_ctx *ctx);     /* ctx may be NULL: returns the message of a failed fsq_create */
const char *fsq_version(void);

/* Tuning knobs (all have defaults; unknown keys are an error): "block_columns" (columns per scan block), "cap"
 * (initial bound on the per-column top list; grown automatically, to what the DP proves sufficient, when it is too
 * small; the bound a shape (m, n, L) ended with is remembered by the context, the option is never overwritten), "keep_dp"
 * (retain the full DP table for fsq_segmentation_dp_table), "store_states" (-1 auto, 0 re-scan for the boundary
 * states, 1 keep every column's state in pass B), "dp_chunks", "dp_max_rounds", "dp_guess_rounds" (DP chaining),
 * "kernel", "chunk", "warps", "stitch_sort" (kernel selection, used by the parity tests), "mem_limit"
 * (plan device memory as if this many bytes were available), "window_blocks" (scan blocks per window of top lists:
 * the exact scan and the DP then alternate over windows that share one list buffer), "halo" (rows of the ranks behind
 * it that a rank's greedy reduction starts with), "trace_window" (entries per rank exported as places where the traceback
 * path may enter it; a longer jump falls back to the traceback over the gathered table), "dp_back_bound" (entries below
 * a rank's first that its DP may look back at, 65,536: the ranks exchange this much of their slices per round, and a
 * longer look-back falls back to whole slices), "exchange_tails" (0: whole slices from the start), "timeline" (1: the
 * statistics "tl_count", "tl_id_<i>", "tl_at_<i>", "tl_end_<i>" list every timed phase of the last call in order). */
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

/* Replaces the column test of remove_identity_columns (remove-identity-columns/main.cc:107-160), the pre-processing tool
 * that drops the columns in which every row carries the same character: identity[i] = 1 for such a column i (n bytes, host;
 * may be NULL), *n_identity their number (may be NULL), and -- when `without` is given -- the alignment that is left, in
 * HBM, ready for fsq_segment (FSQ_ERR_ARG when nothing is left).  The tool's other half, re-inserting the columns into the
 * founders (insert-identity-columns), is host text processing driven by the same 0/1 mask. */
int fsq_identity_columns(fsq_ctx *ctx, const fsq_msa *msa, uint8_t *identity, uint64_t *n_identity, fsq_msa **without);

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
/* A slice of them, straight from the device: the states of segments [first, first + count), count x m each (hundreds of
 * thousands of reduced segments times thousands of rows do not have to fit the host at once). */
int fsq_segmentation_states_range(const fsq_segmentation *seg, uint32_t first, uint32_t count, uint32_t *a, uint32_t *d);
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
 * adjacent reduced segments (segment, segment + 1): W[i * runs_rhs + j] = |S_i ∩ T_j| over the runs in PBWT order.
 * W must hold runs_lhs * runs_rhs ints; a reduced segment has at most max_segment_size runs, so
 * fsq_segmentation_max_segment_size()^2 ints always suffice. */
int fsq_pair_weights(fsq_ctx *ctx, const fsq_segmentation *seg, uint32_t segment, int32_t *W, uint32_t *runs_lhs, uint32_t *runs_rhs);

/* Replaces the pair counting of greedy_matcher::match (founder-sequences/greedy_matcher.cc:293-343: every row's
 * (class in the left segment, class in the right segment) encoded, radix-sorted, runs counted) for the adjacent
 * reduced segments (segment, segment + 1): the distinct pairs in ascending (lhs, rhs) order with their counts -- the
 * non-zero cells of fsq_pair_weights' table in row-major order.  lhs, rhs, count: m entries each (there are at most m
 * distinct pairs).  What the matcher does with them (index_pairs_by_count, the edge drawing, :345-437) is a short
 * sequential loop over at most min(m, K^2) items and stays on the host. */
int fsq_pair_counts(fsq_ctx *ctx, const fsq_segmentation *seg, uint32_t segment, uint32_t *lhs, uint32_t *rhs, uint32_t *count, uint32_t *n_pairs, uint32_t *runs_lhs, uint32_t *runs_rhs);

/* Replaces match_context::match_sequence_and_report of the evaluation tool match_founder_sequences
 * (match-sequences-to-founders/match_founder_sequences.cc:127-214): every row of `sequences` is walked left to right
 * against the rows of `founders` (both packed with the SAME char -> comp table; equal lengths), keeping the founders
 * that agree with it since the current range began; when none is left -- or, with min_segment_length != 0, when the range
 * has that length (:163-165) -- the range is reported with the set as it stood and the next begins.  Output: the ranges
 * in the order the tool prints them single-threaded (by sequence, then left bound): sequence index, [lb, rb), and the
 * founder set as ceil(K / 32) words (bit f of word f / 32).  *count receives the number of ranges; nothing is written
 * when the arrays are NULL or `capacity` is smaller (call again with room).  What the tool writes to stderr (characters
 * no founder has) shows here as ranges with an empty set, as on its stdout. */
int fsq_match_founders(fsq_ctx *ctx, const fsq_msa *sequences, const fsq_msa *founders, uint64_t min_segment_length,
	uint32_t *seq_idx, uint64_t *lb, uint64_t *rb, uint32_t *founder_bits, uint64_t capacity, uint64_t *count);

/* No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The counting step of
 * fsq_pair_counts on the CALLER'S device memory, with nothing allocated: the contingency table it goes through is a
 * static device array of FSQ_PAIR_COUNTS_NOALLOC_MAX_CELLS ints (4 MiB + 256 bytes, held by every GPU that loads the
 * library until the process ends).  On the current device, on `stream` (a cudaStream_t; NULL = the default stream),
 * without waiting for it:
 *   class_lhs, class_rhs [m]   device; row -> its class (run) in the left / right segment, as fsq_segment_sets'
 *                              class_of_row gives them; values < runs_lhs / runs_rhs
 *   lhs, rhs, count            device; min(m, runs_lhs * runs_rhs) entries each: the distinct (lhs, rhs) pairs in ascending
 *                              order with their counts, exactly fsq_pair_counts' output
 *   n_pairs                    device; one word: the number of pairs, 0xffffffff when a class index was out of range
 *                              (nothing else is then written)
 * runs_lhs * runs_rhs > FSQ_PAIR_COUNTS_NOALLOC_MAX_CELLS is refused with FSQ_ERR_UNSUPPORTED (use fsq_pair_counts).
 * Takes no context: a message for a non-zero return goes to `error` (up to error_size bytes, may be NULL).
 * The first call on a device looks the table's address up (cudaGetSymbolAddress): make it outside a stream capture.
 * All calls on a device share the table, cleared at the start of each call: calls that overlap on the GPU race. */
#define FSQ_PAIR_COUNTS_NOALLOC_MAX_CELLS (1u << 20)
int fsq_pair_counts_noalloc(const uint32_t *class_lhs, const uint32_t *class_rhs, uint32_t m, uint32_t runs_lhs, uint32_t runs_rhs,
	uint32_t *lhs, uint32_t *rhs, uint32_t *count, uint32_t *n_pairs, void *stream, char *error, size_t error_size);

/* ---- column-sharded operation (one process -- or thread -- per GPU) -----------------------------------------------
 *
 * The reference runs the path in one process (no counterpart).  Rank r of `size` holds the columns
 * fsq_shard_columns() names of ONE alignment of n_total columns, as an fsq_msa whose first column is set accordingly
 * (fsq_msa_synth with col0, or fsq_msa_set_first_column after an upload), and every rank calls fsq_segment_sharded()
 * with its communicator.  The library does the exchanges itself, as all-gathers enqueued on the context's stream
 * between its kernels (ncclAllGather over NVLink / NVSwitch): the ranks' boundary summaries (8 m bytes each), per DP
 * round the ranks' slices of the table (4 bytes per column, in place) and four control words, the rest of the table
 * once, and for the greedy reduction a few rows per rank.  No host thread takes part between the first and the last
 * stage except to read the control words once per batch of stages.
 * Results are independent of the number of ranks and of the block size (tests/test_sharded*.py). */
typedef struct fsq_comm fsq_comm;
typedef struct fsq_local_group fsq_local_group;
#define FSQ_COMM_ID_BYTES 128

/* The columns [*c0, *c1) rank `rank` of `size` holds: equal shares of ceil(n_total / size) rounded up to 32 columns
 * (the last rank holds what is left).  FSQ_ERR_ARG when a rank would hold no column. */
int fsq_shard_columns(uint64_t n_total, int size, int rank, uint64_t *c0, uint64_t *c1);

/* NCCL communicator, one rank per GPU.  Rank 0 makes the id (an ncclUniqueId) and hands it to the others by whatever
 * means the host has (MPI, torch.distributed, a file); then every rank calls fsq_comm_create, which blocks until all have
 * (it ends with one small all-gather, so that NCCL makes its connections here and not inside the first run).
 * libnccl.so.2 is opened at run time; a process that has it loaded already (PyTorch) shares that copy. */
int fsq_comm_unique_id(uint8_t id[FSQ_COMM_ID_BYTES]);
int fsq_comm_create(fsq_ctx *ctx, const uint8_t id[FSQ_COMM_ID_BYTES], int rank, int size, fsq_comm **out);
/* The same over a communicator the host made itself (an ncclComm_t); it stays the host's. */
int fsq_comm_adopt_nccl(fsq_ctx *ctx, void *nccl_comm, int rank, int size, fsq_comm **out);
/* Ranks as threads of one process, each with its own context, possibly all on one device (how the tests drive several
 * shards through one GPU): the collective is device copies ordered by events, the host threads meet at a barrier. */
int fsq_local_group_create(int size, fsq_local_group **out);
void fsq_local_group_free(fsq_local_group *group);
int fsq_comm_create_local(fsq_ctx *ctx, fsq_local_group *group, int rank, fsq_comm **out);
void fsq_comm_free(fsq_comm *comm);
int fsq_comm_rank(const fsq_comm *comm);
int fsq_comm_size(const fsq_comm *comm);

/* For alignments uploaded shard by shard: the global index of the shard's first column. */
int fsq_msa_set_first_column(fsq_msa *msa, uint64_t first_column);

/* fsq_segment() over column shards; collective: every rank of `comm` calls it with its shard, the same n_total and the
 * same bound.  Every rank receives the whole reduced segmentation (count, bounds, K, traceback); the states it holds
 * are those of the reduced segments whose right bound lies in its columns, fsq_segmentation_own_range().
 * comm == NULL (or a communicator of one rank) is fsq_segment(). */
int fsq_segment_sharded(fsq_ctx *ctx, const fsq_msa *msa, fsq_comm *comm, uint64_t n_total, uint64_t segment_length_bound, fsq_segmentation **out);
/* The reduced segments [*first, *last) whose states fsq_segmentation_states() returns ((last - first) x m each) and
 * which fsq_segment_sets() / fsq_pair_weights() accept; all of them after fsq_segment(). */
int fsq_segmentation_own_range(const fsq_segmentation *seg, uint32_t *first, uint32_t *last);

#ifdef __cplusplus
}
#endif
#endif /* FSQ_H */
