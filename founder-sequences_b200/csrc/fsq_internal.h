// Internal declarations shared by the translation units of libfsq_b200.so.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>

#ifndef FSQ_MAX_SYMS
#define FSQ_MAX_SYMS 16            // 4-bit packed symbols
#endif

// The chunked column step (fsq_cstep.cuh) in its paired form: divergence codes take part in f16x2 arithmetic as bit
// patterns, so they stay below the first non-finite half (0x7C00).
#ifndef FSQ_STEP_HALF2
#define FSQ_STEP_HALF2 1
#endif

namespace fsq {

constexpr uint32_t FSQ_PAD_CODE = FSQ_STEP_HALF2 ? 0x7bffu : 0xffffu;      // the largest code; padding rows carry it
constexpr uint32_t FSQ_CODE_LIMIT = FSQ_STEP_HALF2 ? 0x7bfeu : 0xfffeu;    // codes in use stay below it (chunked step)

enum { MODE_LOCAL = 0, MODE_EXACT = 1, MODE_DUMP = 2 };

// One launch of the column-block scan kernel.  Local column index k <-> global position col0 + k.
struct block_args
{
	uint8_t const *cols;       // packed alignment, column-major, col_stride bytes per column
	uint64_t col_stride;
	uint64_t n_local;          // columns held
	uint64_t col0;             // global index of local column 0
	uint32_t m, n_tiles, nsym;
	uint32_t sym0;             // smallest symbol value that occurs (1 when NUL is absent); blocked kernels use nsym - sym0 symbols
	uint32_t C;                // columns per block
	uint32_t cg;               // columns per TMA stage
	uint32_t first_block;
	uint32_t magic;            // 0x4B000000, passed as an argument so that the step's PRMT reads it from the constant bank
	uint32_t digits;           // chunked step: 1, or 2 = a column is two steps over the base-4 digits of its symbols (alphabets of 9 .. 16 comps)
	int mode;

	// start states (MODE_EXACT, MODE_DUMP)
	uint32_t const *st_a;      // [B][m]
	uint16_t const *code;      // [B][m] rank-compressed divergences
	uint32_t const *oldvals;   // [B][m] code -> value
	uint32_t const *R;         // [B] number of distinct start values

	// MODE_LOCAL output: state at the block end, started from identity
	uint32_t *sum_a, *sum_d;   // [B][m]

	// MODE_EXACT output, per local column
	uint2 *toplist;            // [(cap + 1) entries per column] (v, #{d > v}), v descending; row 0 belongs to local column list_col0
	uint64_t list_col0;
	uint32_t cap;
	uint32_t *cnt;             // entries | complete << 31
	uint32_t *h0;              // #{d == 0}

	uint32_t *states;          // [n_local][state_stride] packed state after every column (optional; chunked kernel only)
	uint64_t state_stride;     // words

	// MODE_DUMP
	uint32_t const *blk_ids;
	uint64_t const *dump_cols; // ascending local right bounds
	uint32_t const *dump_begin;// [B + 1]
	uint32_t *dump_a, *dump_d; // [n_dumps][m]
};

size_t block_smem_bytes(block_args const &A, uint32_t n_warps, bool hist);
cudaError_t launch_block(block_args const &A, uint32_t n_blocks, uint32_t n_warps, uint32_t tpw, bool hist, cudaStream_t st);
// Chunked column step (fsq_cstep.cuh): T threads, E elements per thread, T E >= m; 32-bit histogram counters.
inline size_t cblock_smem_bytes(block_args const &A, uint32_t T, uint32_t E, uint32_t NS, bool hist)
{
	auto r16 = [](size_t b) { return (b + 15) & ~size_t(15); };
	uint32_t const n_warps = T >> 5;
	size_t const words = (size_t) T * E;
	size_t t = 0;
	t += r16(2 * (size_t) A.cg * A.col_stride);
	t += r16(2 * sizeof(uint64_t));
	t += r16((size_t) A.cg * (A.digits > 1 ? 2 : 1) * ((words + 15) & ~size_t(15)));
	t += r16(sizeof(uint32_t) * words);
	t += r16(sizeof(uint32_t) * (size_t) T * ((((6 * (NS - 1)) & ~3u) / 4 + 1) | 1));      // cstep_nsp(NS), fsq_cstep.cuh
	t += r16(sizeof(uint32_t) * n_warps * 2 * NS);
	if (hist) t += r16(sizeof(uint32_t) * ((size_t) A.m + A.C + 8));
	return t;
}
// One translation unit per chunk length (fsq_cscan_e*.cu).
#define FSQ_CHUNK_LENGTHS(X) X(9) X(13) X(17) X(21) X(25) X(27) X(29) X(31) X(33) X(53)
#define FSQ_DECLARE_CBLOCK(E) cudaError_t launch_cblock_e##E(block_args const &A, uint32_t n_blocks, uint32_t T, uint32_t ns, bool hist, cudaStream_t st);
FSQ_CHUNK_LENGTHS(FSQ_DECLARE_CBLOCK)
#undef FSQ_DECLARE_CBLOCK
// Threads per CTA are bounded so that the unrolled chunk (2 E registers and the running maxima) fits the register file:
// 64 registers per thread up to E = 9, 128 up to 33, 168 at 53 (which reaches 20,352 rows).
constexpr uint32_t cblock_thread_limit(uint32_t E) { return E <= 9 ? 1024 : (E <= 33 ? 512 : 384); }
constexpr uint32_t cblock_registers(uint32_t E) { return E <= 9 ? 64 : (E <= 33 ? 128 : 168); }
cudaError_t launch_prep(uint32_t const *st_d, uint32_t m, uint16_t *code, uint32_t *oldvals, uint32_t *R, uint32_t first_block, uint32_t n_blocks, cudaStream_t st);

struct stitch_item
{
	uint32_t const *in_a, *in_d;   // incoming state at s
	uint32_t const *pi, *delta;    // local summary of [s, e)
	uint32_t *out_a, *out_d;       // state at e
	uint32_t s_mark;               // delta <= s_mark: identical to the predecessor over [s, e)
	uint32_t identity_in;          // incoming state is (identity, identity_d)
	uint32_t identity_d;
	uint32_t pad;
};

struct stitch_args
{
	stitch_item const *items;      // device array; CTA g executes items [item_begin[g], item_begin[g + 1]) in order
	uint32_t const *item_begin;    // device array [n_ctas + 1]
	uint32_t n_ctas;
	uint32_t m, n_tiles;
	uint32_t use_sort;             // m <= 1024: the one-sort formulation (stitch_sort_kernel)
};

size_t stitch_smem_bytes(uint32_t m, uint32_t n_warps);
cudaError_t launch_stitch(stitch_args const &A, cudaStream_t st);

// ---- DP ------------------------------------------------------------------------------------------------------
// Control words of a run, in device memory (uint32[16]): the stages are enqueued back to back and read their decisions
// from here; the host reads the block once per batch of stages.
enum
{
	F_CHANGED = 0,      // some stored value changed in this round
	F_OVERFLOW = 1,     // max over the entries whose top list was too short to prove them exact of the value found (0 = none):
	                    // every value computed under a short list is an upper bound of the true one, so lists that reach this far suffice
	F_REACH = 2,        // the farthest any column looked back in the table (entries), over all rounds
	F_UNSETTLED = 3,    // chains of this round whose result a later chain may have read half-way (see dp_chain_kernel)
	F_BOUND = 4,        // F_REACH as it stood when the last guess ran: the look-back bound of the round queued behind it
	F_DONE = 5,         // a round found the table at its fixed point: rounds queued behind it do nothing
	F_ROUNDS = 6,       // rounds that did work
	F_WINDOW_ERR = 7,   // the look-back left the part of the table whose masks and rmq cells are rebuilt between rounds
	I_T = 8,            // traceback entries
	I_K = 9,            // maximum segment size
	I_NRED = 10,        // reduced segments
	I_SEG_FIRST = 11, I_SEG_LAST = 12,      // reduced segments whose state this rank holds
	I_POS_FIRST = 13, I_POS_LAST = 14,      // traceback positions whose right bound lies in this rank's columns
	I_HALO_ERR = 15,    // the greedy reduction ran past the rows it had of the ranks after this one
	I_TRACE_ERR = 16,   // column shards: the path entered a rank below the entries it had exported (a jump longer than the window)
	F_BLIND = 17,       // entries whose top list was too short to hold a single admissible candidate: nothing is known of them but "> cap"
	F_SURE = 18,        // a list proved too short in a chain whose entries are final already (chain c is exact from round c + 1 on)
	N_CONTROL = 20
};

struct dp_args
{
	uint32_t m;
	uint64_t n;                 // columns of the whole alignment
	uint64_t L;
	uint64_t col_base;          // global index of the column cnt[0] / h0[0] belong to (this rank's first column)
	uint64_t list_base;         // global index of the column whose top list is row 0 of toplist (a window of the rank's columns)
	uint32_t t_first, t_last;   // the entries computed by this launch: [t_first, t_last)
	uint2 const *toplist; uint32_t cap; uint32_t const *cnt; uint32_t const *h0;
	uint32_t *tb_max, *tb_lb, *tb_size;   // [n - L + 1]; entry t describes right bound t + L
	uint64_t *mk;               // [n - L + 1] in-block minima masks (see fsq_dp.cu)
	uint32_t *lev;              // [n_levels][n_blk] rmq tables
	uint32_t n_blk, n_levels;
	uint32_t n_entries;         // entries 0 .. n - 2L (the ones the per-column DP writes)
	uint32_t chunk;             // entries per chain (multiple of 64); chain c owns entries [c chunk, (c + 1) chunk)
	uint32_t n_chunks;
	uint32_t *chunk_level;      // [2 n_chunks] scratch for the guess
	uint32_t *flags;            // the control words above
	uint32_t reach_bound;       // the look-back known when the round starts (0 in the first round)
	uint32_t bound_on_device;   // take the bound from flags[F_BOUND] instead (see dp_chain_kernel)
	uint32_t skip_when_done;    // return at once when flags[F_DONE] is set
	uint32_t round_no;          // 1-based; chains chain_base .. chain_base + round_no - 1 (global index) compute final values:
	uint32_t chain_base;        // everything before chain chain_base is final when the rounds start
	uint32_t w_first, w_last;   // masks and rmq cells are rebuilt for entries [w_first, w_last) (w_first a multiple of 64)
	uint32_t lvl_first, lvl_last;   // the guess levels entries [lvl_first, lvl_last) (chains on the global grid of `chunk` entries)
	// Column shards that exchange only what the rank behind reads (the last entries of every slice): a rank then sees the
	// ends of its own chains only, computes their window minima at the end of a round (launch_dp_chunk_minima), and the
	// guess takes every chain's from the gathered copies: lvl_gathered[lvl_ranks][n_chunks], zero where a rank owns nothing.
	uint32_t own_first, own_last;   // chain ends (end - 1) in [own_first, own_last) are this rank's; 0 .. n_entries on one rank
	uint32_t const *lvl_gathered; uint32_t lvl_ranks;
};

cudaError_t launch_dp_chains(dp_args const &A, cudaStream_t st);       // every chain once, against the table as it stands
cudaError_t launch_dp_chunk_minima(dp_args const &A, cudaStream_t st); // chunk_level[n_chunks + c] = the window minimum at the end of chain c (own chains; 0 elsewhere)
// The tails the ranks sent (gathered[size][tail] entries: the last `tail` of each slice of n_per) into the table copy of this rank.
cudaError_t launch_dp_scatter_tails(uint32_t const *gathered, uint32_t size, uint32_t rank, uint32_t tail, uint64_t n_per, uint32_t *tbx_max, cudaStream_t st);
cudaError_t launch_dp_guess(dp_args const &A, cudaStream_t st);        // between rounds: level the table, rebuild masks and rmq tables
cudaError_t launch_dp_final(dp_args const &A, cudaStream_t st);        // the last column (entry n - L)
cudaError_t launch_dp_rebuild(dp_args const &A, cudaStream_t st);      // masks and rmq tables from the table as it stands
// End of a round on one rank: send[0..3] = {another round is needed, F_OVERFLOW, F_REACH, F_BLIND | F_SURE << 31}.
cudaError_t launch_dp_round_end(uint32_t *flags, uint32_t first_round, uint32_t *send, cudaStream_t st);
// ... and with every rank's words at hand (gathered[size][4]): the merged decision into flags; clears the per-round words.
cudaError_t launch_dp_round_merge(uint32_t *flags, uint32_t const *gathered, uint32_t size, uint32_t round_no, uint32_t back_bound, uint32_t one_chain, cudaStream_t st);
cudaError_t launch_dp_overflow_merge(uint32_t *flags, uint32_t const *gathered, uint32_t size, cudaStream_t st);   // F_OVERFLOW = max over the ranks

// Traceback (reference: segmentation_lp_context.cc:191-224) by pointer doubling: jump[k][t] = the entry 2^k hops
// before t on the path t -> lb[t] - L.  The walk from the last entry is then T independent binary descents.
// The number of entries stays on the device (ctl[I_T]); arrays are sized by the bound dp_size / L + 2.
struct trace_args
{
	uint32_t const *tb_lb, *tb_max, *tb_size;
	uint32_t L, dp_size, levels;
	uint32_t *jump;              // [levels][dp_size]
	uint32_t *ctl;               // control words: I_T, I_K, I_POS_FIRST, I_POS_LAST are written here
	uint32_t t_bound;            // capacity of the out arrays
	uint32_t *out_lb, *out_max, *out_size;   // [t_bound] in text order
	uint64_t *out_rb;
	uint64_t c0, c1;             // this rank's columns
};
cudaError_t launch_trace_tables(trace_args const &A, cudaStream_t st);     // jump tables + count + K
cudaError_t launch_trace_gather(trace_args const &A, cudaStream_t st);     // the path in text order + this rank's positions

// The same for column shards, without anyone holding the whole table: every rank builds jump tables over its own entries
// only (a jump that leaves them is an exit: the entry of some rank before), exports (exit, entries on the path) for the
// entries a path can enter it at -- the last `window` ones, a jump being no longer than the DP's farthest look-back --
// and after one all-gather of those exports resolves the chain last rank -> ... -> first: where the path enters each
// rank and how many of its entries lie there.  Each rank then lists its own part of the path.
struct ltrace_args
{
	uint32_t const *tb_lb, *tb_max, *tb_size;      // the table (entry t at [t]); only own entries are read
	uint32_t L, dp_size, levels;
	uint32_t own_first, own_last;                  // own per-column entries [own_first, own_last)
	uint32_t has_final;                            // this rank also owns entry dp_size - 1
	uint32_t own_n;                                // own_last - own_first + has_final: rows of the tables (the final entry is the last row)
	uint32_t *jump, *hops;                         // [levels][own_n]
	uint32_t window;                               // entries exported per rank
	uint2 *xport;                                  // [window + 3] own export: (exit, count) of entry own_last - 1 - i at [i]; [window] the final entry's;
	                                               // [window + 1] = (K, farthest look-back) from the rank with the final entry; [window + 2] = (F_OVERFLOW, F_BLIND)
	uint2 const *xport_all;                        // [size][window + 3] gathered
	uint32_t rank, size;
	uint64_t n_per;
	uint32_t *ctl;                                 // I_T, I_K, I_POS_FIRST, I_POS_LAST written; I_TRACE_ERR raised when a path enters a rank below its window
	uint32_t *pos_first_all;                       // [size + 1] first traceback position of every rank
	uint32_t *entry_all;                           // [size] the entry at which the path enters each rank (~0: it does not)
	uint32_t rows_bound;
	uint32_t *own;                                 // [rows_bound][5] this rank's part of the path in text order: rb (2 words), lb, max, size
	uint32_t const *own_all;                       // [size][rows_bound][5] gathered
	uint32_t t_bound;
	uint32_t *out_lb, *out_max, *out_size;         // [t_bound] the whole path in text order
	uint64_t *out_rb;
};
cudaError_t launch_ltrace_tables(ltrace_args const &A, cudaStream_t st);      // own tables + export
cudaError_t launch_ltrace_resolve(ltrace_args const &A, cudaStream_t st);     // chain over the gathered exports + own part of the path
cudaError_t launch_ltrace_scatter(ltrace_args const &A, cudaStream_t st);     // gathered parts -> the whole path

// Greedy reduction (fsq_dp.cu).  The counts #{d_rb > lb} it needs come from one row per traceback position: either
// `width` entries (v, #{d > v}) of the column's top list with #{d > v} <= K, v descending, padded with (~0, ~0)
// (rows are `width` uint2), or the divergence array itself (rows are m uint32).  A rank holds the rows of its own
// positions [pos_first, pos_last) and the first `halo` rows of every rank (halo_rows[size][halo]).
struct reduce_args
{
	uint32_t *ctl;               // I_T, I_K, I_POS_FIRST, I_POS_LAST are read here; I_HALO_ERR is raised here
	uint32_t m, width;           // row length in the two forms
	uint32_t row_kind;           // 0: list rows, 1: divergence rows
	uint32_t rows_bound;         // capacity of own_rows in rows
	uint64_t const *tb_rb;       // [T] right bounds of the traceback (replicated)
	uint32_t const *tb_size;     // [T]
	void *own_rows;              // [rows_bound] rows of this rank's positions, compact
	void const *halo_rows;       // [size][halo] rows, or null
	uint32_t const *pos_first_all;   // [size + 1] first traceback position of every rank (device), or null
	uint32_t halo, rank, size;
	uint32_t *next;              // [rows_bound + 1] compact: the last traceback entry merged into the segment that starts at own entry u
	uint32_t *next_size;         // [rows_bound + 1]: that segment's size; the last word receives I_HALO_ERR
	// launch_reduce_lists only: the top lists of this rank's columns
	uint2 const *toplist; uint32_t cap; uint32_t const *cnt; uint64_t col_base;
};
cudaError_t launch_reduce_lists(reduce_args const &A, cudaStream_t st);    // fill own_rows from the top lists
cudaError_t launch_reduce(reduce_args const &A, cudaStream_t st);
// pos_first_all[q] = first traceback position whose right bound is > q n_per, q = 0 .. size (device array of size + 1)
cudaError_t launch_rank_positions(uint64_t const *tb_rb, uint32_t const *ctl, uint64_t n_per, uint32_t size, uint32_t *pos_first_all, cudaStream_t st);
// next_all[size][rows_bound + 1] (gathered, compact per rank; the last word of nsize_all's blocks carries I_HALO_ERR) -> next[T], next_size[T]
cudaError_t launch_scatter_next(uint32_t const *next_all, uint32_t const *nsize_all, uint32_t const *pos_first_all, uint32_t size, uint32_t rows_bound, uint32_t t_bound, uint32_t *next, uint32_t *next_size, uint32_t *ctl, cudaStream_t st);

// The walk u -> next[u] + 1 from u = 0 (segmentation_lp_context.cc:335-390 merges in exactly this order), on the
// device by pointer doubling: the reduced segments come out as arrays without a host loop over the traceback.
struct walk_args
{
	uint32_t t_bound, levels;    // 2^levels > t_bound
	uint32_t *ctl;               // reads I_T; writes I_NRED, I_SEG_FIRST, I_SEG_LAST
	uint32_t const *next, *next_size;
	uint64_t const *tb_rb;       // [T]
	uint32_t *jump;              // [levels][t_bound]
	uint64_t *out_lb, *out_rb;   // [t_bound]
	uint32_t *out_size;
	uint32_t *out_pos;           // [t_bound] the traceback position a reduced segment ends at
	uint64_t c0, c1;             // this rank's columns: it holds the state of the segments whose right bound lies in (c0, c1]
};
cudaError_t launch_walk(walk_args const &A, cudaStream_t st);

cudaError_t launch_segment_sets(uint32_t const *a, uint32_t const *d, uint32_t m, uint64_t lb, uint32_t *run_off, uint32_t *ids, uint32_t *class_of_row, uint32_t *n_runs, cudaStream_t st);
cudaError_t launch_pair_weights(uint32_t const *cls_l, uint32_t const *cls_r, uint32_t m, uint32_t runs_r, int32_t *W, cudaStream_t st);
// the non-zero cells of W in row-major order: (lhs class, rhs class, count); at most min(m, runs_l runs_r) of them
cudaError_t launch_pair_counts(int32_t const *W, uint32_t runs_l, uint32_t runs_r, uint32_t *lhs, uint32_t *rhs, uint32_t *count, uint32_t *n_pairs, cudaStream_t st);

// ---- pack / synth ----------------------------------------------------------------------------------------------
cudaError_t launch_pack(uint8_t const *d_rows, uint64_t row_stride, uint32_t m, uint64_t n, uint8_t const *d_comp, uint8_t *cols, uint64_t col_stride, uint64_t col_first, uint32_t *sym_mask, cudaStream_t st);
// match_founder_sequences (fsq_match.cu)
cudaError_t launch_founder_masks(uint8_t const *fcols, uint64_t fstride, uint32_t K, uint64_t n, uint32_t nsym, uint32_t W, uint32_t *masks, cudaStream_t st);
cudaError_t launch_match_founders(uint8_t const *scols, uint64_t sstride, uint32_t m, uint64_t n, uint32_t const *masks, uint32_t nsym, uint32_t K, uint32_t W, uint64_t min_len,
	int mode, uint64_t *row_count, uint64_t *row_off, uint32_t *out_seq, uint64_t *out_lb, uint64_t *out_rb, uint32_t *out_bits, cudaStream_t st);
// identity columns (remove-identity-columns/main.cc:107-160) and the alignment without them
cudaError_t launch_identity_columns(uint8_t const *cols, uint64_t col_stride, uint32_t m, uint64_t n, uint8_t *identity, cudaStream_t st);
cudaError_t launch_kept_offsets(uint8_t const *identity, uint64_t n, uint32_t *block_off, uint64_t *total, cudaStream_t st);
cudaError_t launch_compact_columns(uint8_t const *cols, uint64_t col_stride, uint64_t n, uint8_t const *identity, uint32_t const *block_off, uint8_t *out, cudaStream_t st);
cudaError_t launch_synth_packed(uint8_t *cols, uint64_t col_stride, uint32_t m, uint64_t n, uint64_t col0, uint64_t seed, uint32_t founders, int gaps, cudaStream_t st);
cudaError_t launch_synth_rows(uint8_t *d_rows, uint64_t row_stride, uint32_t m, uint64_t n, uint64_t col0, uint64_t seed, uint32_t founders, int gaps, cudaStream_t st);

} // namespace fsq
