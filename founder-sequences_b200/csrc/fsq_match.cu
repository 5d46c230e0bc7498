// match_founder_sequences on the device (reference: match-sequences-to-founders/match_founder_sequences.cc:127-214):
// every input row is walked left to right against the K founders, keeping the set of founders that agree with it since
// the current range began; when the set runs empty (or the range reaches the given length) the range is reported with
// the set as it stood and a new one begins.  Rows are independent: one warp per row.  The founders are turned first
// into one bit mask per (column, symbol) -- which founders carry that symbol there -- so that a step of the walk is
// one AND per 32 founders.
#include "fsq_internal.h"
#include "fsq_common.cuh"

namespace fsq {

// masks[c][s][w]: bit f of word w set iff founder 32 w + f has comp s at column c.  One warp per (column, word).
__global__ void founder_masks_kernel(uint8_t const *fcols, uint64_t fstride, uint32_t K, uint64_t n, uint32_t nsym, uint32_t W, uint32_t *masks)
{
	uint64_t const item = (uint64_t) blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	if (item >= n * W) return;
	uint64_t const c = item / W;
	uint32_t const w = (uint32_t) (item % W), lane = threadIdx.x & 31u;
	uint32_t const f = 32 * w + lane;
	uint32_t sym = 0xffffffffu;
	if (f < K) sym = (fcols[c * fstride + (f >> 1)] >> ((f & 1u) * 4u)) & 15u;
	for (uint32_t s = 0; s < nsym; ++s)
	{
		uint32_t const b = __ballot_sync(FSQ_FULL, sym == s);
		if (0 == lane) masks[(c * nsym + s) * W + w] = b;
	}
}

// mode 0: count the ranges of every row; mode 1: write them at row_off[row] + i.
__global__ void match_founders_kernel(
	uint8_t const *scols, uint64_t sstride, uint32_t m, uint64_t n, uint32_t const *masks, uint32_t nsym, uint32_t K, uint32_t W, uint64_t min_len,
	int mode, uint64_t *row_count, uint64_t const *row_off, uint32_t *out_seq, uint64_t *out_lb, uint64_t *out_rb, uint32_t *out_bits
)
{
	constexpr int NW = 8;                 // words per lane: K <= 8192 founders
	uint32_t const r = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	if (r >= m) return;
	uint32_t const lane = threadIdx.x & 31u;
	uint32_t all[NW], cand[NW];
#pragma unroll
	for (int i = 0; i < NW; ++i)
	{
		uint32_t const w = lane + 32u * i;
		uint32_t v = 0;
		if (w < W) v = (32 * (w + 1) <= K) ? 0xffffffffu : ((1u << (K - 32 * w)) - 1u);
		all[i] = v; cand[i] = v;
	}
	uint64_t lb = 0, emitted = 0;
	uint64_t const base = mode ? row_off[r] : 0;
	auto emit = [&](uint64_t rb) {
		if (mode)
		{
			uint64_t const at = base + emitted;
			if (0 == lane) { out_seq[at] = r; out_lb[at] = lb; out_rb[at] = rb; }
#pragma unroll
			for (int i = 0; i < NW; ++i)
			{
				uint32_t const w = lane + 32u * i;
				if (w < W) out_bits[at * W + w] = cand[i];
			}
		}
		++emitted;
	};
	for (uint64_t c = 0; c < n; ++c)
	{
		uint32_t const sym = (scols[c * sstride + (r >> 1)] >> ((r & 1u) * 4u)) & 15u;
		uint32_t const *mk = masks + (c * nsym + (sym < nsym ? sym : 0)) * W;
		uint32_t mw[NW];
#pragma unroll
		for (int i = 0; i < NW; ++i)
		{
			uint32_t const w = lane + 32u * i;
			mw[i] = (w < W && sym < nsym) ? mk[w] : 0u;
		}
		bool restart = 0 != min_len && min_len <= c - lb;          // :163-165
		if (!restart)
		{
			uint32_t any = 0;
#pragma unroll
			for (int i = 0; i < NW; ++i) any |= cand[i] & mw[i];
			if (__any_sync(FSQ_FULL, 0 != any))
			{
#pragma unroll
				for (int i = 0; i < NW; ++i) cand[i] &= mw[i];
				continue;
			}
			restart = true;                                            // :169-175 (the message goes to stderr in the reference)
		}
		// re_check, :179-193: report the range with the set as it stood, start anew at this column
		emit(c);
		lb = c;
#pragma unroll
		for (int i = 0; i < NW; ++i) cand[i] = all[i] & mw[i];        // empty when no founder has the character here
	}
	// :206-215: the last range, unless its set is empty
	{
		uint32_t any = 0;
#pragma unroll
		for (int i = 0; i < NW; ++i) any |= cand[i];
		if (__any_sync(FSQ_FULL, 0 != any)) emit(n);
	}
	if (0 == mode && 0 == lane) row_count[r] = emitted;
}

__global__ void row_offsets_kernel(uint64_t const *row_count, uint32_t m, uint64_t *row_off)
{
	// one thread: m is the number of input rows
	uint64_t sum = 0;
	for (uint32_t r = 0; r < m; ++r) { row_off[r] = sum; sum += row_count[r]; }
	row_off[m] = sum;
}

cudaError_t launch_founder_masks(uint8_t const *fcols, uint64_t fstride, uint32_t K, uint64_t n, uint32_t nsym, uint32_t W, uint32_t *masks, cudaStream_t st)
{
	uint64_t const items = n * W;
	founder_masks_kernel <<<(unsigned) ((items + 7) / 8), 256, 0, st>>>(fcols, fstride, K, n, nsym, W, masks);
	return cudaGetLastError();
}

cudaError_t launch_match_founders(uint8_t const *scols, uint64_t sstride, uint32_t m, uint64_t n, uint32_t const *masks, uint32_t nsym, uint32_t K, uint32_t W, uint64_t min_len,
	int mode, uint64_t *row_count, uint64_t *row_off, uint32_t *out_seq, uint64_t *out_lb, uint64_t *out_rb, uint32_t *out_bits, cudaStream_t st)
{
	// two warps per CTA: the rows' walks are long and independent, spread them over the SMs
	match_founders_kernel <<<(m + 1) / 2, 64, 0, st>>>(scols, sstride, m, n, masks, nsym, K, W, min_len, mode, row_count, row_off, out_seq, out_lb, out_rb, out_bits);
	if (0 == mode) row_offsets_kernel <<<1, 1, 0, st>>>(row_count, m, row_off);
	return cudaGetLastError();
}

} // namespace fsq
