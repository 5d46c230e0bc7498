// Pieces shared by the column-block scan kernels: the packed-column symbol source, the divergence-code histogram and
// the per-column top list that feeds the DP.
#pragma once
#include "fsq_internal.h"
#include "fsq_common.cuh"

namespace fsq {

// ---- column symbols: TMA-staged, 4-bit packed ---------------------------------------------------------------
struct nibble_source
{
	uint32_t const *col;   // one packed column in shared memory
	__device__ __forceinline__ uint32_t operator()(uint32_t x) const
	{
		return (col[x >> 3] >> ((x & 7u) * 4u)) & 15u;
	}
};

__device__ __forceinline__ uint32_t hist_get(uint32_t const *H, uint32_t code)
{
	uint32_t const w = H[code >> 1];
	return (code & 1u) ? (w >> 16) : (w & 0xffffu);
}

// Top list of column k for the DP (reference: the pairs of pbwt_context::output_divergence_value_counts(),
// lib/libbio/include/libbio/pbwt.hh:157-199, read from the largest value downwards): entries (v, S) with
// S = #{i : d[i] > v}, for every distinct v with S <= cap.  One warp.
__device__ __forceinline__ void emit_top_list(
	uint32_t const *H, uint32_t const top_code, uint32_t const R, uint32_t const *old, uint64_t const s_glob,
	uint32_t const m, uint32_t const cap, uint2 *out, uint32_t *cnt_out, uint32_t *h0_out, bool const zero_present
)
{
	uint32_t const lane = lane_id();
	uint32_t const lt = (1u << lane) - 1u;
	uint32_t S = 0, E = 0;
	bool dropped = false;
	int base = (int) top_code;
	while (base >= 0 && S <= cap)
	{
		int const code = base - (int) lane;
		uint32_t const c = code >= 0 ? hist_get(H, (uint32_t) code) : 0u;
		uint32_t const inc = warp_incl_sum(c);
		uint32_t const Sx = S + inc - c;
		bool const keep = c > 0 && Sx <= cap;
		uint32_t const mask = __ballot_sync(FSQ_FULL, keep);
		dropped |= 0 != __ballot_sync(FSQ_FULL, c > 0 && !keep);
		if (keep)
		{
			uint32_t const v = (uint32_t) code >= R ? (uint32_t) (s_glob + 1 + ((uint32_t) code - R)) : old[code];
			out[E + __popc(mask & lt)] = make_uint2(v, Sx);
		}
		E += __popc(mask);
		S += __shfl_sync(FSQ_FULL, inc, 31);
		base -= 32;
	}
	if (0 == lane)
	{
		bool const complete = !dropped && S >= m;
		*cnt_out = E | (complete ? 0x80000000u : 0u);
		*h0_out = zero_present ? hist_get(H, 0) : 0u;
	}
}

} // namespace fsq
