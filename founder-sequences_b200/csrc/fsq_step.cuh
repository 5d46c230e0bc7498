// The PBWT column step as a CTA-wide primitive (sm_100a).
//
// One call = one step of libbio::pbwt::build_prefix_and_divergence_arrays
// (reference: lib/libbio/include/libbio/pbwt.hh:66-142) on a state (a, d) that lives in shared memory:
//   * stable partition of the prefix array a into <= 16 symbol buckets, and
//   * d'[dst] = "first" marker for the first element of each symbol, else max(d[prev+1 .. i]) where prev is the
//     previous position with the same symbol.
//
// Layout of the work: positions are cut into tiles of 32; warp w owns TPW consecutive tiles; inside a tile one
// lane owns one position ("striped"), so that the per-symbol work is done once per 32 elements with
// __ballot_sync/__popc instead of once per element:
//   rank inside the bucket   = popc(ballot(sym == mine) & lanes below me)
//   previous same symbol      = highest set bit of that mask (else: carried in from earlier tiles)
//   range max (prev, lane]    = two reads of a 5-level doubling table held in registers (10 shuffles per tile,
//                               independent of the alphabet size)
// Bucket bases and "max since the last occurrence of symbol c" are carried from tile to tile in lane c of the
// warp, and from warp to warp through a small shared-memory scan (skipped when the CTA is a single warp).
// The state is updated in place: every warp first loads its tiles into registers (phase 1), and scatters only
// after the CTA-wide barrier (phase 2).
#pragma once
#include "fsq_common.cuh"

namespace fsq {

struct step_shared
{
	uint16_t *sa;       // [m] row ids in PBWT order
	void *sd;           // [m] divergence (uint16_t codes or uint32_t values)
	uint32_t *wcnt;     // [W][16] per-warp symbol counts, then bucket bases
	uint32_t *wcar;     // [W][16] per-warp "max after last occurrence", then carry-in
	uint32_t *whas;     // [W] bitmask: symbol occurs in the warp's tiles
};

template <typename DT> struct dtraits;
template <> struct dtraits <uint16_t> { static constexpr bool packed = true; };
template <> struct dtraits <uint32_t> { static constexpr bool packed = false; };

// HIST: maintain the divergence-value histogram H (16-bit counters packed two per word, indexed by code).
template <typename DT, int TPW, bool HIST, typename SymF>
__device__ __forceinline__ void column_step(
	step_shared const &sh,
	uint32_t const m,
	uint32_t const n_tiles,
	uint32_t const nsym,
	uint32_t const first_marker,   // value written for the first element of every symbol (k + 1, or its code)
	SymF symf,
	uint32_t *H
)
{
	constexpr bool PACKED = dtraits <DT>::packed;
	uint32_t const lane = lane_id();
	uint32_t const warp = threadIdx.x >> 5;
	uint32_t const n_warps = blockDim.x >> 5;
	bool const multi = n_warps > 1;
	uint32_t const lt = (1u << lane) - 1u;
	DT *sd = static_cast <DT *>(sh.sd);

	uint32_t rx[TPW];                        // packed: x | dv << 16
	uint32_t rdv[PACKED ? 1 : TPW];
	uint32_t rsym[(TPW + 7) / 8];
#pragma unroll
	for (int q = 0; q < (TPW + 7) / 8; ++q) rsym[q] = 0;

	// ---- phase 1: load, count, per-warp carry summary -----------------------------------------------------
	uint32_t cnt_c = 0, car_c = 0, has_c = 0;
#pragma unroll
	for (int j = 0; j < TPW; ++j)
	{
		uint32_t const tile = warp * TPW + j;
		if (tile < n_tiles)
		{
			uint32_t const i = tile * 32 + lane;
			bool const valid = i < m;
			uint32_t const x = valid ? sh.sa[i] : 0u;
			uint32_t const dv = valid ? (uint32_t) sd[i] : 0u;
			uint32_t const sym = valid ? symf(x) : 0u;
			if (PACKED) rx[j] = x | (dv << 16); else { rx[j] = x; rdv[PACKED ? 0 : j] = dv; }
			rsym[j >> 3] |= sym << ((j & 7) * 4);

			uint32_t myB = 0;
			for (uint32_t c = 0; c < nsym; ++c)
			{
				uint32_t const B = __ballot_sync(FSQ_FULL, valid && sym == c);
				if (lane == c) myB = B;
			}
			cnt_c += __popc(myB);

			if (multi)
			{
				uint32_t s = dv;        // suffix max over lanes lane..31
#pragma unroll
				for (int o = 1; o < 32; o <<= 1)
					s = max(s, __shfl_down_sync(FSQ_FULL, s, o));
				uint32_t const tmax = __shfl_sync(FSQ_FULL, s, 0);
				uint32_t const last = myB ? 31u - __clz(myB) : 0u;
				uint32_t const after = __shfl_sync(FSQ_FULL, s, min(last + 1u, 31u));
				if (myB) { car_c = (31u == last) ? 0u : after; has_c = 1; }
				else car_c = max(car_c, tmax);
			}
		}
	}

	uint32_t base_c, carry_c;
	if (multi)
	{
		if (lane < FSQ_MAX_SYMS)
		{
			sh.wcnt[warp * FSQ_MAX_SYMS + lane] = cnt_c;
			sh.wcar[warp * FSQ_MAX_SYMS + lane] = car_c;
		}
		uint32_t const hasmask = __ballot_sync(FSQ_FULL, has_c);
		if (0 == lane) sh.whas[warp] = hasmask;
		__syncthreads();
		if (0 == warp)
		{
			uint32_t run = 0;
			if (lane < FSQ_MAX_SYMS)
				for (uint32_t w = 0; w < n_warps; ++w)
				{
					uint32_t const t = sh.wcnt[w * FSQ_MAX_SYMS + lane];
					sh.wcnt[w * FSQ_MAX_SYMS + lane] = run;
					run += t;
				}
			uint32_t const sbase = warp_incl_sum(run) - run;
			if (lane < FSQ_MAX_SYMS)
			{
				uint32_t cin = first_marker;
				for (uint32_t w = 0; w < n_warps; ++w)
				{
					sh.wcnt[w * FSQ_MAX_SYMS + lane] += sbase;
					uint32_t const t = sh.wcar[w * FSQ_MAX_SYMS + lane];
					bool const h = (sh.whas[w] >> lane) & 1u;
					sh.wcar[w * FSQ_MAX_SYMS + lane] = cin;
					cin = h ? t : max(cin, t);
				}
			}
		}
		__syncthreads();
		base_c = (lane < FSQ_MAX_SYMS) ? sh.wcnt[warp * FSQ_MAX_SYMS + lane] : 0u;
		carry_c = (lane < FSQ_MAX_SYMS) ? sh.wcar[warp * FSQ_MAX_SYMS + lane] : 0u;
	}
	else
	{
		// A single warp owns the whole column: bucket bases are the exclusive scan of its own counts.
		base_c = warp_incl_sum(cnt_c) - cnt_c;
		carry_c = first_marker;
		__syncwarp();
	}

	// ---- phase 2: ranks, divergence, scatter ---------------------------------------------------------------
#pragma unroll
	for (int j = 0; j < TPW; ++j)
	{
		uint32_t const tile = warp * TPW + j;
		if (tile < n_tiles)
		{
			uint32_t const i = tile * 32 + lane;
			bool const valid = i < m;
			uint32_t const x = PACKED ? (rx[j] & 0xffffu) : rx[j];
			uint32_t const dv = PACKED ? (rx[j] >> 16) : rdv[PACKED ? 0 : j];
			uint32_t const sym = (rsym[j >> 3] >> ((j & 7) * 4)) & 15u;

			uint32_t Bs = 0, myB = 0;
			for (uint32_t c = 0; c < nsym; ++c)
			{
				uint32_t const B = __ballot_sync(FSQ_FULL, valid && sym == c);
				if (sym == c) Bs = B;
				if (lane == c) myB = B;
			}
			uint32_t const prevmask = Bs & lt;
			uint32_t const dst = __shfl_sync(FSQ_FULL, base_c, sym) + __popc(prevmask);
			base_c += __popc(myB);

			// doubling table: t_q = max dv over lanes (lane - 2^q, lane]
			uint32_t const t0 = dv;
			uint32_t const t1 = max(t0, __shfl_up_sync(FSQ_FULL, t0, 1));
			uint32_t const t2 = max(t1, __shfl_up_sync(FSQ_FULL, t1, 2));
			uint32_t const t3 = max(t2, __shfl_up_sync(FSQ_FULL, t2, 4));
			uint32_t const t4 = max(t3, __shfl_up_sync(FSQ_FULL, t3, 8));
			uint32_t const t5 = max(t4, __shfl_up_sync(FSQ_FULL, t4, 16));   // prefix max over lanes 0..lane

			uint32_t q = 5, src = lane;
			if (prevmask)
			{
				uint32_t const pl = 31u - __clz(prevmask);
				q = 31u - __clz(lane - pl);
				src = pl + (1u << q);
			}
			uint32_t const v0 = __shfl_sync(FSQ_FULL, t0, src);
			uint32_t const v1 = __shfl_sync(FSQ_FULL, t1, src);
			uint32_t const v2 = __shfl_sync(FSQ_FULL, t2, src);
			uint32_t const v3 = __shfl_sync(FSQ_FULL, t3, src);
			uint32_t const v4 = __shfl_sync(FSQ_FULL, t4, src);
			uint32_t const cin = __shfl_sync(FSQ_FULL, carry_c, sym);
			uint32_t const first = (0 == q) ? t0 : (1 == q) ? t1 : (2 == q) ? t2 : (3 == q) ? t3 : (4 == q) ? t4 : t5;
			uint32_t const second = (0 == q) ? v0 : (1 == q) ? v1 : (2 == q) ? v2 : (3 == q) ? v3 : (4 == q) ? v4 : cin;
			uint32_t const out = max(first, second);

			// carry for the next tile: max after the last occurrence of symbol `lane` (lane c keeps symbol c)
			{
				uint32_t s = dv;
#pragma unroll
				for (int o = 1; o < 32; o <<= 1)
					s = max(s, __shfl_down_sync(FSQ_FULL, s, o));
				uint32_t const tmax = __shfl_sync(FSQ_FULL, s, 0);
				uint32_t const last = myB ? 31u - __clz(myB) : 0u;
				uint32_t const after = __shfl_sync(FSQ_FULL, s, min(last + 1u, 31u));
				if (myB) carry_c = (31u == last) ? 0u : after;
				else carry_c = max(carry_c, tmax);
			}

			if (valid)
			{
				sh.sa[dst] = (uint16_t) x;
				sd[dst] = (DT) out;
				if (HIST && out != dv)
				{
					atomicAdd(&H[out >> 1], (out & 1u) ? 0x10000u : 1u);
					atomicSub(&H[dv >> 1], (dv & 1u) ? 0x10000u : 1u);
				}
			}
		}
	}
	__syncthreads();
}

// ---- CTA-wide exclusive scan of one value per thread (any blockDim <= 1024) -------------------------------
__device__ __forceinline__ uint32_t block_excl_scan(uint32_t v, uint32_t *scratch /* >= 33 words */, uint32_t &total)
{
	uint32_t const lane = lane_id();
	uint32_t const warp = threadIdx.x >> 5;
	uint32_t const n_warps = (blockDim.x + 31) >> 5;
	uint32_t const inc = warp_incl_sum(v);
	if (31 == lane) scratch[warp] = inc;
	__syncthreads();
	if (0 == warp)
	{
		uint32_t const t = lane < n_warps ? scratch[lane] : 0u;
		uint32_t const ti = warp_incl_sum(t);
		scratch[lane] = ti - t;
		if (31 == lane) scratch[32] = ti;
	}
	__syncthreads();
	uint32_t const res = scratch[warp] + inc - v;
	total = scratch[32];
	__syncthreads();
	return res;
}

} // namespace fsq
