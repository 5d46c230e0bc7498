// Min-max segmentation DP with the reference's exact tie-breaking and RMQ behaviour, the greedy reduction, and
// the segment-set / edge-weight kernels (sm_100a).
#include "fsq_internal.h"
#include "fsq_common.cuh"
#include <cooperative_groups.h>
#include <algorithm>

namespace fsq {

// ---- founder_sequences::rmq <…, 64> emulated bit for bit (reference: include/founder_sequences/rmq.hh:61-118) ----
// lev[0][b]  = leftmost argmin of entries [64 b, 64 b + 64)
// lev[p][i]  = (M[lev[0][i + 2^p - 1]] < M[lev[p-1][i]]) ? lev[0][i + 2^p - 1] : lev[p-1][i]
//              -- rmq.hh:76-77 reads the same cell twice, so level p covers blocks {i, i+1, i+3, ..., i+2^p-1} only.
// The reference's linear scans inside a 64-block (naive_min, rmq.hh:86-118) are answered in O(1) from a per-entry
// mask: bit j of mk[t] is set iff entry 64 b + j (j <= t mod 64, b = t / 64) is <= every later entry up to t; the
// lowest set bit at or above a position is then the leftmost minimum of [position, t].

// Where the table is read from.  table_global: HBM / L2.  table_ring: the last DP_RING entries behind a chain's wavefront
// in shared memory (entries, masks and the first DP_SLEV rmq levels), global memory for anything older.
struct table_global
{
	uint32_t const *Mg; uint64_t const *mkg; uint32_t const *levg; uint32_t n_blk;
	__device__ __forceinline__ uint32_t M(uint32_t i) const { return Mg[i]; }
	__device__ __forceinline__ uint32_t M_front(uint32_t i) const { return Mg[i]; }
	__device__ __forceinline__ unsigned long long mask(uint32_t i) const { return mkg[i]; }
	__device__ __forceinline__ uint32_t lev(uint32_t p, uint32_t i) const { return levg[(size_t) p * n_blk + i]; }
};

constexpr uint32_t DP_RING = 2048;            // entries; a multiple of 64 and a power of two
constexpr uint32_t DP_RING_BLK = DP_RING / 64;
constexpr uint32_t DP_SLEV = 6;               // levels 0 .. 5: ranges of up to 63 full blocks

struct table_ring
{
	table_global g;
	uint32_t *sM; unsigned long long *smk; uint32_t *slev;      // [DP_RING], [DP_RING], [DP_SLEV][DP_RING_BLK]
	uint32_t lo;                                                  // entries >= lo are in the ring (lo is a multiple of 64)
	__device__ __forceinline__ uint32_t M(uint32_t i) const { return i >= lo ? sM[i & (DP_RING - 1)] : g.Mg[i]; }
	// an entry of the 64-block the wavefront is in: always in the ring (lo trails the wavefront by DP_RING - 128 at least)
	__device__ __forceinline__ uint32_t M_front(uint32_t i) const { return sM[i & (DP_RING - 1)]; }
	__device__ __forceinline__ unsigned long long mask(uint32_t i) const { return i >= lo ? smk[i & (DP_RING - 1)] : g.mkg[i]; }
	__device__ __forceinline__ uint32_t lev(uint32_t p, uint32_t i) const
	{
		return (p < DP_SLEV && 64u * i >= lo) ? slev[p * DP_RING_BLK + (i & (DP_RING_BLK - 1))] : g.lev(p, i);
	}
};

template <typename V>
__device__ __forceinline__ uint32_t inblock_min(V const &v, uint32_t first, uint32_t last)
{
	unsigned long long const msk = v.mask(last) & (~0ull << (first & 63u));
	return (last & ~63u) + (uint32_t) __ffsll((long long) msk) - 1u;
}

template <typename V>
__device__ __forceinline__ uint32_t rmq_query(V const &v, uint32_t beg, uint32_t end)
{
	uint32_t const bb = beg / 64 + 1;
	uint32_t const eb = end / 64;
	uint32_t const last = end - 1;
	if (bb >= eb)
	{
		// naive_min(beg, end): at most two consecutive blocks
		if ((beg >> 6) == (last >> 6))
			return inblock_min(v, beg, last);
		uint32_t const left = inblock_min(v, beg, beg | 63u);
		uint32_t const right = inblock_min(v, last & ~63u, last);
		return (v.M(right) < v.M(left)) ? right : left;
	}
	uint32_t const p = 31u - __clz(eb - bb);
	uint32_t const s1 = v.lev(p, bb);
	uint32_t const s2 = v.lev(p, eb - (1u << p));
	uint32_t smp = (v.M(s2) < v.M(s1)) ? s2 : s1;
	uint32_t const left = inblock_min(v, beg, beg | 63u);
	smp = (v.M(left) < v.M(smp)) ? left : smp;
	if (end == eb * 64)
		return smp;
	uint32_t const right = inblock_min(v, eb * 64, last);
	return (v.M(right) < v.M(smp)) ? right : smp;
}

// mk[t] from the table; one warp.  Returns the mask in every lane.
template <typename V>
__device__ __forceinline__ unsigned long long make_mask(V const &v, uint32_t t)
{
	uint32_t const lane = lane_id();
	uint32_t const blk = t & ~63u, tt = t & 63u;
	uint32_t const v0 = lane <= tt ? v.M_front(blk + lane) : 0xffffffffu;
	uint32_t const v1 = lane + 32 <= tt ? v.M_front(blk + 32 + lane) : 0xffffffffu;
	uint32_t s0 = v0, s1 = v1;          // inclusive suffix minima inside each half (shfl_down returns the own value past lane 31)
#pragma unroll
	for (int o = 1; o < 32; o <<= 1)
	{
		s0 = min(s0, __shfl_down_sync(FSQ_FULL, s0, o));
		s1 = min(s1, __shfl_down_sync(FSQ_FULL, s1, o));
	}
	uint32_t const hi_all = __shfl_sync(FSQ_FULL, s1, 0);
	uint32_t e0 = __shfl_down_sync(FSQ_FULL, s0, 1);
	uint32_t e1 = __shfl_down_sync(FSQ_FULL, s1, 1);
	e0 = (31 == lane) ? hi_all : min(e0, hi_all);
	e1 = (31 == lane) ? 0xffffffffu : e1;
	uint32_t const b0 = __ballot_sync(FSQ_FULL, lane <= tt && v0 <= e0);
	uint32_t const b1 = __ballot_sync(FSQ_FULL, lane + 32 <= tt && v1 <= e1);
	return (unsigned long long) b0 | ((unsigned long long) b1 << 32);
}

// The head of a column's top list, loaded ahead of time (the list lives in HBM; the DP itself runs out of shared memory).
struct list_head
{
	uint32_t cw, h0;        // entry count | complete << 31; #{d == 0}
	uint2 top;              // entry 0
	uint2 e0, e1;           // entries 1 + lane and 33 + lane
	uint32_t old_max;       // what the table holds for this entry (change detection without a round trip at write time)
	unsigned long long old_mask;
};

__device__ __forceinline__ list_head load_list_head(dp_args const &A, uint64_t k)
{
	uint32_t const lane = lane_id();
	uint64_t const kl = k - A.col_base;
	uint2 const *list = A.toplist + (k - A.list_base) * (A.cap + 1);
	list_head h;
	h.cw = A.cnt[kl];
	h.h0 = A.h0[kl];
	uint32_t const E = h.cw & 0x7fffffffu;
	// a column's slots all exist (cap + 1 of them), so these loads need not wait for the count; slots past E are ignored later
	(void) E;
	h.top = list[0];
	h.e0 = 1 + lane <= A.cap ? list[1 + lane] : make_uint2(0, 0);
	h.e1 = 33 + lane <= A.cap ? list[33 + lane] : make_uint2(0, 0);
	h.old_max = 0; h.old_mask = 0;
	return h;
}

// One column of calculate_segmentation_lp_dp_arg (reference: founder-sequences/segmentation_lp_context.cc:393-481),
// evaluated by one warp over the column's top list.  Only ranges whose segment size S is <= cap are listed; any
// other range has a candidate value > cap, so the result is exact whenever it is <= cap (or the list is complete).
// The list is in descending order of the divergence value, i.e. ascending S, and a candidate is worth at least its
// S: once S exceeds the best value so far nothing later can win or tie.
// Returns (in every lane) the entry's max; lb and size through the references.
template <typename V>
__device__ __forceinline__ uint32_t dp_column(dp_args const &A, V const &v, list_head const &h, uint64_t const k, uint64_t const t_out, uint32_t &reach, uint32_t &lowest, uint32_t &lb_out, uint32_t &size_out, bool const sure = false)
{
	uint32_t const lane = lane_id();
	uint32_t const L = (uint32_t) A.L;
	uint32_t const E = h.cw & 0x7fffffffu;
	bool const complete = h.cw >> 31;
	uint2 const *list = A.toplist + (k - A.list_base) * (A.cap + 1);
	uint32_t const hi_clip = (uint32_t) (k + 2 - L);

	// candidates compare as (value, range index) packed into 32 bits -- both are < 2^16 (m <= 65000) -- so that the
	// minimum over the warp is one redux instruction; ties: the lowest range (largest t) wins
	uint32_t best = 0xffffffffu;
	uint32_t best_idx = 0, best_S = 0;
	uint32_t far = (uint32_t) t_out;      // the lowest table entry a candidate looked at
	uint32_t carry_x = h.top.x;            // value of the entry before this round's first
	uint32_t carry_S = 0;                  // ... and its S: a lower bound for every later S
	for (uint32_t t0 = 1; t0 < E; t0 += 32)
	{
		if ((best >> 16) < carry_S)
			break;                          // before touching the next part of the list
		uint32_t const t = t0 + lane;
		uint2 en;
		if (1 == t0) en = h.e0;
		else if (33 == t0) en = h.e1;
		else en = t < E ? list[t] : make_uint2(0, 0);
		uint32_t const first_S = __shfl_sync(FSQ_FULL, en.y, 0);
		if ((best >> 16) < first_S)
			break;
		uint32_t up = __shfl_up_sync(FSQ_FULL, en.x, 1);
		if (0 == lane) up = carry_x;
		carry_x = __shfl_sync(FSQ_FULL, en.x, 31);
		carry_S = t0 + 31 < E ? __shfl_sync(FSQ_FULL, en.y, 31) : 0xffffffffu;
		uint32_t key = 0xffffffffu;
		uint32_t idx = 0;
		uint32_t const S = en.y;
		if (t < E && 0 != en.x)        // the range starting at value 0 is the whole-range candidate, below
		{
			uint32_t const lo = max(en.x, L);
			uint32_t const hi = min(up, hi_clip);
			if (lo < hi)
			{
				far = min(far, lo - L);
				idx = rmq_query(v, lo - L, hi - L);
				uint32_t const val = max(v.M(idx), S);
				key = (val << 16) | (0xffffu - t);
			}
		}
		uint32_t const kmin = __reduce_min_sync(FSQ_FULL, key);
		if (kmin < best)
		{
			uint32_t const owner = __ffs(__ballot_sync(FSQ_FULL, key == kmin)) - 1;
			best = kmin;
			best_idx = __shfl_sync(FSQ_FULL, idx, owner);
			best_S = __shfl_sync(FSQ_FULL, S, owner);
		}
	}
	far = __reduce_min_sync(FSQ_FULL, far);
	reach = max(reach, (uint32_t) t_out - far);
	lowest = min(lowest, far);
	uint32_t const whole = A.m - h.h0;          // == m when value 0 is absent
	uint32_t const best_val = best >> 16;
	uint32_t lb = 0, mx = whole, sz = whole;
	if (0xffffffffu != best && best_val < whole) { lb = best_idx + L; mx = best_val; sz = best_S; }
	// Under a list that is too short every value found is an upper bound of the true one (the candidates left out are
	// worth more than cap), and so is everything computed from it: lists that reach as far as the largest such value
	// prove the whole table.
	if (0 == lane && !complete && mx > A.cap)
	{
		// ... unless the list did not reach a single admissible candidate (all of it lies within the last L columns): then
		// the value is the whole-range one and says nothing about how long the lists need to be
		if (0xffffffffu == best) atomicAdd(&A.flags[F_BLIND], 1u);
		else atomicMax(&A.flags[F_OVERFLOW], mx);
		if (sure) A.flags[F_SURE] = 1;
	}
	lb_out = lb; size_out = sz;
	return mx;
}

// First pass staging (reference: segmentation_lp_context.cc:26-188).  Entry t (column t + L - 1) reads only
// entries <= t - L, so L consecutive entries are independent of one another: a wavefront.  The n / L wavefronts
// form a serial chain; to run it in parallel the table is cut into chunks and every chunk's chain is run at the same
// time against the table as it stands -- its own entries as it computes them, earlier chunks' entries as a previous
// round left them.  A round that changes no stored value has computed every entry from a table it agrees with,
// which by induction over the entries is the reference's table; until then rounds repeat (the first chunk is exact
// after one round, so at most n_chunks rounds are needed; the guess after the first round makes it two or three).
// Everything stored is written by exactly one chain: entries, masks and level-0 cells by the chunk's owner, a cell
// lev[p][i] by the owner of block i + 2^p - 1, whose completion defines it (rmq.hh:61-84).
//
// A chain is latency-bound (a wavefront is a handful of dependent look-ups), so it works out of shared memory: the
// last DP_RING entries, their masks and the low rmq levels sit in a ring behind the wavefront (a column looks back a
// few hundred entries at most on real data; anything older is read from global memory), every result is written
// through to global memory for the other chains and the later stages, and the head of the next wavefront's top lists
// is loaded while the current one is computed.  One warp per column of the wavefront.
#ifndef FSQ_DP_MIN_CTAS
#define FSQ_DP_MIN_CTAS 2
#endif
__global__ void __launch_bounds__(512, FSQ_DP_MIN_CTAS) dp_chain_kernel(dp_args const A)
{
	__shared__ uint32_t sM[DP_RING];
	__shared__ unsigned long long smk[DP_RING];
	__shared__ uint32_t slev[DP_SLEV * DP_RING_BLK];
	if (A.skip_when_done && A.flags[F_DONE]) return;
	uint32_t const warp = threadIdx.x >> 5;
	uint32_t const W = blockDim.x >> 5;           // columns per wavefront, <= min(L, 32)
	uint32_t const lane = lane_id();
	uint32_t const L = (uint32_t) A.L;
	// chains sit on a global grid of `chunk` entries; this rank runs the ones that meet [t_first, t_last)
	uint32_t const c = A.t_first / A.chunk + blockIdx.x;
	uint32_t const t_begin = max(A.t_first, c * A.chunk);
	uint32_t const t_end = min(A.t_last, (c + 1) * A.chunk);
	uint32_t reach = 0;
	// What a later chain may have read of this one while it ran: the last reach_bound entries (the look-back seen in the
	// rounds before).  A round in which no chain changed anything there, and no chain read below that window, has
	// computed every entry from values that are final: flags[3] counts the chains for which this cannot be said.
	// (bound_on_device: the bound is the copy of flags[2] the guess made before this round, flags[4], so that two rounds
	// can be queued without the host in between; the live flags[2] moves while the round runs and cannot serve)
	uint32_t const reach_bound = A.bound_on_device ? A.flags[F_BOUND] : A.reach_bound;
	uint32_t const tail_from = t_end > reach_bound ? t_end - reach_bound : 0u;
	uint32_t lowest = 0xffffffffu;
	bool unsettled = false, changed = false;      // published once per warp at the end (a store per entry would pile up on one address)

	table_ring v;
	v.g.Mg = A.tb_max; v.g.mkg = A.mk; v.g.levg = A.lev; v.g.n_blk = A.n_blk;
	v.sM = sM; v.smk = smk; v.slev = slev;
	// what the table holds behind the chain, as the previous round (or the other chains) left it
	{
		uint32_t const hi = (t_begin + 63u) & ~63u;
		uint32_t const lo = hi > DP_RING ? hi - DP_RING : 0u;
		for (uint32_t i = lo + threadIdx.x; i < t_begin; i += blockDim.x)
		{
			sM[i & (DP_RING - 1)] = A.tb_max[i];
			smk[i & (DP_RING - 1)] = A.mk[i];
		}
		for (uint32_t j = threadIdx.x; j < DP_SLEV * DP_RING_BLK; j += blockDim.x)
		{
			uint32_t const p = j / DP_RING_BLK, i = lo / 64 + j % DP_RING_BLK;
			if (p < A.n_levels && i < A.n_blk) slev[p * DP_RING_BLK + (i & (DP_RING_BLK - 1))] = A.lev[(size_t) p * A.n_blk + i];
		}
		v.lo = lo;
	}
	__syncthreads();

	auto load_head = [&](uint32_t t) {
		list_head h = load_list_head(A, (uint64_t) t + L - 1);
		h.old_max = A.tb_max[t];
		h.old_mask = A.mk[t];
		return h;
	};
	list_head next;
	if (t_begin + warp < t_end) next = load_head(t_begin + warp);
	uint32_t pend_old = 0, pend_new = 0;      // warp 0: a level cell's previous and new value, compared one completion later

	for (uint32_t t0 = t_begin; t0 < t_end; t0 += W)
	{
		uint32_t const t1 = min(t_end, t0 + W);
		uint32_t const t = t0 + warp;
		// the ring keeps [lo, t1): entries that would share a slot with this wavefront's are read from global memory
		uint32_t const hi = (t1 + 63u) & ~63u;
		v.lo = max(v.lo, hi > DP_RING ? hi - DP_RING : 0u);
		list_head const cur = next;
		if (t + W < t_end) next = load_head(t + W);
		if (t < t1)
		{
			uint32_t lb = 0, sz = 0;
			uint32_t const mx = dp_column(A, v, cur, (uint64_t) t + L - 1, t, reach, lowest, lb, sz, c >= A.chain_base && c - A.chain_base < A.round_no);
			if (0 == lane)
			{
				A.tb_max[t] = mx;
				A.tb_lb[t] = lb;
				A.tb_size[t] = sz;
				if (cur.old_max != mx) { changed = true; unsettled |= t >= tail_from; }
				sM[t & (DP_RING - 1)] = mx;       // this slot's previous tenant is older than anything the wavefront reads from the ring
			}
		}
		__syncthreads();

		// masks of the new entries; then rmq.update for the block that completed, if one did (W <= 64: at most one)
		if (t < t1)
		{
			unsigned long long const msk = make_mask(v, t);
			if (0 == lane)
			{
				smk[t & (DP_RING - 1)] = msk;
				A.mk[t] = msk;
				if (cur.old_mask != msk) { changed = true; unsettled |= t >= tail_from; }
			}
		}
		__syncthreads();
		if ((t1 >> 6) != (t0 >> 6))
		{
			if (0 == warp)
			{
				uint32_t const b = (t1 >> 6) - 1;
				uint32_t const smp0 = 64 * b + (uint32_t) __ffsll((long long) v.mask(64 * b + 63)) - 1u;      // leftmost minimal entry of the block
				uint32_t const val0 = v.M(smp0);
				if (pend_old != pend_new) { changed = true; unsettled = true; }
				pend_old = pend_new = 0;
				// lev[p][b + 1 - 2^p] from lev[p - 1][b + 1 - 2^p], which an earlier block's completion wrote: independent across p
				uint32_t const p = lane;
				if (p < A.n_levels && (0 == p || (1ull << p) <= (uint64_t) b + 1))
				{
					uint32_t const i = b + 1 - (1u << p);
					uint32_t res = smp0;
					if (p)
					{
						uint32_t const prev = v.lev(p - 1, i);
						res = (val0 < v.M(prev)) ? smp0 : prev;
					}
					pend_new = res;
					pend_old = atomicExch(&A.lev[(size_t) p * A.n_blk + i], res);     // the old value is looked at a block later
					if (p < DP_SLEV && 64u * i >= v.lo) slev[p * DP_RING_BLK + (i & (DP_RING_BLK - 1))] = res;
				}
			}
			// the next wavefront's queries may read what was just written
			__syncthreads();
		}
	}
	if (0 == warp && pend_old != pend_new) { changed = true; unsettled = true; }
	if (__any_sync(FSQ_FULL, changed) && 0 == lane) A.flags[F_CHANGED] = 1;
	if (0 == lane && reach) atomicMax(&A.flags[F_REACH], reach);
	// reads below the own entries must stay inside the window the chains before this one watch
	if (0xffffffffu != lowest && lowest < t_begin && t_begin - lowest > reach_bound) unsettled = true;
	if (__any_sync(FSQ_FULL, unsettled) && 0 == lane) atomicAdd(&A.flags[F_UNSETTLED], 1u);
}

cudaError_t launch_dp_chains(dp_args const &A, cudaStream_t st)
{
	// one warp per column of a wavefront
	uint32_t const n_warps = (uint32_t) (A.L < 16 ? A.L : 16);
	if (A.t_first >= A.t_last) return cudaSuccess;
	uint32_t const own_chunks = (A.t_last - 1) / A.chunk - A.t_first / A.chunk + 1;
	dp_chain_kernel <<<own_chunks, n_warps * 32, 0, st>>>(A);
	return cudaGetLastError();
}

// segmentation_lp_context.cc:165-185: the final column, from the final histogram
__global__ void dp_final_kernel(dp_args const A)
{
	uint32_t reach = 0, lowest = 0xffffffffu, lb = 0, sz = 0;
	table_global v;
	v.Mg = A.tb_max; v.mkg = A.mk; v.levg = A.lev; v.n_blk = A.n_blk;
	uint64_t const t = A.n - A.L;
	list_head const h = load_list_head(A, A.n - 1);
	uint32_t const mx = dp_column(A, v, h, A.n - 1, t, reach, lowest, lb, sz, 0 != A.flags[F_DONE]);
	if (0 == lane_id())
	{
		A.tb_max[t] = mx;
		A.tb_lb[t] = lb;
		A.tb_size[t] = sz;
	}
}

cudaError_t launch_dp_final(dp_args const &A, cudaStream_t st)
{
	dp_final_kernel <<<1, 32, 0, st>>>(A);
	return cudaGetLastError();
}

// ---- the guess between rounds -------------------------------------------------------------------------------------------
// A chain that starts from a wrong idea of the entries before its chunk only repairs its own chunk, so a correction
// would travel one chunk per round.  What travels is a level: an entry is max(an earlier entry within reach, a segment
// size) >= the smallest entry within reach, so the minimum over a window as long as the reach never decreases along the
// table and bounds every later entry from below.  After a round, every entry is therefore raised to the window minimum
// at the end of the chunk before it (computed from values that are themselves at most the true ones).  Any guess is
// allowed -- only a round that changes nothing ends the iteration -- a good one saves rounds.
__global__ void __launch_bounds__(128) dp_chunk_level_kernel(dp_args const A)
{
	__shared__ uint32_t wmin[4];
	if (A.skip_when_done && A.flags[F_DONE]) return;
	uint32_t const c = blockIdx.x;
	uint32_t const end = min(A.n_entries, (c + 1) * A.chunk);
	if (end - 1 < A.own_first || end - 1 >= A.own_last)
	{
		// another rank's chain: its minimum arrives with the exchange
		if (0 == threadIdx.x) A.chunk_level[A.n_chunks + c] = 0;
		return;
	}
	uint32_t const D = 2 * A.flags[F_REACH] + 64;          // twice the farthest look-back seen so far
	uint32_t const begin = end > D ? end - D : 0;
	uint32_t v = 0xffffffffu;
	for (uint32_t t = begin + threadIdx.x; t < end; t += blockDim.x) v = min(v, A.tb_max[t]);
	v = __reduce_min_sync(FSQ_FULL, v);
	if (0 == lane_id()) wmin[threadIdx.x >> 5] = v;
	__syncthreads();
	if (0 == threadIdx.x) A.chunk_level[A.n_chunks + c] = min(min(wmin[0], wmin[1]), min(wmin[2], wmin[3]));
}

// chunk_level[c] = max over c' < c of the window minimum at the end of chunk c'
__global__ void __launch_bounds__(1024) dp_level_kernel(dp_args const A)
{
	__shared__ uint32_t carry;
	__shared__ uint32_t wmax[32];
	uint32_t const lane = lane_id(), warp = threadIdx.x >> 5;
	if (A.skip_when_done && A.flags[F_DONE]) return;
	if (0 == threadIdx.x) { carry = 0; A.flags[F_BOUND] = A.flags[F_REACH]; }      // the look-back bound of the next round (no chain is running now)
	__syncthreads();
	for (uint32_t c0 = 0; c0 < A.n_chunks; c0 += blockDim.x)
	{
		uint32_t const c = c0 + threadIdx.x;
		uint32_t v = 0;
		if (c < A.n_chunks)
		{
			if (A.lvl_gathered) { for (uint32_t r = 0; r < A.lvl_ranks; ++r) v = max(v, A.lvl_gathered[(size_t) r * A.n_chunks + c]); }
			else v = A.chunk_level[A.n_chunks + c];
		}
		uint32_t s = v;                                    // inclusive running maximum inside the warp
#pragma unroll
		for (int o = 1; o < 32; o <<= 1)
			s = max(s, __shfl_up_sync(FSQ_FULL, s, o));      // shfl_up returns the own value below lane o
		if (31 == lane) wmax[warp] = s;
		__syncthreads();
		uint32_t before = carry;
		for (uint32_t w = 0; w < warp; ++w) before = max(before, wmax[w]);
		uint32_t const incl = max(before, s);
		uint32_t const prev = __shfl_up_sync(FSQ_FULL, incl, 1);
		if (c < A.n_chunks) A.chunk_level[c] = (0 == lane) ? before : prev;      // exclusive
		__syncthreads();
		if (blockDim.x - 1 == threadIdx.x) carry = incl;
		__syncthreads();
	}
}

__global__ void dp_raise_kernel(dp_args const A)
{
	uint32_t const t = A.lvl_first + blockIdx.x * blockDim.x + threadIdx.x;
	if (t >= A.lvl_last) return;
	if (A.skip_when_done && A.flags[F_DONE]) return;
	uint32_t const lvl = A.chunk_level[t / A.chunk];
	if (A.tb_max[t] < lvl) A.tb_max[t] = lvl;
}

// Masks and rmq cells are rebuilt for entries [A.w_first, A.w_last) (w_first a multiple of 64): a column-sharded rank
// reads the table only from its own last entry back to the look-back bound before its first, and a chain that reads
// below that bound asks for another round.
__global__ void dp_masks_kernel(dp_args const A)
{
	// One warp per block of 64 entries, the masks of the block one after the other: the entries that are <= every later
	// one up to t are those of t - 1 that are also <= entry t, and t itself.  (A third of the instructions of one suffix
	// scan per entry.)
	uint32_t const b = A.w_first / 64 + blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	uint32_t const base = 64 * b;
	if (base >= A.w_last) return;
	if (A.skip_when_done && A.flags[F_DONE]) return;
	uint32_t const lane = lane_id();
	uint32_t const count = min(64u, A.w_last - base);          // entries of this block that exist
	uint32_t const v0 = lane < count ? A.tb_max[base + lane] : 0xffffffffu;
	uint32_t const v1 = lane + 32 < count ? A.tb_max[base + 32 + lane] : 0xffffffffu;
	unsigned long long run = 0, m0 = 0, m1 = 0;
	for (uint32_t t = 0; t < count; ++t)
	{
		uint32_t const vt = __shfl_sync(FSQ_FULL, t < 32 ? v0 : v1, t & 31u);
		unsigned long long const le = (unsigned long long) __ballot_sync(FSQ_FULL, v0 <= vt) | ((unsigned long long) __ballot_sync(FSQ_FULL, v1 <= vt) << 32);
		run = (run & le) | (1ull << t);
		if (lane == (t & 31u)) { if (t < 32) m0 = run; else m1 = run; }
	}
	if (lane < count) A.mk[base + lane] = m0;
	if (lane + 32 < count) A.mk[base + 32 + lane] = m1;
}

// Level p of the rmq tables for every cell whose defining block is complete.
__global__ void dp_levels_kernel(dp_args const A, uint32_t p)
{
	uint32_t const i = A.w_first / 64 + blockIdx.x * blockDim.x + threadIdx.x;
	uint32_t const n_complete = A.w_last / 64;
	uint64_t const b = (uint64_t) i + (1ull << p) - 1;       // the block whose completion defines lev[p][i]
	if (b >= n_complete) return;
	uint32_t const smp0 = 64 * (uint32_t) b + (uint32_t) __ffsll((long long) A.mk[64 * b + 63]) - 1u;
	if (0 == p) { A.lev[i] = smp0; return; }
	uint32_t const prev = A.lev[(size_t) (p - 1) * A.n_blk + i];
	A.lev[(size_t) p * A.n_blk + i] = (A.tb_max[smp0] < A.tb_max[prev]) ? smp0 : prev;
}

cudaError_t launch_dp_chunk_minima(dp_args const &A, cudaStream_t st)
{
	dp_chunk_level_kernel <<<A.n_chunks, 128, 0, st>>>(A);
	return cudaGetLastError();
}

__global__ void dp_scatter_tails_kernel(uint32_t const *gathered, uint32_t rank, uint32_t tail, uint64_t n_per, uint32_t *tbx_max)
{
	// blockIdx.y: the rank whose tail this is (the ranks before this one)
	uint32_t const q = blockIdx.y;
	uint32_t const i = blockIdx.x * blockDim.x + threadIdx.x;
	if (q >= rank || i >= tail) return;
	tbx_max[(uint64_t) (q + 1) * n_per - tail + i] = gathered[(size_t) q * tail + i];
}

cudaError_t launch_dp_scatter_tails(uint32_t const *gathered, uint32_t size, uint32_t rank, uint32_t tail, uint64_t n_per, uint32_t *tbx_max, cudaStream_t st)
{
	(void) size;
	if (0 == rank || 0 == tail) return cudaSuccess;
	dp_scatter_tails_kernel <<<dim3((tail + 255) / 256, rank), 256, 0, st>>>(gathered, rank, tail, n_per, tbx_max);
	return cudaGetLastError();
}

cudaError_t launch_dp_guess(dp_args const &A, cudaStream_t st)
{
	// (launch_dp_rebuild is declared in fsq_internal.h)
	if (!A.lvl_gathered) dp_chunk_level_kernel <<<A.n_chunks, 128, 0, st>>>(A);
	dp_level_kernel <<<1, 1024, 0, st>>>(A);
	if (A.lvl_last > A.lvl_first) dp_raise_kernel <<<(A.lvl_last - A.lvl_first + 255) / 256, 256, 0, st>>>(A);
	return launch_dp_rebuild(A, st);
}


// Masks and rmq tables as a function of the table as it stands (no levelling): what a rank does after it has received the
// other ranks' entries.
// All levels of the rmq tables in one launch: level p of cell i needs level p - 1 of the same cell only, so the levels
// of a cell are computed by one thread, in order, without any barrier.
__global__ void dp_levels_all_kernel(dp_args const A)
{
	uint32_t const i = A.w_first / 64 + blockIdx.x * blockDim.x + threadIdx.x;
	uint32_t const n_complete = A.w_last / 64;
	if (A.skip_when_done && A.flags[F_DONE]) return;
	uint32_t prev = 0;
	for (uint32_t p = 0; p < A.n_levels; ++p)
	{
		uint64_t const b = (uint64_t) i + (1ull << p) - 1;       // the block whose completion defines lev[p][i]
		if (b >= n_complete) return;
		uint32_t const smp0 = 64 * (uint32_t) b + (uint32_t) __ffsll((long long) A.mk[64 * b + 63]) - 1u;
		uint32_t const res = (0 == p) ? smp0 : ((A.tb_max[smp0] < A.tb_max[prev]) ? smp0 : prev);
		A.lev[(size_t) p * A.n_blk + i] = res;
		prev = res;
	}
}

cudaError_t launch_dp_rebuild(dp_args const &A, cudaStream_t st)
{
	if (A.w_first >= A.w_last) return cudaSuccess;
	dp_masks_kernel <<<((A.w_last + 63) / 64 - A.w_first / 64 + 7) / 8, 256, 0, st>>>(A);
	uint32_t const n_complete = A.w_last / 64;
	uint32_t const first_blk = A.w_first / 64;
	if (first_blk < n_complete)
		dp_levels_all_kernel <<<(n_complete - first_blk + 255) / 256, 256, 0, st>>>(A);
	return cudaGetLastError();
}


// ---- end of a DP round: the decision stays on the device ----------------------------------------------------------------
__global__ void dp_round_end_kernel(uint32_t *flags, uint32_t first_round, uint32_t *send)
{
	// another round is needed: in the first round when anything changed at all; later, when a value changed that a chain after
	// its owner may have read while it was still the old one (dp_chain_kernel)
	bool const cont = !flags[F_DONE] && flags[F_CHANGED] && (first_round || flags[F_UNSETTLED]);
	send[0] = cont ? 1u : 0u;
	send[1] = flags[F_OVERFLOW];
	send[2] = flags[F_REACH];
	send[3] = min(flags[F_BLIND], 0x7fffffffu) | (flags[F_SURE] ? 0x80000000u : 0u);
}

__global__ void dp_round_merge_kernel(uint32_t *flags, uint32_t const *gathered, uint32_t size, uint32_t round_no, uint32_t back_bound, uint32_t one_chain)
{
	if (flags[F_DONE]) return;
	uint32_t cont = 0, ovf = 0, reach = 0, blind = 0, sure = 0;
	for (uint32_t r = 0; r < size; ++r)
	{
		cont |= gathered[4 * r];
		ovf = max(ovf, gathered[4 * r + 1]);
		reach = max(reach, gathered[4 * r + 2]);
		blind = max(blind, gathered[4 * r + 3] & 0x7fffffffu);
		sure |= gathered[4 * r + 3] >> 31;
	}
	flags[F_REACH] = reach;
	flags[F_ROUNDS] = round_no;
	if (sure) flags[F_SURE] = 1;
	if (2 * (uint64_t) reach + 128 > back_bound) flags[F_WINDOW_ERR] = 1;
	// (one chain over everything is the reference's own order: exact as it stands)
	if ((round_no > 1 && !cont) || one_chain)
	{
		flags[F_DONE] = 1;
		flags[F_OVERFLOW] = ovf;         // of the round that found the fixed point: its values are final
		flags[F_BLIND] = blind;
		return;
	}
	// a table that is still moving says nothing about the lists: values computed from guesses may exceed the final ones
	// (what the chains that are final already found out stays in F_SURE)
	if (!sure) { flags[F_OVERFLOW] = 0; flags[F_BLIND] = 0; }
	else { flags[F_OVERFLOW] = ovf; flags[F_BLIND] = blind; }
	flags[F_CHANGED] = 0;
	flags[F_UNSETTLED] = 0;
}

// after the last entry: whether any rank's lists were too short
__global__ void dp_overflow_merge_kernel(uint32_t *flags, uint32_t const *gathered, uint32_t size)
{
	uint32_t ovf = flags[F_OVERFLOW], blind = flags[F_BLIND];
	for (uint32_t r = 0; r < size; ++r)
	{
		ovf = max(ovf, gathered[4 * r + 1]);
		blind = max(blind, gathered[4 * r + 3] & 0x7fffffffu);
	}
	flags[F_OVERFLOW] = ovf;
	flags[F_BLIND] = blind;
}

cudaError_t launch_dp_overflow_merge(uint32_t *flags, uint32_t const *gathered, uint32_t size, cudaStream_t st)
{
	dp_overflow_merge_kernel <<<1, 1, 0, st>>>(flags, gathered, size);
	return cudaGetLastError();
}

cudaError_t launch_dp_round_end(uint32_t *flags, uint32_t first_round, uint32_t *send, cudaStream_t st)
{
	dp_round_end_kernel <<<1, 1, 0, st>>>(flags, first_round, send);
	return cudaGetLastError();
}

cudaError_t launch_dp_round_merge(uint32_t *flags, uint32_t const *gathered, uint32_t size, uint32_t round_no, uint32_t back_bound, uint32_t one_chain, cudaStream_t st)
{
	dp_round_merge_kernel <<<1, 1, 0, st>>>(flags, gathered, size, round_no, back_bound, one_chain);
	return cudaGetLastError();
}


// ---- traceback (reference: segmentation_lp_context.cc:191-224) --------------------------------------------------------------
constexpr uint32_t TRACE_END = 0xffffffffu;

__global__ void trace_init_kernel(trace_args const A)
{
	uint32_t const t = blockIdx.x * blockDim.x + threadIdx.x;
	if (t >= A.dp_size) return;
	uint32_t const lb = A.tb_lb[t];
	A.jump[t] = lb ? lb - A.L : TRACE_END;
}

__global__ void trace_double_kernel(trace_args const A, uint32_t k)
{
	uint32_t const t = blockIdx.x * blockDim.x + threadIdx.x;
	if (t >= A.dp_size) return;
	uint32_t const *prev = A.jump + (size_t) (k - 1) * A.dp_size;
	uint32_t const j = prev[t];
	A.jump[(size_t) k * A.dp_size + t] = (TRACE_END == j) ? TRACE_END : prev[j];
}

__device__ __forceinline__ void trace_count(trace_args const &A)
{
	uint32_t cur = A.dp_size - 1, hops = 0;
	for (int k = (int) A.levels - 1; k >= 0; --k)
	{
		uint32_t const j = A.jump[(size_t) k * A.dp_size + cur];
		if (TRACE_END != j) { cur = j; hops += 1u << k; }
	}
	A.ctl[I_T] = hops + 1;
	A.ctl[I_K] = A.tb_max[A.dp_size - 1];
}

__global__ void trace_count_kernel(trace_args const A) { trace_count(A); }

__global__ void trace_gather_kernel(trace_args const A)
{
	uint32_t const count = A.ctl[I_T];
	uint32_t const s = blockIdx.x * blockDim.x + threadIdx.x;      // hops back from the last entry
	if (s >= count || s >= A.t_bound) return;
	uint32_t cur = A.dp_size - 1;
	for (uint32_t k = 0; k < A.levels; ++k)
		if ((s >> k) & 1u) cur = A.jump[(size_t) k * A.dp_size + cur];
	uint32_t const o = count - 1 - s;
	A.out_lb[o] = A.tb_lb[cur];
	A.out_rb[o] = (uint64_t) cur + A.L;
	A.out_max[o] = A.tb_max[cur];
	A.out_size[o] = A.tb_size[cur];
}

// the traceback positions whose right bound lies in (c0, c1]: this rank holds the state there
__global__ void trace_own_kernel(trace_args const A)
{
	if (threadIdx.x >= 2) return;
	uint64_t const c = threadIdx.x ? A.c1 : A.c0;
	uint32_t lo = 0, hi = min(A.ctl[I_T], A.t_bound);
	while (lo < hi)
	{
		uint32_t const mid = (lo + hi) >> 1;
		if (A.out_rb[mid] <= c) lo = mid + 1; else hi = mid;
	}
	A.ctl[I_POS_FIRST + threadIdx.x] = lo;
}

// Level-by-level phases (pointer doubling, rmq levels) are a dozen or two launches of kernels that each run for a few
// microseconds: the launch rate of the host, not the device, sets their time, and on a busy host it varies by the
// millisecond.  They run instead as ONE cooperative launch with a grid-wide barrier between levels; the grid is what is
// resident at once.  (The one-launch-per-level forms stay as the fallback.)
template <typename K, typename A>
static bool launch_cooperative(K kernel, A const &args, uint32_t threads, uint64_t work_items, cudaStream_t st, cudaError_t *err)
{
	static int coop = -1, sms = 0;
	if (coop < 0)
	{
		int dev = 0;
		cudaGetDevice(&dev);
		cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev);
		cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
	}
	if (coop <= 0) return false;
	int per_sm = 0;
	if (cudaSuccess != cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, (int) threads, 0) || per_sm <= 0) { cudaGetLastError(); return false; }
	uint64_t const want = (work_items + threads - 1) / threads;
	uint32_t const grid = (uint32_t) std::max <uint64_t>(1, std::min <uint64_t>(want, (uint64_t) per_sm * sms));
	A a = args;
	void *params[] = {&a};
	*err = cudaLaunchCooperativeKernel((void const *) kernel, dim3(grid), dim3(threads), params, 0, st);
	return true;
}

__global__ void __launch_bounds__(256) trace_tables_coop_kernel(trace_args const A)
{
	cooperative_groups::grid_group grid = cooperative_groups::this_grid();
	uint32_t const stride = gridDim.x * blockDim.x, first = blockIdx.x * blockDim.x + threadIdx.x;
	for (uint32_t t = first; t < A.dp_size; t += stride)
	{
		uint32_t const lb = A.tb_lb[t];
		A.jump[t] = lb ? lb - A.L : TRACE_END;
	}
	for (uint32_t k = 1; k < A.levels; ++k)
	{
		grid.sync();
		uint32_t const *prev = A.jump + (size_t) (k - 1) * A.dp_size;
		uint32_t *cur = A.jump + (size_t) k * A.dp_size;
		for (uint32_t t = first; t < A.dp_size; t += stride)
		{
			uint32_t const j = prev[t];
			cur[t] = (TRACE_END == j) ? TRACE_END : prev[j];
		}
	}
	grid.sync();
	if (0 == first) trace_count(A);
}

cudaError_t launch_trace_tables(trace_args const &A, cudaStream_t st)
{
	cudaError_t err = cudaSuccess;
	if (launch_cooperative(trace_tables_coop_kernel, A, 256, A.dp_size, st, &err)) return err;
	uint32_t const grid = (A.dp_size + 255) / 256;
	trace_init_kernel <<<grid, 256, 0, st>>>(A);
	for (uint32_t k = 1; k < A.levels; ++k)
		trace_double_kernel <<<grid, 256, 0, st>>>(A, k);
	trace_count_kernel <<<1, 1, 0, st>>>(A);
	return cudaGetLastError();
}

cudaError_t launch_trace_gather(trace_args const &A, cudaStream_t st)
{
	trace_gather_kernel <<<(A.t_bound + 255) / 256, 256, 0, st>>>(A);
	trace_own_kernel <<<1, 32, 0, st>>>(A);
	return cudaGetLastError();
}


// ---- traceback over column shards ---------------------------------------------------------------------------------------
// Row i of the own tables is entry own_first + i; the last row is the final entry when this rank owns it.
__device__ __forceinline__ uint32_t ltrace_row(ltrace_args const &A, uint32_t t)
{
	return (A.has_final && t == A.dp_size - 1) ? A.own_n - 1 : t - A.own_first;
}

__device__ __forceinline__ uint32_t ltrace_entry(ltrace_args const &A, uint32_t row)
{
	return (A.has_final && row == A.own_n - 1) ? A.dp_size - 1 : A.own_first + row;
}

__device__ __forceinline__ bool ltrace_own(ltrace_args const &A, uint32_t t)
{
	return TRACE_END != t && t >= A.own_first && (t < A.own_last || (A.has_final && t == A.dp_size - 1));
}

__global__ void __launch_bounds__(256) ltrace_tables_coop_kernel(ltrace_args const A)
{
	cooperative_groups::grid_group grid = cooperative_groups::this_grid();
	uint32_t const stride = gridDim.x * blockDim.x, first = blockIdx.x * blockDim.x + threadIdx.x;
	for (uint32_t i = first; i < A.own_n; i += stride)
	{
		uint32_t const lb = A.tb_lb[ltrace_entry(A, i)];
		A.jump[i] = lb ? lb - A.L : TRACE_END;
		A.hops[i] = 1;
	}
	for (uint32_t k = 1; k < A.levels; ++k)
	{
		grid.sync();
		uint32_t const *pj = A.jump + (size_t) (k - 1) * A.own_n, *ph = A.hops + (size_t) (k - 1) * A.own_n;
		uint32_t *cj = A.jump + (size_t) k * A.own_n, *ch = A.hops + (size_t) k * A.own_n;
		for (uint32_t i = first; i < A.own_n; i += stride)
		{
			uint32_t const j = pj[i];
			if (ltrace_own(A, j)) { uint32_t const r = ltrace_row(A, j); cj[i] = pj[r]; ch[i] = ph[i] + ph[r]; }
			else { cj[i] = j; ch[i] = ph[i]; }
		}
	}
	grid.sync();
	// exports: where a path that enters at one of the last `window` entries leaves this rank, and how many entries it has here
	uint32_t const *tj = A.jump + (size_t) (A.levels - 1) * A.own_n, *th = A.hops + (size_t) (A.levels - 1) * A.own_n;
	uint32_t const own_cols = A.own_last - A.own_first;
	for (uint32_t i = first; i < A.window + 3; i += stride)
	{
		uint2 x = make_uint2(TRACE_END, 0);
		if (i < A.window) { if (i < own_cols) { uint32_t const r = own_cols - 1 - i; x = make_uint2(tj[r], th[r]); } }
		else if (i == A.window) { if (A.has_final) x = make_uint2(tj[A.own_n - 1], th[A.own_n - 1]); }
		else if (i == A.window + 1) { if (A.has_final) x = make_uint2(A.tb_max[A.dp_size - 1], A.ctl[F_REACH]); }
		else x = make_uint2(A.ctl[F_OVERFLOW], A.ctl[F_BLIND]);      // this rank's verdict on its lists rides along (the last entry's included)
		A.xport[i] = x;
	}
}

cudaError_t launch_ltrace_tables(ltrace_args const &A, cudaStream_t st)
{
	cudaError_t err = cudaSuccess;
	if (launch_cooperative(ltrace_tables_coop_kernel, A, 256, std::max <uint64_t>(A.own_n, A.window + 3), st, &err)) return err;
	return cudaErrorNotSupported;
}

// One thread walks the chain of exports (every rank does, and arrives at the same answer); then the rank's own part of the
// path is listed by all threads: the s-th entry before the one the path enters at is a binary descent through the tables.
constexpr uint32_t LTRACE_MAX_RANKS = 256;

__global__ void __launch_bounds__(256) ltrace_resolve_kernel(ltrace_args const A)
{
	__shared__ uint32_t s_first[LTRACE_MAX_RANKS + 1], s_enter[LTRACE_MAX_RANKS];
	__shared__ uint32_t s_bad;
	if (0 == threadIdx.x)
	{
		size_t const xs = (size_t) A.window + 3;
		for (uint32_t q = 0; q < A.size; ++q) { s_enter[q] = TRACE_END; s_first[q] = 0; }
		uint32_t const last = A.size - 1;
		uint2 const fin = A.xport_all[last * xs + A.window];
		uint2 const kk = A.xport_all[last * xs + A.window + 1];
		s_enter[last] = A.dp_size - 1;
		s_first[last] = fin.y;                      // counts for now; turned into offsets below
		uint32_t cur = fin.x;
		bool bad = false;
		while (TRACE_END != cur)
		{
			uint64_t const col = (uint64_t) cur + A.L - 1;
			uint32_t const q = (uint32_t) min((uint64_t) last, col / A.n_per);
			// rank q's per-column entries end at min(n_entries, (q + 1) n_per + 1 - L)
			uint64_t const hi = (uint64_t) (q + 1) * A.n_per + 1 > A.L ? (uint64_t) (q + 1) * A.n_per + 1 - A.L : 0;
			uint32_t const q_last = (uint32_t) min((uint64_t) (A.dp_size - A.L), hi);      // n_entries = dp_size - L
			if (cur >= q_last || q_last - 1 - cur >= A.window) { bad = true; break; }
			uint2 const x = A.xport_all[q * xs + (q_last - 1 - cur)];
			s_enter[q] = cur;
			s_first[q] = x.y;
			cur = x.x;
		}
		uint32_t sum = 0;
		for (uint32_t q = 0; q < A.size; ++q) { uint32_t const c = s_first[q]; s_first[q] = sum; sum += c; }
		s_first[A.size] = sum;
		if (sum > A.t_bound) bad = true;
		s_bad = bad ? 1u : 0u;
		if (0 == blockIdx.x)
		{
			for (uint32_t q = 0; q <= A.size; ++q) A.pos_first_all[q] = bad ? 0u : s_first[q];
			for (uint32_t q = 0; q < A.size; ++q) A.entry_all[q] = s_enter[q];
			A.ctl[I_T] = bad ? 0u : sum;
			A.ctl[I_K] = kk.x;
			uint32_t ovf = 0, blind = 0;
			for (uint32_t q = 0; q < A.size; ++q)
			{
				uint2 const v = A.xport_all[q * xs + A.window + 2];
				ovf = max(ovf, v.x); blind = max(blind, v.y);
			}
			A.ctl[F_OVERFLOW] = ovf;
			A.ctl[F_BLIND] = blind;
			A.ctl[I_POS_FIRST] = bad ? 0u : s_first[A.rank];
			A.ctl[I_POS_LAST] = bad ? 0u : s_first[A.rank + 1];
			if (bad) A.ctl[I_TRACE_ERR] = 1;
		}
	}
	__syncthreads();
	uint32_t const count = s_bad ? 0u : min(s_first[A.rank + 1] - s_first[A.rank], A.rows_bound);
	uint32_t const enter = s_enter[A.rank];
	for (uint32_t s = blockIdx.x * blockDim.x + threadIdx.x; s < count; s += gridDim.x * blockDim.x)
	{
		uint32_t row = ltrace_row(A, enter);
		for (uint32_t k = 0; k < A.levels; ++k)
			if ((s >> k) & 1u) row = ltrace_row(A, A.jump[(size_t) k * A.own_n + row]);
		uint32_t const t = ltrace_entry(A, row);
		uint32_t *o = A.own + (size_t) (count - 1 - s) * 5;
		uint64_t const rb = (uint64_t) t + A.L;
		o[0] = (uint32_t) rb; o[1] = (uint32_t) (rb >> 32);
		o[2] = A.tb_lb[t]; o[3] = A.tb_max[t]; o[4] = A.tb_size[t];
	}
}

cudaError_t launch_ltrace_resolve(ltrace_args const &A, cudaStream_t st)
{
	uint32_t const grid = std::max <uint32_t>(1, std::min <uint32_t>(1024, (A.rows_bound + 255) / 256));
	ltrace_resolve_kernel <<<grid, 256, 0, st>>>(A);
	return cudaGetLastError();
}

__global__ void ltrace_scatter_kernel(ltrace_args const A)
{
	uint32_t const q = blockIdx.y;
	uint32_t const i = blockIdx.x * blockDim.x + threadIdx.x;
	uint32_t const first = A.pos_first_all[q], last = A.pos_first_all[q + 1];
	if (i >= A.rows_bound || first + i >= last || first + i >= A.t_bound) return;
	uint32_t const *o = A.own_all + ((size_t) q * A.rows_bound + i) * 5;
	A.out_rb[first + i] = (uint64_t) o[0] | ((uint64_t) o[1] << 32);
	A.out_lb[first + i] = o[2];
	A.out_max[first + i] = o[3];
	A.out_size[first + i] = o[4];
}

cudaError_t launch_ltrace_scatter(ltrace_args const &A, cudaStream_t st)
{
	ltrace_scatter_kernel <<<dim3((A.rows_bound + 255) / 256, A.size), 256, 0, st>>>(A);
	return cudaGetLastError();
}


// ---- greedy reduction (reference: segmentation_lp_context.cc:335-390; pbwt.hh:206-216) ------------------------------
// The reference merges traceback segments while #{d_rb > cur_lb} <= K.  That count is the number of distinct rows over
// [cur_lb, rb), which cannot decrease as rb grows, so the run merged from a given start is a function of the start
// alone: one warp per start u (u = 0: left bound 0; u >= 1: left bound rb[u - 1]) extends while the count stays <= K;
// the walk below then follows next[] from u = 0.
// The count itself comes from the top list of column rb - 1: with v* the largest divergence value <= lb,
// #{d > lb} = #{d > v*} (no value lies in (v*, lb]), which is the S of the first list entry whose v is <= lb.  Only
// entries with S <= K matter -- no such entry means the count exceeds K (K < m here) -- so each traceback position
// keeps a row of the entries with S <= K.  Where the top lists are gone by then (the scan ran in windows over a reused
// list buffer) the rows are the divergence arrays at the traceback positions instead and the count is taken directly.
// A column-sharded rank does this for its own positions; a run that leaves them continues over the first rows of the
// ranks after it (the halo), which is all the ranks exchange.
__global__ void reduce_lists_kernel(reduce_args const A)
{
	uint32_t const first = A.ctl[I_POS_FIRST], last = A.ctl[I_POS_LAST], K = A.ctl[I_K];
	uint32_t const i = blockIdx.x;
	if (first + i >= last) return;
	uint64_t const kl = A.tb_rb[first + i] - 1 - A.col_base;
	uint32_t const E = A.cnt[kl] & 0x7fffffffu;
	uint2 const *list = A.toplist + kl * (A.cap + 1);
	uint2 *out = (uint2 *) A.own_rows + (size_t) i * A.width;
	for (uint32_t j = threadIdx.x; j < A.width; j += blockDim.x)
	{
		uint2 e = make_uint2(0xffffffffu, 0xffffffffu);
		if (j < E) { uint2 const x = list[j]; if (x.y <= K) e = x; }
		out[j] = e;
	}
}

cudaError_t launch_reduce_lists(reduce_args const &A, cudaStream_t st)
{
	if (0 == A.rows_bound) return cudaSuccess;
	reduce_lists_kernel <<<A.rows_bound, 64, 0, st>>>(A);
	return cudaGetLastError();
}

__global__ void rank_positions_kernel(uint64_t const *tb_rb, uint32_t const *ctl, uint64_t n_per, uint32_t size, uint32_t *pos_first_all)
{
	uint32_t const q = blockIdx.x * blockDim.x + threadIdx.x;
	if (q > size) return;
	uint32_t const T = ctl[I_T];
	uint64_t const c = (uint64_t) q * n_per;
	uint32_t lo = 0, hi = T;
	if (q == size) lo = T;
	while (lo < hi)
	{
		uint32_t const mid = (lo + hi) >> 1;
		if (tb_rb[mid] <= c) lo = mid + 1; else hi = mid;
	}
	pos_first_all[q] = lo;
}

cudaError_t launch_rank_positions(uint64_t const *tb_rb, uint32_t const *ctl, uint64_t n_per, uint32_t size, uint32_t *pos_first_all, cudaStream_t st)
{
	rank_positions_kernel <<<(size + 1 + 63) / 64, 64, 0, st>>>(tb_rb, ctl, n_per, size, pos_first_all);
	return cudaGetLastError();
}

// #{d > lb} for the row of traceback position t, one warp; ~0 when the row is out of reach (halo exhausted)
__device__ __forceinline__ uint32_t reduce_count(reduce_args const &A, uint32_t t, uint32_t first, uint32_t last, uint32_t cur_lb, uint32_t K, bool &missing)
{
	uint32_t const lane = lane_id();
	size_t const row_words = A.row_kind ? (size_t) A.m : 2 * (size_t) A.width;
	uint32_t const *row = nullptr;
	if (t < last) row = (uint32_t const *) A.own_rows + (size_t) (t - first) * row_words;
	else if (A.halo_rows)
	{
		// the rank after this one that holds position t, and its local index there
		uint32_t q = A.rank + 1;
		while (q < A.size && A.pos_first_all[q + 1] <= t) ++q;
		uint32_t const i = t - A.pos_first_all[q];
		if (q < A.size && i < A.halo) row = (uint32_t const *) A.halo_rows + ((size_t) q * A.halo + i) * row_words;
	}
	if (!row) { missing = true; return 0xffffffffu; }
	if (A.row_kind)
	{
		uint32_t c = 0;
		for (uint32_t i = lane; i < A.m; i += 32) c += row[i] > cur_lb;
		return __reduce_add_sync(FSQ_FULL, c);
	}
	uint2 const *list = (uint2 const *) row;
	for (uint32_t i0 = 0; i0 < A.width; i0 += 32)
	{
		uint32_t const i = i0 + lane;
		uint2 const e = i < A.width ? list[i] : make_uint2(0xffffffffu, 0xffffffffu);
		uint32_t const hit = __ballot_sync(FSQ_FULL, e.x <= cur_lb);       // the padding (v = ~0) never matches
		if (hit) return __shfl_sync(FSQ_FULL, e.y, __ffs(hit) - 1);
		if (__any_sync(FSQ_FULL, e.y > K)) break;                          // S ascends: nothing further counts
	}
	return 0xffffffffu;
}

__global__ void reduce_kernel(reduce_args const A)
{
	uint32_t const lane = lane_id();
	uint32_t const first = A.ctl[I_POS_FIRST], last = A.ctl[I_POS_LAST], T = A.ctl[I_T], K = A.ctl[I_K];
	uint32_t const i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	uint32_t const u = first + i;
	if (u >= last) return;
	uint32_t const cur_lb = (0 == u) ? 0u : (uint32_t) A.tb_rb[u - 1];
	// The rows at hand are a prefix of the positions from `first` on: the own ones, then the first `halo` rows of every rank
	// behind this one until a rank has more positions than that.
	uint32_t avail = last;
	if (A.halo_rows)
		for (uint32_t q = A.rank + 1; q < A.size; ++q)
		{
			uint32_t const cnt = A.pos_first_all[q + 1] - A.pos_first_all[q];
			avail += min(cnt, A.halo);
			if (cnt > A.halo) break;
		}
	avail = min(avail, T);
	// The count #{d_rb(t) > cur_lb} cannot decrease with t (it is the number of distinct rows over [cur_lb, rb(t))), so the
	// last position that still merges is found by galloping and bisection: a logarithmic number of counts per start where
	// walking the run costs its length (runs are hundreds of positions long when K is in the thousands, and with
	// divergence rows a count reads m values).
	bool missing = false;
	uint32_t lo = u, size = A.tb_size[u];             // count(lo) <= K
	uint32_t hi = u + 1, step = 1;                    // first position not yet known to merge
	while (hi < avail)
	{
		uint32_t const s = reduce_count(A, hi, first, last, cur_lb, K, missing);
		if (s > K) break;
		lo = hi; size = s;
		hi = min(avail, hi + step);
		step <<= 1;
		if (lo + 1 == avail) { hi = avail; break; }
	}
	if (hi >= avail) hi = avail;                      // everything at hand up to avail - 1 may merge: bisect (lo, avail)
	else { /* count(hi) > K */ }
	// invariant: count(lo) <= K; hi == avail or count(hi) > K
	{
		uint32_t l = lo, h = hi;                      // answer in [l, h)
		while (l + 1 < h)
		{
			uint32_t const mid = l + (h - l) / 2;
			uint32_t const s = reduce_count(A, mid, first, last, cur_lb, K, missing);
			if (s <= K) { l = mid; size = s; } else h = mid;
		}
		lo = l; hi = h;
	}
	// the run ends at lo; that is certain when the position behind it was counted (> K) or does not exist
	bool const ran_out = hi == avail && avail < T && lo + 1 == avail;
	if (0 == lane)
	{
		A.next[i] = lo; A.next_size[i] = size;
		if (ran_out || missing) A.ctl[I_HALO_ERR] = 1;
	}
}

// the "ran out of rows" flag rides behind the runs (word rows_bound) so that every rank learns of it
__global__ void reduce_flag_kernel(reduce_args const A)
{
	A.next_size[A.rows_bound] = A.ctl[I_HALO_ERR];
	A.next[A.rows_bound] = 0;
}

cudaError_t launch_reduce(reduce_args const &A, cudaStream_t st)
{
	if (0 == A.rows_bound) return cudaSuccess;
	uint32_t const wpb = 8;
	reduce_kernel <<<(A.rows_bound + wpb - 1) / wpb, wpb * 32, 0, st>>>(A);
	reduce_flag_kernel <<<1, 1, 0, st>>>(A);
	return cudaGetLastError();
}

// (blocks of rows_bound + 1 words per rank: the last word of next_size is the rank's "ran out of rows" flag)
__global__ void scatter_next_kernel(uint32_t const *next_all, uint32_t const *nsize_all, uint32_t const *pos_first_all, uint32_t size, uint32_t rows_bound, uint32_t t_bound, uint32_t *next, uint32_t *next_size, uint32_t *ctl)
{
	uint32_t const q = blockIdx.y;
	uint32_t const i = blockIdx.x * blockDim.x + threadIdx.x;
	size_t const stride = (size_t) rows_bound + 1;
	if (0 == i && nsize_all[q * stride + rows_bound]) ctl[I_HALO_ERR] = 1;
	uint32_t const first = pos_first_all[q], last = pos_first_all[q + 1];
	if (i >= rows_bound || first + i >= last || first + i >= t_bound) return;
	next[first + i] = next_all[q * stride + i];
	next_size[first + i] = nsize_all[q * stride + i];
}

cudaError_t launch_scatter_next(uint32_t const *next_all, uint32_t const *nsize_all, uint32_t const *pos_first_all, uint32_t size, uint32_t rows_bound, uint32_t t_bound, uint32_t *next, uint32_t *next_size, uint32_t *ctl, cudaStream_t st)
{
	if (0 == rows_bound) return cudaSuccess;
	scatter_next_kernel <<<dim3((rows_bound + 255) / 256, size), 256, 0, st>>>(next_all, nsize_all, pos_first_all, size, rows_bound, t_bound, next, next_size, ctl);
	return cudaGetLastError();
}


// ---- the reduced segments from next[] -----------------------------------------------------------------------------------
__device__ __forceinline__ void walk_count(walk_args const &A, uint32_t T)
{
	uint32_t cur = 0, hops = 0;
	for (int k = (int) A.levels - 1; k >= 0; --k)
	{
		uint32_t const j = A.jump[(size_t) k * A.t_bound + cur];
		if (TRACE_END != j) { cur = j; hops += 1u << k; }
	}
	A.ctl[I_NRED] = T ? hops + 1 : 0;
}

__device__ __forceinline__ void walk_own(walk_args const &A, uint32_t which)
{
	uint64_t const c = which ? A.c1 : A.c0;
	uint32_t lo = 0, hi = A.ctl[I_NRED];
	while (lo < hi)
	{
		uint32_t const mid = (lo + hi) >> 1;
		if (A.out_rb[mid] <= c) lo = mid + 1; else hi = mid;
	}
	A.ctl[I_SEG_FIRST + which] = lo;
}

__global__ void walk_init_kernel(walk_args const A)
{
	uint32_t const T = min(A.ctl[I_T], A.t_bound);
	uint32_t const u = blockIdx.x * blockDim.x + threadIdx.x;
	if (u >= T) return;
	uint32_t const v = A.next[u] + 1;
	A.jump[u] = v < T ? v : TRACE_END;
}

__global__ void walk_double_kernel(walk_args const A, uint32_t k)
{
	uint32_t const T = min(A.ctl[I_T], A.t_bound);
	uint32_t const u = blockIdx.x * blockDim.x + threadIdx.x;
	if (u >= T) return;
	uint32_t const *prev = A.jump + (size_t) (k - 1) * A.t_bound;
	uint32_t const j = prev[u];
	A.jump[(size_t) k * A.t_bound + u] = (TRACE_END == j) ? TRACE_END : prev[j];
}

__global__ void walk_count_kernel(walk_args const A) { walk_count(A, min(A.ctl[I_T], A.t_bound)); }

__global__ void walk_gather_kernel(walk_args const A)
{
	uint32_t const s = blockIdx.x * blockDim.x + threadIdx.x;      // segments in text order = hops from u = 0
	if (s >= A.ctl[I_NRED]) return;
	uint32_t cur = 0;
	for (uint32_t k = 0; k < A.levels; ++k)
		if ((s >> k) & 1u) cur = A.jump[(size_t) k * A.t_bound + cur];
	A.out_lb[s] = cur ? A.tb_rb[cur - 1] : 0ull;
	A.out_rb[s] = A.tb_rb[A.next[cur]];
	A.out_size[s] = A.next_size[cur];
	A.out_pos[s] = A.next[cur];
}

// upper bounds of c0 and c1 among the right bounds: the segments whose end state this rank holds
__global__ void walk_own_kernel(walk_args const A)
{
	if (threadIdx.x < 2) walk_own(A, threadIdx.x);
}

__global__ void __launch_bounds__(256) walk_coop_kernel(walk_args const A)
{
	cooperative_groups::grid_group grid = cooperative_groups::this_grid();
	uint32_t const T = min(A.ctl[I_T], A.t_bound);
	uint32_t const stride = gridDim.x * blockDim.x, first = blockIdx.x * blockDim.x + threadIdx.x;
	for (uint32_t u = first; u < T; u += stride)
	{
		uint32_t const v = A.next[u] + 1;
		A.jump[u] = v < T ? v : TRACE_END;
	}
	for (uint32_t k = 1; k < A.levels; ++k)
	{
		grid.sync();
		uint32_t const *prev = A.jump + (size_t) (k - 1) * A.t_bound;
		uint32_t *cur = A.jump + (size_t) k * A.t_bound;
		for (uint32_t u = first; u < T; u += stride)
		{
			uint32_t const j = prev[u];
			cur[u] = (TRACE_END == j) ? TRACE_END : prev[j];
		}
	}
	grid.sync();
	if (0 == first) walk_count(A, T);
	grid.sync();
	uint32_t const count = A.ctl[I_NRED];
	for (uint32_t s = first; s < count; s += stride)
	{
		uint32_t cur = 0;
		for (uint32_t k = 0; k < A.levels; ++k)
			if ((s >> k) & 1u) cur = A.jump[(size_t) k * A.t_bound + cur];
		A.out_lb[s] = cur ? A.tb_rb[cur - 1] : 0ull;
		A.out_rb[s] = A.tb_rb[A.next[cur]];
		A.out_size[s] = A.next_size[cur];
		A.out_pos[s] = A.next[cur];
	}
	grid.sync();
	if (first < 2) walk_own(A, first);
}

cudaError_t launch_walk(walk_args const &A, cudaStream_t st)
{
	cudaError_t err = cudaSuccess;
	if (launch_cooperative(walk_coop_kernel, A, 256, A.t_bound, st, &err)) return err;
	uint32_t const grid = (A.t_bound + 255) / 256;
	walk_init_kernel <<<grid, 256, 0, st>>>(A);
	for (uint32_t k = 1; k < A.levels; ++k)
		walk_double_kernel <<<grid, 256, 0, st>>>(A, k);
	walk_count_kernel <<<1, 1, 0, st>>>(A);
	walk_gather_kernel <<<grid, 256, 0, st>>>(A);
	walk_own_kernel <<<1, 32, 0, st>>>(A);
	return cudaGetLastError();
}


// ---- segment sets (reference: pbwt.hh:253-280, create_segment_texts_task.cc:15-35) ----------------------------------
// Runs of the prefix array w.r.t. the segment's left bound, in PBWT order; the rows of a run in ascending order.
__global__ void __launch_bounds__(1024) segment_sets_kernel(
	uint32_t const *a, uint32_t const *d, uint32_t m, uint32_t lb,
	uint32_t *run_off, uint32_t *ids, uint32_t *class_of_row, uint32_t *n_runs
)
{
	extern __shared__ __align__(16) uint8_t smem_raw[];
	uint32_t *cursor = (uint32_t *) smem_raw;          // [m + 1]
	__shared__ uint32_t scratch[34];
	uint32_t const tid = threadIdx.x;
	uint32_t const per = (m + blockDim.x - 1) / blockDim.x;
	uint32_t const lo = min(tid * per, m), hi = min(lo + per, m);
	uint32_t cnt = 0;
	for (uint32_t i = lo; i < hi; ++i) cnt += (0 == i || lb < d[i]);
	// inline block scan
	uint32_t const lane = lane_id(), warp = tid >> 5, n_warps = (blockDim.x + 31) >> 5;
	uint32_t const inc = warp_incl_sum(cnt);
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
	uint32_t c = scratch[warp] + inc - cnt;
	uint32_t const total = scratch[32];
	for (uint32_t i = lo; i < hi; ++i)
	{
		if (0 == i || lb < d[i]) { run_off[c] = i; cursor[c] = i; ++c; }
		class_of_row[a[i]] = c - 1;
	}
	if (0 == tid) { run_off[total] = m; *n_runs = total; }
	__syncthreads();
	// rows in ascending order are appended to their run: a stable multi-split done by one warp, 32 rows at a time
	if (0 == warp)
	{
		for (uint32_t x0 = 0; x0 < m; x0 += 32)
		{
			uint32_t const x = x0 + lane;
			bool const valid = x < m;
			uint32_t const cl = valid ? class_of_row[x] : 0xffffffffu;
			uint32_t const peers = __match_any_sync(FSQ_FULL, cl);
			uint32_t const rank = __popc(peers & ((1u << lane) - 1u));
			uint32_t base = 0;
			if (valid) base = cursor[cl];
			__syncwarp();
			if (valid)
			{
				ids[base + rank] = x;
				if (0 == rank) cursor[cl] = base + __popc(peers);
			}
			__syncwarp();
		}
	}
}

cudaError_t launch_segment_sets(uint32_t const *a, uint32_t const *d, uint32_t m, uint64_t lb, uint32_t *run_off, uint32_t *ids, uint32_t *class_of_row, uint32_t *n_runs, cudaStream_t st)
{
	size_t const smem = sizeof(uint32_t) * ((size_t) m + 1);
	cudaError_t err = cudaFuncSetAttribute(segment_sets_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
	if (cudaSuccess != err) return err;
	uint32_t threads = 1024;
	while (threads > 32 && threads / 2 >= m) threads /= 2;
	segment_sets_kernel <<<1, threads, smem, st>>>(a, d, m, (uint32_t) lb, run_off, ids, class_of_row, n_runs);
	return cudaGetLastError();
}


// ---- edge weights (reference: merge_segments_task.cc:133-176, set_count.hh:61-89) -----------------------------------
// |S_i ∩ T_j| over all pairs of runs is the contingency table of the two class maps.
__global__ void pair_weights_kernel(uint32_t const *cls_l, uint32_t const *cls_r, uint32_t m, uint32_t runs_r, int32_t *W)
{
	uint32_t const x = blockIdx.x * blockDim.x + threadIdx.x;
	if (x < m)
		atomicAdd(&W[(size_t) cls_l[x] * runs_r + cls_r[x]], 1);
}

cudaError_t launch_pair_weights(uint32_t const *cls_l, uint32_t const *cls_r, uint32_t m, uint32_t runs_r, int32_t *W, cudaStream_t st)
{
	pair_weights_kernel <<<(m + 255) / 256, 256, 0, st>>>(cls_l, cls_r, m, runs_r, W);
	return cudaGetLastError();
}


// ---- class pairs of adjacent segments with their counts (reference: greedy_matcher.cc:293-343) --------------------------
// The greedy matcher encodes, for every row, (class in the left segment, class in the right segment), radix-sorts the
// codes and counts runs: the distinct pairs in ascending (lhs, rhs) order with their multiplicities.  That list is the
// non-zero cells of the contingency table W in row-major order; one CTA streams over W and compacts them.
__global__ void __launch_bounds__(1024) pair_counts_kernel(int32_t const *W, uint32_t runs_l, uint32_t runs_r, uint32_t *lhs, uint32_t *rhs, uint32_t *count, uint32_t *n_pairs)
{
	__shared__ uint32_t wsum[32];
	__shared__ uint32_t carry;
	if (0 == threadIdx.x) carry = 0;
	__syncthreads();
	uint64_t const cells = (uint64_t) runs_l * runs_r;
	uint32_t const lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
	for (uint64_t c0 = 0; c0 < cells; c0 += 1024)
	{
		uint64_t const c = c0 + threadIdx.x;
		int32_t const v = c < cells ? W[c] : 0;
		uint32_t const b = __ballot_sync(FSQ_FULL, v > 0);
		if (0 == lane) wsum[warp] = __popc(b);
		__syncthreads();
		uint32_t before = carry;
		for (uint32_t w = 0; w < warp; ++w) before += wsum[w];
		if (v > 0)
		{
			uint32_t const at = before + __popc(b & ((1u << lane) - 1u));
			lhs[at] = (uint32_t) (c / runs_r);
			rhs[at] = (uint32_t) (c % runs_r);
			count[at] = (uint32_t) v;
		}
		__syncthreads();
		if (1023 == threadIdx.x) carry = before + __popc(b);
		__syncthreads();
	}
	if (0 == threadIdx.x) *n_pairs = carry;
}

cudaError_t launch_pair_counts(int32_t const *W, uint32_t runs_l, uint32_t runs_r, uint32_t *lhs, uint32_t *rhs, uint32_t *count, uint32_t *n_pairs, cudaStream_t st)
{
	pair_counts_kernel <<<1, 1024, 0, st>>>(W, runs_l, runs_r, lhs, rhs, count, n_pairs);
	return cudaGetLastError();
}

} // namespace fsq
