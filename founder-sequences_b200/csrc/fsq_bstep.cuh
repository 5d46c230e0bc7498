// The PBWT column step, "blocked" formulation (sm_100a): the kernel the scan passes spend their time in.
//
// One call = one step of libbio::pbwt::build_prefix_and_divergence_arrays
// (reference: lib/libbio/include/libbio/pbwt.hh:66-142) on a state that lives in shared memory as packed words
// (row id | divergence code << 16), updated in place.
//
// Work layout: thread t owns the E consecutive positions [t E, t E + E) of the prefix array and walks them
// sequentially, so every quantity the step needs inside a chunk is plain per-thread arithmetic:
//   * rank of an element among the same symbol      -> bit-field counters in one (or two) registers
//   * max of d since the previous same symbol        -> NS running maxima in registers (the textbook PBWT update)
// Only the chunk summaries cross threads: per symbol the count (-> bucket bases, exclusive scan) and
// "max of d after the last occurrence" with a has-flag (-> carry-in, a segmented max scan); both are warp
// shuffles, plus one shared-memory exchange when the CTA has more than one warp.  E is odd, so the E loads of a
// thread (stride E words between lanes) are bank-conflict free without padding.
//
// A warp instruction here advances 32 elements by one operation (about 50 instructions per element in total),
// where the striped formulation (fsq_step.cuh, still used by the stitch) spends about 300 warp instructions per 32
// elements on ballots and shuffles.
#pragma once
#include "fsq_common.cuh"

namespace fsq {

struct bstep_shared
{
	uint32_t *st;       // [T E] packed state: row id | code << 16
	uint32_t *tb;       // [NS][T] bucket base of each thread, per symbol
	uint32_t *tc;       // [NS][T] carry-in of each thread, per symbol
	uint32_t *wcnt;     // [W][NS] per-warp symbol counts
	uint32_t *wcar;     // [W][NS] per-warp "max after the last occurrence"
	uint32_t *whas;     // [W] bitmask: symbol occurs in the warp's chunks
};

// Counters: one bit field per symbol, wide enough for E.
template <int E> struct field_bits { static constexpr int value = E < 16 ? 4 : (E < 32 ? 5 : (E < 64 ? 6 : 7)); };

template <int NS, int FB, bool WIDE = (NS * FB > 32)> struct sym_counters;

template <int NS, int FB> struct sym_counters <NS, FB, false>
{
	uint32_t v{0};
	__device__ __forceinline__ uint32_t get(uint32_t s) const { return (v >> (s * FB)) & ((1u << FB) - 1u); }
	__device__ __forceinline__ void add(uint32_t s, uint32_t one) { v += one << (s * FB); }
	template <int C> __device__ __forceinline__ uint32_t field() const { return (v >> (C * FB)) & ((1u << FB) - 1u); }
};

template <int NS, int FB> struct sym_counters <NS, FB, true>
{
	unsigned long long v{0};
	__device__ __forceinline__ uint32_t get(uint32_t s) const { return (uint32_t) (v >> (s * FB)) & ((1u << FB) - 1u); }
	__device__ __forceinline__ void add(uint32_t s, uint32_t one) { v += (unsigned long long) one << (s * FB); }
	template <int C> __device__ __forceinline__ uint32_t field() const { return (uint32_t) (v >> (C * FB)) & ((1u << FB) - 1u); }
};

// Scan of per-lane chunk summaries across a warp, stage A (independent of what comes in from the left):
//   ex[c]    exclusive prefix of the counts          tot[c]  warp total
//   prev[c]  segmented max of `af` over the lanes (last lane below me that has c .. me - 1]; with no such lane: over
//            lanes [0, me - 1] -- to be combined with the carry-in by the caller (nohas bit set)
//   waf[c]   the same over the whole warp (lane 31 inclusive); whas: symbols that occur in the warp
template <int NS>
__device__ __forceinline__ void lane_scan(
	uint32_t const lane, uint32_t const (&n)[NS], uint32_t const (&af)[NS], uint32_t const hs,
	uint32_t (&ex)[NS], uint32_t (&tot)[NS], uint32_t (&prev)[NS], uint32_t &nohas, uint32_t (&waf)[NS], uint32_t &whas
)
{
	nohas = 0; whas = 0;
	uint32_t const le = 0xffffffffu >> (31u - lane);     // lanes <= me
#pragma unroll
	for (int c = 0; c < NS; ++c)
	{
		uint32_t v = n[c];
#pragma unroll
		for (int o = 1; o < 32; o <<= 1)
		{
			uint32_t const t = __shfl_up_sync(FSQ_FULL, v, o);
			if (lane >= (uint32_t) o) v += t;
		}
		ex[c] = v - n[c];
		tot[c] = __shfl_sync(FSQ_FULL, v, 31);

		uint32_t const hm = __ballot_sync(FSQ_FULL, (hs >> c) & 1u);
		uint32_t const bi = hm & le;
		uint32_t const seg = bi ? 31u - __clz(bi) : 0u;  // the inclusive scan at this lane starts here
		uint32_t s = af[c];
#pragma unroll
		for (int o = 1; o < 32; o <<= 1)
		{
			uint32_t const t = __shfl_up_sync(FSQ_FULL, s, o);
			if (lane >= (uint32_t) o + seg) s = max(s, t);
		}
		uint32_t const p = __shfl_up_sync(FSQ_FULL, s, 1);
		prev[c] = lane ? p : 0u;
		if (0 == (hm & (le >> 1))) nohas |= 1u << c;     // no lane below me has c
		waf[c] = __shfl_sync(FSQ_FULL, s, 31);
		if (hm) whas |= 1u << c;
	}
}

// The step.  NS symbols (values sym0 .. sym0 + NS - 1 as produced by symf), E elements per thread, T = blockDim.x
// threads with T E >= m.  HIST: maintain the divergence-code histogram H (16-bit counters, two per word).
template <int E, int NS, bool HIST, typename SymF>
__device__ __forceinline__ void column_step_blocked(
	bstep_shared const &sh,
	uint32_t const m,
	uint32_t const sym0,
	uint32_t const first_marker,   // code written for the first element of every symbol (k + 1)
	SymF symf,
	uint32_t *H
)
{
	constexpr int FB = field_bits <E>::value;
	uint32_t const tid = threadIdx.x;
	uint32_t const T = blockDim.x;
	uint32_t const lane = tid & 31u;
	uint32_t const warp = tid >> 5;
	uint32_t const n_warps = T >> 5;
	uint32_t const i0 = tid * E;

	uint32_t wd[E];     // row id | out << 16
	uint32_t meta[E];   // sym | local rank << 4 | old code << 16

	// ---- phase 1: walk the chunk -------------------------------------------------------------------------------
	sym_counters <NS, FB> cnt;
	uint32_t rm[NS];
#pragma unroll
	for (int c = 0; c < NS; ++c) rm[c] = 0;

#pragma unroll
	for (int j = 0; j < E; ++j)
	{
		uint32_t const i = i0 + j;
		bool const valid = i < m;
		uint32_t const w = sh.st[i];
		uint32_t const x = w & 0xffffu;
		uint32_t const dv = valid ? (w >> 16) : 0u;
		uint32_t const sym = valid ? symf(x) - sym0 : 0u;
		uint32_t const lr = cnt.get(sym);
		cnt.add(sym, valid ? 1u : 0u);
		uint32_t out = 0;
#pragma unroll
		for (int c = 0; c < NS; ++c)
		{
			bool const is = sym == (uint32_t) c;
			uint32_t const nm = max(rm[c], dv);
			out = is ? nm : out;
			rm[c] = is ? 0u : nm;
		}
		wd[j] = x | (out << 16);
		meta[j] = sym | (lr << 4) | (dv << 16);
	}

	// ---- chunk summaries across the warp (and the CTA) ----------------------------------------------------------
	uint32_t n[NS], hs = 0;
	{
		uint32_t tmp[NS];
		// unpack the counters (compile-time field positions)
#pragma unroll
		for (int c = 0; c < NS; ++c) tmp[c] = cnt.get((uint32_t) c);
#pragma unroll
		for (int c = 0; c < NS; ++c) { n[c] = tmp[c]; if (tmp[c]) hs |= 1u << c; }
	}
	uint32_t ex[NS], tot[NS], prev[NS], waf[NS], nohas, whas;
	lane_scan <NS>(lane, n, rm, hs, ex, tot, prev, nohas, waf, whas);

	uint32_t wex[NS], wcin[NS];     // this warp's exclusive counts and carry-in
	if (n_warps > 1)
	{
		if (31 == lane)
		{
#pragma unroll
			for (int c = 0; c < NS; ++c)
			{
				sh.wcnt[warp * NS + c] = tot[c];
				sh.wcar[warp * NS + c] = waf[c];
			}
			sh.whas[warp] = whas;
		}
		__syncthreads();        // also orders every thread's phase-1 loads before any scatter
		uint32_t n2[NS], af2[NS], hs2 = 0;
		bool const in = lane < n_warps;
#pragma unroll
		for (int c = 0; c < NS; ++c)
		{
			n2[c] = in ? sh.wcnt[lane * NS + c] : 0u;
			af2[c] = in ? sh.wcar[lane * NS + c] : 0u;
		}
		hs2 = in ? sh.whas[lane] : 0u;
		uint32_t ex2[NS], tot2[NS], prev2[NS], waf2[NS], nohas2, whas2;
		lane_scan <NS>(lane, n2, af2, hs2, ex2, tot2, prev2, nohas2, waf2, whas2);
#pragma unroll
		for (int c = 0; c < NS; ++c)
		{
			uint32_t const cin = ((nohas2 >> c) & 1u) ? max(prev2[c], first_marker) : prev2[c];
			wex[c] = __shfl_sync(FSQ_FULL, ex2[c], warp);
			wcin[c] = __shfl_sync(FSQ_FULL, cin, warp);
			tot[c] = tot2[c];
		}
	}
	else
	{
#pragma unroll
		for (int c = 0; c < NS; ++c) { wex[c] = 0; wcin[c] = first_marker; }
		__syncwarp();
	}

	{
		uint32_t base = 0;
#pragma unroll
		for (int c = 0; c < NS; ++c)
		{
			sh.tb[c * T + tid] = base + wex[c] + ex[c];
			sh.tc[c * T + tid] = ((nohas >> c) & 1u) ? max(prev[c], wcin[c]) : prev[c];
			base += tot[c];
		}
	}
	// tb / tc entries are private to the thread that wrote them: no barrier needed before reading them back.

	// ---- phase 2: scatter ---------------------------------------------------------------------------------------
#pragma unroll
	for (int j = 0; j < E; ++j)
	{
		uint32_t const i = i0 + j;
		uint32_t const mt = meta[j];
		uint32_t const sym = mt & 15u;
		uint32_t const lr = (mt >> 4) & 0xfffu;
		uint32_t const dst = sh.tb[sym * T + tid] + lr;
		uint32_t word = wd[j];
		if (0 == lr)
		{
			uint32_t const cv = sh.tc[sym * T + tid];
			uint32_t const o = max(word >> 16, cv);
			word = (word & 0xffffu) | (o << 16);
		}
		if (i < m)
		{
			sh.st[dst] = word;
			if (HIST)
			{
				uint32_t const out = word >> 16, dv = mt >> 16;
				if (out != dv)
				{
					atomicAdd(&H[out >> 1], (out & 1u) ? 0x10000u : 1u);
					atomicSub(&H[dv >> 1], (dv & 1u) ? 0x10000u : 1u);
				}
			}
		}
	}
	__syncthreads();
}

} // namespace fsq
