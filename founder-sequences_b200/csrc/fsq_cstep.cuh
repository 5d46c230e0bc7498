// The PBWT column step, "chunked" formulation (sm_100a): the kernel the scan passes spend their time in.
//
// One call = one step of libbio::pbwt::build_prefix_and_divergence_arrays
// (reference: lib/libbio/include/libbio/pbwt.hh:66-142) on a state that lives in shared memory as packed words
// (row id | divergence code << 16), updated in place.
//
// Work layout: thread t owns the E consecutive positions [t E, t E + E) of the prefix array and walks them
// sequentially, so every quantity the step needs inside a chunk is per-thread arithmetic:
//   * rank of an element among the same symbol   -> bit-field counters in one register
//   * max of d since the previous same symbol     -> NS running maxima in registers (the textbook PBWT update)
// Only the chunk summaries cross threads: per symbol the count (-> bucket bases; an exclusive scan over counts packed
// two to a word) and "max of d after the last occurrence" (-> carry-in; a segmented max scan done as a plain max scan
// over keys `segment index << 16 | value`, which needs no predicates).
//
// Instruction budget.  On sm_100 integer ALU work (ISETP, SEL, VIMNMX, LOP3, SHF, PRMT) and FMA-pipe work (FFMA, FADD,
// IMAD) issue through two different pipes of 16 lanes each, so a loop made of compares and selects alone runs at half
// the issue rate.  The running maxima are therefore kept as floats (codes < 2^16 are exact) and the per-symbol
// select/reset is arithmetic on a one-hot mask: oh = (sym == c) as 1.0f / 0.0f (one FSET), nm = max(rm, d) (FMNMX),
// out += nm oh (FFMA), rm = nm - nm oh (FFMA): two ALU-pipe and two FMA-pipe instructions per symbol.  Conversions
// ride on the 2^23 trick (0x4B000000 | v is the float 2^23 + v) so they are one PRMT plus one FADD.
//
// Rows beyond m are padding: they carry the largest symbol and the code 0xffff, so they stay behind every real row,
// never change (max(..., 0xffff) = 0xffff) and the loops need no bounds checks.
#pragma once
#include "fsq_common.cuh"

namespace fsq {

__device__ __forceinline__ uint32_t lds32(uint32_t addr)
{
	uint32_t v;
	asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
	return v;
}

__device__ __forceinline__ void sts32(uint32_t addr, uint32_t v)
{
	asm volatile("st.shared.u32 [%0], %1;" :: "r"(addr), "r"(v) : "memory");
}

__device__ __forceinline__ uint32_t lds8(uint32_t addr)
{
	uint32_t v;
	asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
	return v;
}

__device__ __forceinline__ void reds_add(uint32_t addr, uint32_t v)
{
	asm volatile("red.shared.add.u32 [%0], %1;" :: "r"(addr), "r"(v) : "memory");
}

// 1.0f when a == b, else 0.0f, as one FSET.BF (a C conditional is compiled back to an integer compare plus a select)
__device__ __forceinline__ float onehot(float a, float b)
{
	float r;
	asm("set.eq.f32.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
	return r;
}

// When a != b: one count moves from the histogram cell of code word `from` to that of `to` (cells are 32 bits wide at
// base + 4 code).  One branch around both updates, with the address arithmetic inside it: ptxas turns a predicated
// red.shared into a branch region of its own, which costs nine instructions per element even when nothing changes.
__device__ __forceinline__ void hist_move_if_ne(float a, float b, uint32_t to, uint32_t from, uint32_t base)
{
	asm volatile(
		"{\n"
		".reg .pred q;\n"
		".reg .b32 t0, t1;\n"
		"setp.ne.f32 q, %0, %1;\n"
		"@!q bra SKIP;\n"
		"shl.b32 t0, %2, 2;\n"
		"shl.b32 t1, %3, 2;\n"
		"add.u32 t0, t0, %4;\n"
		"add.u32 t1, t1, %4;\n"
		"red.shared.add.u32 [t0], 1;\n"
		"red.shared.add.u32 [t1], -1;\n"
		"SKIP:\n"
		"}\n" :: "f"(a), "f"(b), "r"(to), "r"(from), "r"(base) : "memory");
}

// bytes (a2, a3, b2, b3) -> result bytes 0..3: the high half of a under the high half of b
__device__ __forceinline__ uint32_t prmt_7632(uint32_t a, uint32_t b)
{
	uint32_t r;
	asm("prmt.b32 %0, %1, %2, 0x7632;" : "=r"(r) : "r"(a), "r"(b));
	return r;
}

// ---- paired formulation (FSQ_STEP_HALF2): two symbols per instruction in the halves of f16x2 operations --------------
// A divergence code c < 0x7C00 read as the bit pattern of a half is a finite non-negative number, and the order of such
// patterns as halves is their order as integers (subnormals included: none of the instructions below flushes them), so
// max, compare-equal and multiplication by 1.0 / 0.0 work on the codes as they stand.  The running maxima of symbols
// (2p, 2p + 1) share a register; per element and pair: one HSET2 (one-hot, 1.0 / 0.0 per half), one HMNMX2, two HFMA2 --
// twelve instructions for up to six symbols where the scalar form takes four per symbol.
// (FSQ_STEP_HALF2, FSQ_PAD_CODE, FSQ_CODE_LIMIT: fsq_internal.h)

__device__ __forceinline__ uint32_t h2_set_eq(uint32_t a, uint32_t b)
{
	uint32_t r;
	asm("set.eq.f16x2.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
	return r;
}

__device__ __forceinline__ uint32_t h2_max(uint32_t a, uint32_t b)
{
	uint32_t r;
	asm("max.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
	return r;
}

__device__ __forceinline__ uint32_t h2_fma(uint32_t a, uint32_t b, uint32_t c)
{
	uint32_t r;
	asm("fma.rn.f16x2 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
	return r;
}

// c - a b
__device__ __forceinline__ uint32_t h2_fnma(uint32_t a, uint32_t b, uint32_t c)
{
	uint32_t r;
	asm("{\n.reg .b32 t;\nneg.f16x2 t, %1;\nfma.rn.f16x2 %0, t, %2, %3;\n}" : "=r"(r) : "r"(a), "r"(b), "r"(c));
	return r;
}

__device__ __forceinline__ uint32_t h2_add(uint32_t a, uint32_t b)
{
	uint32_t r;
	asm("add.rn.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
	return r;
}

// One count moves from the old code's histogram cell to the new one's for each of up to three elements (old codes in
// the low halves of o*, new ones in the high halves of the rebuilt words w*; equal codes: nothing to do).
static __device__ __noinline__ void hist_moves(uint32_t H, uint32_t o0, uint32_t w0, uint32_t o1, uint32_t w1, uint32_t o2, uint32_t w2)
{
	uint32_t const od[3] = {o0 & 0xffffu, o1 & 0xffffu, o2 & 0xffffu}, nw[3] = {w0 >> 16, w1 >> 16, w2 >> 16};
#pragma unroll
	for (int i = 0; i < 3; ++i)
		if (nw[i] != od[i])
		{
			reds_add(H + 4u * nw[i], 1u);
			reds_add(H + 4u * od[i], 0xffffffffu);
		}
}

struct cstep_shared
{
	uint32_t st;        // shared-window byte address of the packed state [T E]
	uint32_t tb;        // ... of the per-thread bucket table [T][NSP]: byte address of the thread's next slot per symbol
	uint32_t *wx;       // [W][2 NS] per-warp totals and (after-value | has << 16), W > 1 only
	uint32_t magic;     // 0x4B000000 from a kernel argument (see column_step_chunked)
};

// The column's symbols reach the step as u = 6 s (one byte per row): u is the counter shift as it stands, and u & ~3
// (0, 4, 12, 16, 24, 28, 36, 40 for s = 0 .. 7: distinct and word-aligned) is the symbol's byte offset in the thread's
// bucket-table row, so neither needs arithmetic per element.
constexpr int cstep_tb_off(int c) { return (6 * c) & ~3; }
constexpr int cstep_nsp(int NS) { return (cstep_tb_off(NS - 1) / 4 + 1) | 1; }     // words per table row, odd: conflict-free

// Bit-field counters, 6 bits per symbol (E <= 63).
template <int NS, bool WIDE = (NS * 6 > 32)> struct chunk_counters;

template <int NS> struct chunk_counters <NS, false>
{
	uint32_t v{0};
	__device__ __forceinline__ uint32_t bump(uint32_t u)       // u = 6 s
	{
		uint32_t const r = (v >> u) & 63u;
		v += 1u << u;
		return r;
	}
	__device__ __forceinline__ uint32_t field(int c) const { return (v >> (c * 6)) & 63u; }
};

template <int NS> struct chunk_counters <NS, true>
{
	uint32_t lo{0}, hi{0};      // symbols 0..4 and 5..
	__device__ __forceinline__ uint32_t bump(uint32_t u)       // u = 6 s
	{
		bool const up = u >= 30u;
		uint32_t const sh = up ? u - 30u : u;
		uint32_t const one = 1u << sh;
		uint32_t const r = ((up ? hi : lo) >> sh) & 63u;
		lo += up ? 0u : one;
		hi += up ? one : 0u;
		return r;
	}
	__device__ __forceinline__ uint32_t field(int c) const { return c < 5 ? (lo >> (c * 6)) & 63u : (hi >> ((c - 5) * 6)) & 63u; }
};

// Scans of the per-lane chunk summaries across 32 lanes.
//   n[c]   count of symbol c in the lane's chunk                 -> ex[c] exclusive prefix, tot[c] total
//   af[c]  max of d after the lane's last c (whole chunk if none) -> pv[c]: max over (last lane below me that has c .. me - 1],
//          or over [0, me - 1] when no lane below has c (nohas bit c set); lane 0: 0 with nohas.
//   wkey[c] the inclusive key of lane 31: value | has-count << 16 for the whole warp
template <int NS>
__device__ __forceinline__ void lane_scans(
	uint32_t const lane, uint32_t const (&n)[NS], uint32_t const (&af)[NS],
	uint32_t (&ex)[NS], uint32_t (&tot)[NS], uint32_t (&pv)[NS], uint32_t &nohas, uint32_t (&wkey)[NS]
)
{
	constexpr int NP = (NS + 1) / 2;
	uint32_t np[2 * NP];
#pragma unroll
	for (int c = 0; c < 2 * NP; ++c) np[c] = c < NS ? n[c < NS ? c : 0] : 0u;
	uint32_t pk[NP];
#pragma unroll
	for (int i = 0; i < NP; ++i) pk[i] = np[2 * i] | (np[2 * i + 1] << 16);
#pragma unroll
	for (int o = 1; o < 32; o <<= 1)
	{
		bool const take = lane >= (uint32_t) o;
#pragma unroll
		for (int i = 0; i < NP; ++i)
		{
			uint32_t const t = __shfl_up_sync(FSQ_FULL, pk[i], o);
			pk[i] += take ? t : 0u;
		}
	}
	uint32_t exp_[2 * NP], totp[2 * NP];
#pragma unroll
	for (int i = 0; i < NP; ++i)
	{
		uint32_t const all = __shfl_sync(FSQ_FULL, pk[i], 31);
		exp_[2 * i] = (pk[i] & 0xffffu) - np[2 * i];
		exp_[2 * i + 1] = (pk[i] >> 16) - np[2 * i + 1];
		totp[2 * i] = all & 0xffffu;
		totp[2 * i + 1] = all >> 16;
	}
#pragma unroll
	for (int c = 0; c < NS; ++c) { ex[c] = exp_[c]; tot[c] = totp[c]; }

	uint32_t const le = 0xffffffffu >> (31u - lane);     // lanes <= me
	nohas = 0;
#pragma unroll
	for (int c = 0; c < NS; ++c)
	{
		uint32_t const hm = __ballot_sync(FSQ_FULL, 0 != n[c]);
		uint32_t key = ((uint32_t) __popc(hm & le) << 16) | af[c];
#pragma unroll
		for (int o = 1; o < 32; o <<= 1)
			key = max(key, __shfl_up_sync(FSQ_FULL, key, o));     // lanes below o get their own key back: a no-op
		uint32_t const p = __shfl_up_sync(FSQ_FULL, key, 1);
		pv[c] = lane ? (p & 0xffffu) : 0u;
		if (0 == lane || 0 == (p >> 16)) nohas |= 1u << c;
		wkey[c] = __shfl_sync(FSQ_FULL, key, 31);
	}
}

// The step.  NS symbols s = 0 .. NS - 1 (already offset by the smallest symbol that occurs), E elements per thread,
// blockDim.x threads.  sym8: shared-window byte address of this column's symbols as u = 6 s, one byte per row id
// (and 6 (NS - 1) for the padding ids).  HIST: maintain the divergence-code histogram H (32-bit counters).
template <int E, int NS, bool HIST>
__device__ __forceinline__ void column_step_chunked(
	cstep_shared const &sh,
	uint32_t const sym8,
	uint32_t const first_marker,   // code written for the first element of every symbol (k + 1)
	uint32_t const H               // shared-window byte address of the code histogram (HIST)
)
{
	static_assert(E <= 63, "counters are 6 bits wide");
	constexpr int NSP = cstep_nsp(NS);
	uint32_t const tid = threadIdx.x;
	uint32_t const lane = tid & 31u;
	uint32_t const warp = tid >> 5;
	uint32_t const n_warps = blockDim.x >> 5;
	uint32_t const mine = sh.st + tid * (E * 4u);
	uint32_t const mytb = sh.tb + tid * (NSP * 4u);

	uint32_t wd[E];     // row id | out << 16
	uint32_t mt[E];     // u | local rank << 8
	// 0x4B00hhll is the float 2^23 + code.  The maxima live in that biased form (0 = "nothing seen": below every
	// biased value), so a code goes in with one PRMT and comes out with none.  The constant comes from a kernel argument:
	// as a literal, ptxas gives it the PRMT's immediate slot and re-materialises the selector in a register per element.
	uint32_t const magic = sh.magic;

	// ---- phase 1: walk the chunk -------------------------------------------------------------------------------
	chunk_counters <NS> cnt;
#if !FSQ_STEP_HALF2
	float rm[NS];
#pragma unroll
	for (int c = 0; c < NS; ++c) rm[c] = 0.0f;
#endif

	// The chunk's words first, then their symbols, each as one batch of independent loads: the loads are volatile asm and
	// keep their program order, so interleaving them with the arithmetic (load word, load symbol, compute, next element)
	// put two shared-memory latencies on every element's critical path (ncu r2_v0: short-scoreboard 15-18 % of samples).
#pragma unroll
	for (int j = 0; j < E; ++j) wd[j] = lds32(mine + 4u * j);
#pragma unroll
	for (int j = 0; j < E; ++j) mt[j] = lds8(sym8 + (wd[j] & 0xffffu));           // 6 s

	// Histogram moves are rare (a divergence changes only when an element's predecessor of the same symbol is not its
	// neighbour): one test per HG elements instead of a branch per element.
	constexpr int HG = 3;
	uint32_t chg = 0, h_old[HG], h_new[HG];
#if FSQ_STEP_HALF2
	constexpr int NP2 = (NS + 1) / 2;
	uint32_t rm2[NP2];
#pragma unroll
	for (int p = 0; p < NP2; ++p) rm2[p] = 0u;
#pragma unroll
	for (int j = 0; j < E; ++j)
	{
		uint32_t const w = wd[j];
		uint32_t const u = mt[j];
		uint32_t const dd = __byte_perm(w, w, 0x3232);               // the code in both halves
		uint32_t const uu = __byte_perm(u, 0u, 0x0404);              // u << 8 in both halves: distinct finite halves for distinct symbols
		uint32_t const lr = cnt.bump(u);
		uint32_t out2 = 0u;
#pragma unroll
		for (int p = 0; p < NP2; ++p)
		{
			// symbols 2 p and 2 p + 1; a pair's second half without a symbol compares against a byte no symbol has
			uint32_t const want = ((uint32_t) (6 * (2 * p)) << 8) | ((2 * p + 1 < NS ? (uint32_t) (6 * (2 * p + 1)) : 0x7bu) << 24);
			uint32_t const oh = h2_set_eq(uu, want);
			uint32_t const nm = h2_max(rm2[p], dd);
			out2 = h2_fma(nm, oh, out2);
			rm2[p] = h2_fnma(nm, oh, nm);
		}
		// exactly one half of out2 holds the element's new code, the other is +0
		uint32_t const ob = h2_add(out2, __byte_perm(out2, out2, 0x1032));
		wd[j] = __byte_perm(w, ob, 0x5410);
		mt[j] = lr * 256u + u;
		if (HIST)
		{
			h_old[j % HG] = dd;
			chg |= dd ^ ob;
			if (HG - 1 == j % HG || E - 1 == j)
			{
				if (chg)
				{
					// rare, and out of line: eleven inlined copies of it were a tenth of the loop body, which is fetched from
					// beyond the instruction cache as it is (the new codes are read back from the words just built)
					hist_moves(H, h_old[0], wd[j - j % HG], j % HG >= 1 ? h_old[1] : 0u, j % HG >= 1 ? wd[j - j % HG + 1] : 0u,
						j % HG >= 2 ? h_old[2] : 0u, j % HG >= 2 ? wd[j - j % HG + 2] : 0u);
				}
				chg = 0;
			}
		}
	}
#else
#pragma unroll
	for (int j = 0; j < E; ++j)
	{
		uint32_t const w = wd[j];
		uint32_t const u = mt[j];
		uint32_t const db = prmt_7632(w, magic);                     // 0x4B00hhll
		float const dvf = __uint_as_float(db);
		float const uf = (float) u;
		uint32_t const lr = cnt.bump(u);
		float out = 0.0f;
#pragma unroll
		for (int c = 0; c < NS; ++c)
		{
			float const oh = onehot(uf, (float) (6 * c));
			float const nm = fmaxf(rm[c], dvf);
			out = fmaf(nm, oh, out);
			rm[c] = fmaf(-nm, oh, nm);
		}
		uint32_t const ob = __float_as_uint(out);                    // 0x4B00hhll
		wd[j] = __byte_perm(w, ob, 0x5410);
		mt[j] = lr * 256u + u;
		if (HIST)
		{
			h_old[j % HG] = db; h_new[j % HG] = ob;
			chg |= db ^ ob;
			if (HG - 1 == j % HG || E - 1 == j)
			{
				if (chg)
				{
					// one count moves from the old code to the new one; 4 (0x4B000000 + c) = 0x2C000000 + 4 c (mod 2^32)
					uint32_t const Hb = H - 0x2C000000u;
#pragma unroll
					for (int i = 0; i <= j % HG; ++i)
						if (h_old[i] != h_new[i])
						{
							reds_add(Hb + 4u * h_new[i], 1u);
							reds_add(Hb + 4u * h_old[i], 0xffffffffu);
						}
				}
				chg = 0;
			}
		}
	}
#endif

	// ---- chunk summaries across the warp (and the CTA) ----------------------------------------------------------
	uint32_t n[NS], af[NS];
	{
#pragma unroll
		for (int c = 0; c < NS; ++c)
		{
			n[c] = cnt.field(c);
#if FSQ_STEP_HALF2
			af[c] = (c & 1) ? (rm2[c / 2] >> 16) : (rm2[c / 2] & 0xffffu);    // 0 when nothing was seen
#else
			af[c] = __float_as_uint(fmaxf(rm[c], 8388608.0f)) & 0xffffu;      // 0 when nothing was seen
#endif
		}
	}
	uint32_t ex[NS], tot[NS], pv[NS], wkey[NS], nohas;
	lane_scans <NS>(lane, n, af, ex, tot, pv, nohas, wkey);

	uint32_t wex[NS], wcin[NS];     // this warp's exclusive counts and carry-in
	if (n_warps > 1)
	{
		if (31 == lane)
		{
#pragma unroll
			for (int c = 0; c < NS; ++c)
			{
				sh.wx[warp * (2 * NS) + c] = tot[c];
				sh.wx[warp * (2 * NS) + NS + c] = wkey[c];
			}
		}
		__syncthreads();        // also orders every thread's phase-1 loads before any scatter
		uint32_t n2[NS], af2[NS];
		bool const in = lane < n_warps;
#pragma unroll
		for (int c = 0; c < NS; ++c)
		{
			uint32_t const k = in ? sh.wx[lane * (2 * NS) + NS + c] : 0u;
			n2[c] = in ? sh.wx[lane * (2 * NS) + c] : 0u;
			af2[c] = k & 0xffffu;
			// lane_scans() derives "has" from a non-zero count; a warp has c exactly when its count of c is non-zero
		}
		uint32_t ex2[NS], tot2[NS], pv2[NS], wkey2[NS], nohas2;
		lane_scans <NS>(lane, n2, af2, ex2, tot2, pv2, nohas2, wkey2);
#pragma unroll
		for (int c = 0; c < NS; ++c)
		{
			uint32_t const cin = ((nohas2 >> c) & 1u) ? max(pv2[c], first_marker) : pv2[c];
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

	uint32_t slot[NS], cin[NS];
	{
		uint32_t base = 0;
#pragma unroll
		for (int c = 0; c < NS; ++c)
		{
			slot[c] = sh.st + 4u * (base + wex[c] + ex[c]);
			cin[c] = ((nohas >> c) & 1u) ? max(pv[c], wcin[c]) : pv[c];
			sts32(mytb + (uint32_t) cstep_tb_off(c), slot[c]);
			base += tot[c];
		}
	}
	// table rows are private to the thread that wrote them: no barrier needed before reading them back

	// ---- phase 2: scatter ---------------------------------------------------------------------------------------
	// (bucket bases in batches ahead of the stores that use them: see the note on the loads of phase 1)
	constexpr int PB = 11;
#pragma unroll
	for (int j0 = 0; j0 < E; j0 += PB)
	{
		uint32_t base[PB];
#pragma unroll
		for (int i = 0; i < PB; ++i)
			if (j0 + i < E) base[i] = lds32(mytb + (mt[j0 + i] & 0x3cu));
#pragma unroll
		for (int i = 0; i < PB; ++i)
			if (j0 + i < E) sts32(base[i] + (mt[j0 + i] >> 6), wd[j0 + i]);
	}

	// ---- the first element of each symbol in the chunk takes the carry-in ------------------------------------------
#pragma unroll
	for (int c = 0; c < NS; ++c)
	{
		if (n[c])
		{
			uint32_t const w = lds32(slot[c]);
			uint32_t const o = w >> 16;
			uint32_t const o2 = max(o, cin[c]);
			if (o2 != o)
			{
				sts32(slot[c], (w & 0xffffu) | (o2 << 16));
				if (HIST)
				{
					reds_add(H + 4u * o2, 1u);
					reds_add(H + 4u * o, 0xffffffffu);
				}
			}
		}
	}
}

} // namespace fsq
