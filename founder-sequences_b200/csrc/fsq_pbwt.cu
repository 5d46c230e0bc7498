// PBWT column-block kernels (sm_100a): local scan (pass A), exact scan with divergence histogram and
// per-column top lists (pass B), re-scan to requested columns (pass C), rank compression of a block's
// start state, and the boundary stitch.  See DESIGN.md for the scheme and fsq_step.cuh for the column step.
#include "fsq_internal.h"
#include "fsq_step.cuh"
#include "fsq_scan_common.cuh"

namespace fsq {

template <int TPW, bool HIST>
__global__ void __launch_bounds__(1024) pbwt_block_kernel(block_args const A)
{
	extern __shared__ __align__(128) uint8_t smem_raw[];
	uint32_t const tid = threadIdx.x;
	uint32_t const n_warps = blockDim.x >> 5;
	uint32_t const warp = tid >> 5;
	uint32_t const m = A.m;

	uint32_t const b = (MODE_DUMP == A.mode) ? A.blk_ids[blockIdx.x] : blockIdx.x + A.first_block;
	uint64_t const s = (uint64_t) b * A.C;                          // local column index of the block start
	uint64_t e = s + A.C < A.n_local ? s + A.C : A.n_local;
	uint64_t const s_glob = A.col0 + s;

	uint32_t dump_next = 0, dump_end = 0;
	if (MODE_DUMP == A.mode)
	{
		dump_next = A.dump_begin[b];
		dump_end = A.dump_begin[b + 1];
		if (dump_next == dump_end) return;
		e = A.dump_cols[dump_end - 1];                                // last right bound needed in this block
	}

	// ---- carve shared memory (mirrors block_smem_bytes()) ----
	size_t off = 0;
	auto carve = [&](size_t bytes) { void *p = smem_raw + off; off += (bytes + 15) & ~size_t(15); return p; };
	uint8_t *colbuf = (uint8_t *) carve(2 * (size_t) A.cg * A.col_stride);
	uint64_t *mbar = (uint64_t *) carve(2 * sizeof(uint64_t));
	step_shared sh;
	sh.sa = (uint16_t *) carve(sizeof(uint16_t) * m);
	sh.sd = carve(sizeof(uint16_t) * m);
	sh.wcnt = (uint32_t *) carve(sizeof(uint32_t) * n_warps * FSQ_MAX_SYMS);
	sh.wcar = (uint32_t *) carve(sizeof(uint32_t) * n_warps * FSQ_MAX_SYMS);
	sh.whas = (uint32_t *) carve(sizeof(uint32_t) * n_warps);
	uint32_t *H = HIST ? (uint32_t *) carve(sizeof(uint32_t) * ((m + A.C + 2) / 2 + 1)) : nullptr;
	uint16_t *sd = (uint16_t *) sh.sd;

	// ---- start state ----
	uint32_t R = 1;
	uint32_t const *old = nullptr;
	if (MODE_LOCAL == A.mode)
	{
		for (uint32_t i = tid; i < m; i += blockDim.x) { sh.sa[i] = (uint16_t) i; sd[i] = 0; }
	}
	else
	{
		R = A.R[b];
		old = A.oldvals + (size_t) b * m;
		uint32_t const *a0 = A.st_a + (size_t) b * m;
		uint16_t const *c0 = A.code + (size_t) b * m;
		for (uint32_t i = tid; i < m; i += blockDim.x) { sh.sa[i] = (uint16_t) a0[i]; sd[i] = c0[i]; }
	}
	if (HIST)
	{
		uint32_t const words = (m + A.C + 2) / 2 + 1;
		for (uint32_t i = tid; i < words; i += blockDim.x) H[i] = 0;
	}
	if (0 == tid)
	{
		mbar_init(&mbar[0], 1);
		mbar_init(&mbar[1], 1);
		mbar_fence_init();
	}
	__syncthreads();
	if (HIST)
	{
		for (uint32_t i = tid; i < m; i += blockDim.x)
		{
			uint32_t const c = sd[i];
			atomicAdd(&H[c >> 1], (c & 1u) ? 0x10000u : 1u);
		}
	}
	bool const zero_present = (MODE_LOCAL == A.mode) ? (0 == s_glob) : (0 == old[0]);

	auto decode = [&](uint32_t c) -> uint32_t {
		if (c >= R) return (uint32_t) (s_glob + 1 + (c - R));
		return (MODE_LOCAL == A.mode) ? (uint32_t) s_glob : old[c];
	};

	// ---- column loop; columns arrive through a two-stage TMA ring of cg columns each ----
	uint64_t const n_cols = e - s;
	uint32_t const n_stages = (uint32_t) ((n_cols + A.cg - 1) / A.cg);
	auto issue = [&](uint32_t g) {
		uint64_t const c_first = s + (uint64_t) g * A.cg;
		uint64_t const c_cnt = (c_first + A.cg <= e) ? A.cg : e - c_first;
		uint32_t const bytes = (uint32_t) (c_cnt * A.col_stride);
		mbar_expect_tx(&mbar[g & 1], bytes);
		tma_load_1d(colbuf + (size_t) (g & 1) * A.cg * A.col_stride, A.cols + c_first * A.col_stride, bytes, &mbar[g & 1]);
	};
	if (0 == tid && n_stages) issue(0);

	for (uint64_t j = 0; j < n_cols; ++j)
	{
		uint32_t const g = (uint32_t) (j / A.cg);
		uint32_t const jj = (uint32_t) (j % A.cg);
		if (0 == jj)
		{
			if (0 == tid && g + 1 < n_stages) issue(g + 1);
			mbar_wait(&mbar[g & 1], (g >> 1) & 1u);
		}
		if (HIST && j > 0 && warp == n_warps - 1)
		{
			uint64_t const kk = s + j - 1;
			emit_top_list(H, R + (uint32_t) (j - 1), R, old, s_glob, m, A.cap,
				A.toplist + (kk - A.list_col0) * (A.cap + 1), A.cnt + kk, A.h0 + kk, zero_present);
		}
		nibble_source src{(uint32_t const *) (colbuf + ((size_t) (g & 1) * A.cg + jj) * A.col_stride)};
		column_step <uint16_t, TPW, HIST>(sh, m, A.n_tiles, A.nsym, R + (uint32_t) j, src, H);

		if (MODE_DUMP == A.mode)
		{
			while (dump_next < dump_end && A.dump_cols[dump_next] == s + j + 1)
			{
				uint32_t *oa = A.dump_a + (size_t) dump_next * m;
				uint32_t *od = A.dump_d + (size_t) dump_next * m;
				for (uint32_t i = tid; i < m; i += blockDim.x) { oa[i] = sh.sa[i]; od[i] = decode(sd[i]); }
				++dump_next;
			}
		}
	}
	if (HIST && n_cols && warp == n_warps - 1)
	{
		uint64_t const kk = s + n_cols - 1;
		emit_top_list(H, R + (uint32_t) (n_cols - 1), R, old, s_glob, m, A.cap,
			A.toplist + (kk - A.list_col0) * (A.cap + 1), A.cnt + kk, A.h0 + kk, zero_present);
	}
	if (MODE_LOCAL == A.mode)
	{
		uint32_t *oa = A.sum_a + (size_t) b * m;
		uint32_t *od = A.sum_d + (size_t) b * m;
		for (uint32_t i = tid; i < m; i += blockDim.x) { oa[i] = sh.sa[i]; od[i] = decode(sd[i]); }
	}
}


size_t block_smem_bytes(block_args const &A, uint32_t n_warps, bool hist)
{
	auto r16 = [](size_t b) { return (b + 15) & ~size_t(15); };
	size_t t = 0;
	t += r16(2 * (size_t) A.cg * A.col_stride);
	t += r16(2 * sizeof(uint64_t));
	t += r16(sizeof(uint16_t) * A.m) * 2;
	t += r16(sizeof(uint32_t) * n_warps * FSQ_MAX_SYMS) * 2;
	t += r16(sizeof(uint32_t) * n_warps);
	if (hist) t += r16(sizeof(uint32_t) * ((A.m + A.C + 2) / 2 + 1));
	return t;
}

template <int TPW, bool HIST>
static cudaError_t launch_block_t(block_args const &A, uint32_t n_blocks, uint32_t n_warps, cudaStream_t st)
{
	size_t const smem = block_smem_bytes(A, n_warps, HIST);
	auto kern = pbwt_block_kernel <TPW, HIST>;
	cudaError_t err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
	if (cudaSuccess != err) return err;
	kern <<<n_blocks, n_warps * 32, smem, st>>>(A);
	return cudaGetLastError();
}

// Tiles per warp are a compile-time bound (the state is staged in registers); pick the smallest instantiation.
cudaError_t launch_block(block_args const &A, uint32_t n_blocks, uint32_t n_warps, uint32_t tpw, bool hist, cudaStream_t st)
{
#define FSQ_CASE(T) if (tpw <= T) return hist ? launch_block_t <T, true>(A, n_blocks, n_warps, st) : launch_block_t <T, false>(A, n_blocks, n_warps, st);
	FSQ_CASE(2) FSQ_CASE(4) FSQ_CASE(8) FSQ_CASE(16) FSQ_CASE(24) FSQ_CASE(32)
#undef FSQ_CASE
	return cudaErrorInvalidValue;
}


// ---- rank compression of a block's start divergences ------------------------------------------------------------
// d values at a block start are arbitrary positions <= s; the scan kernels keep them as 16-bit order-preserving
// codes (rank among the distinct start values), which double as histogram indices.
__global__ void __launch_bounds__(1024) prep_kernel(uint32_t const *st_d, uint32_t m, uint32_t p2, uint16_t *code, uint32_t *oldvals, uint32_t *R, uint32_t first_block)
{
	extern __shared__ __align__(128) uint8_t smem_raw[];
	uint32_t *buf = (uint32_t *) smem_raw;                  // [p2]
	uint16_t *rank = (uint16_t *) (buf + p2);               // [p2]
	__shared__ uint32_t scratch[34];
	uint32_t const b = blockIdx.x + first_block;
	uint32_t const tid = threadIdx.x;
	uint32_t const *d = st_d + (size_t) b * m;

	if (p2 == blockDim.x)
	{
		// one value per thread (m <= 1024): the network runs in registers below 32 lanes apart and through shared memory above;
		// a value takes its row along, so that the row's code is written straight from the sorted position
		uint32_t *xi = (uint32_t *) (rank + p2);             // [p2] the rows, exchanged beside the values
		uint32_t v = tid < m ? d[tid] : 0xffffffffu, row = tid;
		for (uint32_t k = 2; k <= p2; k <<= 1)
			for (uint32_t j = k >> 1; j > 0; j >>= 1)
			{
				uint32_t ov, orow;
				if (j >= 32)
				{
					__syncthreads();                                // the exchange before this one has been read
					buf[tid] = v; xi[tid] = row;
					__syncthreads();
					ov = buf[tid ^ j]; orow = xi[tid ^ j];
				}
				else
				{
					ov = __shfl_xor_sync(FSQ_FULL, v, j);
					orow = __shfl_xor_sync(FSQ_FULL, row, j);
				}
				bool const keep_min = (0 == (tid & j)) == (0 == (tid & k));
				// (equal values: both sides keep their own, so no row is lost or doubled)
				if (keep_min ? ov < v : ov > v) { v = ov; row = orow; }
			}
		__syncthreads();
		buf[tid] = v;
		__syncthreads();
		bool const head = tid < m && (0 == tid || buf[tid - 1] != v);
		uint32_t total;
		uint32_t const pos = block_excl_scan(head ? 1u : 0u, scratch, total);      // distinct values before this position
		if (head) oldvals[(size_t) b * m + pos] = v;
		if (tid < m) code[(size_t) b * m + row] = (uint16_t) (pos + (head ? 1u : 0u) - 1u);
		if (0 == tid) R[b] = total;
		return;
	}
	for (uint32_t i = tid; i < p2; i += blockDim.x) buf[i] = i < m ? d[i] : 0xffffffffu;
	__syncthreads();
	for (uint32_t k = 2; k <= p2; k <<= 1)
		for (uint32_t j = k >> 1; j > 0; j >>= 1)
		{
			for (uint32_t i = tid; i < p2; i += blockDim.x)
			{
				uint32_t const l = i ^ j;
				if (l > i)
				{
					uint32_t const x = buf[i], y = buf[l];
					bool const up = 0 == (i & k);
					if ((x > y) == up) { buf[i] = y; buf[l] = x; }
				}
			}
			__syncthreads();
		}

	// distinct values → ranks
	uint32_t const per = (p2 + blockDim.x - 1) / blockDim.x;
	uint32_t const lo = tid * per, hi = min(lo + per, m);
	uint32_t cnt = 0;
	for (uint32_t i = lo; i < hi; ++i) cnt += (0 == i || buf[i] != buf[i - 1]);
	uint32_t total;
	uint32_t pos = block_excl_scan(cnt, scratch, total);
	uint32_t *ov = oldvals + (size_t) b * m;
	for (uint32_t i = lo; i < hi; ++i)
	{
		if (0 == i || buf[i] != buf[i - 1]) ov[pos++] = buf[i];
		rank[i] = (uint16_t) (pos - 1);
	}
	if (0 == tid) R[b] = total;
	__syncthreads();
	uint16_t *co = code + (size_t) b * m;
	for (uint32_t i = tid; i < m; i += blockDim.x)
	{
		uint32_t const v = d[i];
		uint32_t l = 0, h = m;          // lower bound in buf[0..m)
		while (l < h) { uint32_t const mid = (l + h) >> 1; if (buf[mid] < v) l = mid + 1; else h = mid; }
		co[i] = rank[l];
	}
}

cudaError_t launch_prep(uint32_t const *st_d, uint32_t m, uint16_t *code, uint32_t *oldvals, uint32_t *R, uint32_t first_block, uint32_t n_blocks, cudaStream_t st)
{
	uint32_t p2 = 32;
	while (p2 < m) p2 <<= 1;
	size_t const smem = (size_t) p2 * 6 + (p2 <= 1024 ? (size_t) p2 * 4 : 0);
	cudaError_t err = cudaFuncSetAttribute(prep_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
	if (cudaSuccess != err) return err;
	uint32_t const threads = p2 < 1024 ? p2 : 1024;
	prep_kernel <<<n_blocks, threads, smem, st>>>(st_d, m, p2, code, oldvals, R, first_block);
	return cudaGetLastError();
}


// ---- boundary stitch ------------------------------------------------------------------------------
// true state at e  =  one generalised PBWT step on the true state at s whose "symbol" is the class of the row in the
// local summary (rows with identical columns [s, e) form a class, classes in local PBWT order).  The class id is
// fed to the ordinary 16-way column step one hex digit at a time, least significant first (a stable LSD radix sort,
// i.e. up to four virtual columns); divergences of class starts are then taken from the summary.
struct digit_source
{
	uint16_t const *cls;
	uint32_t shift;
	__device__ __forceinline__ uint32_t operator()(uint32_t x) const { return (cls[x] >> shift) & 15u; }
};

template <int TPW>
__global__ void __launch_bounds__(1024) stitch_kernel(stitch_args const A)
{
	extern __shared__ __align__(128) uint8_t smem_raw[];
	__shared__ uint32_t scratch[34];
	uint32_t const tid = threadIdx.x;
	uint32_t const n_warps = blockDim.x >> 5;
	uint32_t const m = A.m;
	size_t off = 0;
	auto carve = [&](size_t bytes) { void *p = smem_raw + off; off += (bytes + 15) & ~size_t(15); return p; };
	step_shared sh;
	sh.sa = (uint16_t *) carve(sizeof(uint16_t) * m);
	sh.sd = carve(sizeof(uint32_t) * m);
	uint16_t *cls = (uint16_t *) carve(sizeof(uint16_t) * m);
	sh.wcnt = (uint32_t *) carve(sizeof(uint32_t) * n_warps * FSQ_MAX_SYMS);
	sh.wcar = (uint32_t *) carve(sizeof(uint32_t) * n_warps * FSQ_MAX_SYMS);
	sh.whas = (uint32_t *) carve(sizeof(uint32_t) * n_warps);
	uint32_t *sd = (uint32_t *) sh.sd;

	uint32_t const per = (m + blockDim.x - 1) / blockDim.x;
	uint32_t const lo = min(tid * per, m), hi = min(lo + per, m);

	uint32_t const step_end = A.item_begin[blockIdx.x + 1];
	for (uint32_t step = A.item_begin[blockIdx.x]; step < step_end; ++step)
	{
		stitch_item const it = A.items[step];
		uint32_t const *in_a = it.in_a, *in_d = it.in_d;           // incoming state (true state or a summary)
		uint32_t const *pi = it.pi, *delta = it.delta;             // local summary of [s, e)
		uint32_t const s_mark = it.s_mark;                         // delta <= s_mark  <=>  same class as predecessor

		// classes of the summary
		uint32_t cnt = 0;
		for (uint32_t i = lo; i < hi; ++i) cnt += (0 == i || delta[i] > s_mark);
		uint32_t total;
		uint32_t c = block_excl_scan(cnt, scratch, total);
		for (uint32_t i = lo; i < hi; ++i)
		{
			if (0 == i || delta[i] > s_mark) ++c;
			cls[pi[i]] = (uint16_t) (c - 1);
		}
		if (it.identity_in)
			for (uint32_t i = tid; i < m; i += blockDim.x) { sh.sa[i] = (uint16_t) i; sd[i] = it.identity_d; }
		else
			for (uint32_t i = tid; i < m; i += blockDim.x) { sh.sa[i] = (uint16_t) in_a[i]; sd[i] = in_d[i]; }
		__syncthreads();

		uint32_t digits = 1;
		while (digits < 4 && (total - 1) >> (4 * digits)) ++digits;
		for (uint32_t p = 0; p < digits; ++p)
		{
			digit_source src{cls, 4 * p};
			column_step <uint32_t, TPW, false>(sh, m, A.n_tiles, 16, 0xffffffffu, src, nullptr);
		}

		uint32_t *oa = it.out_a, *od = it.out_d;
		for (uint32_t i = tid; i < m; i += blockDim.x)
		{
			uint32_t const dl = delta[i];
			oa[i] = sh.sa[i];
			od[i] = (0 == i || dl > s_mark) ? dl : sd[i];
		}
		__syncthreads();
		__threadfence();
	}
}


// ---- boundary stitch, small m: one sort instead of radix passes ---------------------------------------------------
// The same composition for m <= 1024 rows with one row per thread: the incoming positions are sorted by
// (class in the summary, incoming position) -- a bitonic network, in registers below 32 lanes apart and through shared
// memory above -- and the divergence of a row that follows a row of its own class is the maximum of the incoming
// divergences between the two incoming positions, read from a sparse table (two look-ups).  About a quarter of the
// time of the radix formulation, which stays for larger m.
__global__ void __launch_bounds__(1024) stitch_sort_kernel(stitch_args const A)
{
	extern __shared__ __align__(128) uint8_t smem_raw[];
	__shared__ uint32_t wtot[32];
	uint32_t const tid = threadIdx.x;
	uint32_t const lane = tid & 31u, warp = tid >> 5;
	uint32_t const p2 = blockDim.x;                        // power of two >= m, >= 32
	uint32_t const m = A.m;
	uint32_t levels = 1;                                   // sparse-table levels: spans of 2^l positions, 2^(levels - 1) <= m
	while ((2u << (levels - 1)) <= m) ++levels;
	uint32_t *ks0 = (uint32_t *) smem_raw;                 // [p2] exchange buffers of the sorting network
	uint32_t *ks1 = ks0 + p2;                              // [p2]
	uint32_t *st = ks1 + p2;                               // [levels][p2] range maxima of the incoming divergences
	uint16_t *cls = (uint16_t *) (st + (size_t) levels * p2);      // [p2] class of a row id
	uint16_t *sa = cls + p2;                               // [p2] incoming prefix array

	uint32_t const step_end = A.item_begin[blockIdx.x + 1];
	for (uint32_t step = A.item_begin[blockIdx.x]; step < step_end; ++step)
	{
		stitch_item const it = A.items[step];
		uint32_t const s_mark = it.s_mark;
		bool const live = tid < m;
		// every global read of the item happens before its first write (outputs may alias the summary)
		uint32_t const dl = live ? it.delta[tid] : 0u;
		uint32_t const pi = live ? it.pi[tid] : 0u;
		uint32_t const a_in = live ? (it.identity_in ? tid : it.in_a[tid]) : 0u;
		uint32_t const d_in = live ? (it.identity_in ? it.identity_d : it.in_d[tid]) : 0u;
		bool const start = live && (0 == tid || dl > s_mark);

		// class of every row: inclusive count of class starts in summary order
		uint32_t const bal = __ballot_sync(FSQ_FULL, start);
		if (0 == lane) wtot[warp] = __popc(bal);
		st[tid] = d_in;
		if (live) sa[tid] = (uint16_t) a_in;
		__syncthreads();
		{
			uint32_t const w = lane < (p2 >> 5) ? wtot[lane] : 0u;
			uint32_t const incl = warp_incl_sum(w);
			uint32_t const before = __shfl_sync(FSQ_FULL, incl - w, warp);
			uint32_t const c = before + __popc(bal & (0xffffffffu >> (31u - lane)));      // >= 1 for every live row
			if (live) cls[pi] = (uint16_t) (c - 1);
		}
		__syncthreads();

		// sparse table over the incoming divergences (positions past m hold 0)
		for (uint32_t l = 1; l < levels; ++l)
		{
			uint32_t const half = 1u << (l - 1);
			uint32_t const x = st[(size_t) (l - 1) * p2 + tid];
			uint32_t const y = tid + half < p2 ? st[(size_t) (l - 1) * p2 + tid + half] : 0u;
			st[(size_t) l * p2 + tid] = max(x, y);
			__syncthreads();
		}

		// sort by (class, incoming position); padding keys sort last
		uint32_t key = live ? (((uint32_t) cls[a_in] << 16) | tid) : 0xffffffffu;
		uint32_t flip = 0;
		for (uint32_t k = 2; k <= p2; k <<= 1)
			for (uint32_t j = k >> 1; j > 0; j >>= 1)
			{
				uint32_t other;
				if (j >= 32)
				{
					uint32_t *buf = flip ? ks1 : ks0;
					flip ^= 1u;
					buf[tid] = key;
					__syncthreads();
					other = buf[tid ^ j];
				}
				else
					other = __shfl_xor_sync(FSQ_FULL, key, j);
				bool const keep_min = (0 == (tid & j)) == (0 == (tid & k));
				key = keep_min ? min(key, other) : max(key, other);
			}
		uint32_t *buf = flip ? ks1 : ks0;
		buf[tid] = key;
		__syncthreads();

		if (live)
		{
			uint32_t const i = key & 0xffffu;
			uint32_t od = dl;
			if (!start)
			{
				uint32_t const ip = buf[tid - 1] & 0xffffu;      // the row before in the same class: ip < i
				uint32_t const len = i - ip;                      // max over incoming positions ip + 1 .. i
				uint32_t const l = 31u - __clz(len);
				od = max(st[(size_t) l * p2 + ip + 1], st[(size_t) l * p2 + i + 1 - (1u << l)]);
			}
			it.out_a[tid] = sa[i];
			it.out_d[tid] = od;
		}
		__syncthreads();
		__threadfence();
	}
}

// ---- boundary stitch, small m: counting instead of sorting ---------------------------------------------------------
// The order wanted is the stable partition of the incoming positions by class.  A position's place is (rows in classes
// before its own) + (earlier positions of its class): the second term is, inside a warp, the lanes below it that
// __match_any_sync pairs it with, and across warps a prefix over one byte per (class, warp) -- a warp holds at most 32 of
// a class -- summed four at a time with dp4a.  A third of the sorting network's instructions and five barriers
// instead of eleven; the divergences come from the same sparse table.
__global__ void __launch_bounds__(1024) stitch_count_kernel(stitch_args const A)
{
	extern __shared__ __align__(128) uint8_t smem_raw[];
	__shared__ uint32_t scratch[34];
	uint32_t const tid = threadIdx.x;
	uint32_t const lane = tid & 31u, warp = tid >> 5;
	uint32_t const p2 = blockDim.x;                        // power of two >= m, >= 32
	uint32_t const m = A.m;
	uint32_t levels = 1;                                   // sparse-table levels: spans of 2^l positions, 2^(levels - 1) <= m
	while ((2u << (levels - 1)) <= m) ++levels;
	uint32_t *tbl = (uint32_t *) smem_raw;                 // [p2][8] words = [class][warp] bytes: positions of the class in the warp
	uint32_t *cstart = tbl + (size_t) p2 * 8;              // [p2] rows in the classes before this one
	uint32_t *st = cstart + p2;                            // [levels][p2] range maxima of the incoming divergences
	uint16_t *cls = (uint16_t *) (st + (size_t) levels * p2);      // [p2] class of a row id
	uint16_t *sa = cls + p2;                               // [p2] incoming prefix array
	uint16_t *srt = sa + p2;                               // [p2] incoming position at every outgoing one

	uint32_t const step_end = A.item_begin[blockIdx.x + 1];
	for (uint32_t step = A.item_begin[blockIdx.x]; step < step_end; ++step)
	{
		stitch_item const it = A.items[step];
		uint32_t const s_mark = it.s_mark;
		bool const live = tid < m;
		// every global read of the item happens before its first write (outputs may alias the summary)
		uint32_t const dl = live ? it.delta[tid] : 0u;
		uint32_t const pi = live ? it.pi[tid] : 0u;
		uint32_t const a_in = live ? (it.identity_in ? tid : it.in_a[tid]) : 0u;
		uint32_t const d_in = live ? (it.identity_in ? it.identity_d : it.in_d[tid]) : 0u;
		bool const start = live && (0 == tid || dl > s_mark);

		// class of every row: inclusive count of class starts in summary order
		uint32_t const bal = __ballot_sync(FSQ_FULL, start);
		if (0 == lane) scratch[warp] = __popc(bal);
		st[tid] = d_in;
		if (live) sa[tid] = (uint16_t) a_in;
#pragma unroll
		for (uint32_t i = 0; i < 8; ++i) tbl[i * p2 + tid] = 0;
		__syncthreads();
		{
			uint32_t const w = lane < (p2 >> 5) ? scratch[lane] : 0u;
			uint32_t const incl = warp_incl_sum(w);
			uint32_t const before = __shfl_sync(FSQ_FULL, incl - w, warp);
			uint32_t const c = before + __popc(bal & (0xffffffffu >> (31u - lane)));      // >= 1 for every live row
			if (live) cls[pi] = (uint16_t) (c - 1);
		}
		__syncthreads();

		// sparse table over the incoming divergences (positions past m hold 0)
		for (uint32_t l = 1; l < levels; ++l)
		{
			uint32_t const half = 1u << (l - 1);
			uint32_t const x = st[(size_t) (l - 1) * p2 + tid];
			uint32_t const y = tid + half < p2 ? st[(size_t) (l - 1) * p2 + tid + half] : 0u;
			st[(size_t) l * p2 + tid] = max(x, y);
			__syncthreads();
		}

		// positions of the own class in the own warp
		uint32_t const c = live ? (uint32_t) cls[a_in] : 0xffffu;
		uint32_t const same = __match_any_sync(FSQ_FULL, c);
		uint32_t const below = __popc(same & ((1u << lane) - 1u));
		if (live && 0 == below) ((uint8_t *) tbl)[c * 32u + warp] = (uint8_t) __popc(same);
		__syncthreads();
		{
			// thread q: the size of class q, then the rows before it
			uint32_t size_q = 0;
#pragma unroll
			for (uint32_t i = 0; i < 8; ++i) size_q = __dp4a(tbl[tid * 8 + i], 0x01010101u, size_q);
			uint32_t total;
			cstart[tid] = block_excl_scan(size_q, scratch, total);
		}
		__syncthreads();
		if (live)
		{
			uint32_t at = cstart[c] + below;
			uint32_t const *row = tbl + c * 8;
#pragma unroll
			for (uint32_t i = 0; i < 8; ++i)
			{
				// the bytes of the warps before this one
				uint32_t const keep = 4 * i + 4 <= warp ? 0xffffffffu : (4 * i >= warp ? 0u : (1u << (8 * (warp - 4 * i))) - 1u);
				at = __dp4a(row[i] & keep, 0x01010101u, at);
			}
			srt[at] = (uint16_t) tid;
		}
		__syncthreads();

		if (live)
		{
			uint32_t const i = srt[tid];
			uint32_t od = dl;
			if (!start)
			{
				uint32_t const ip = srt[tid - 1];                 // the row before in the same class: ip < i
				uint32_t const len = i - ip;                      // max over incoming positions ip + 1 .. i
				uint32_t const l = 31u - __clz(len);
				od = max(st[(size_t) l * p2 + ip + 1], st[(size_t) l * p2 + i + 1 - (1u << l)]);
			}
			it.out_a[tid] = sa[i];
			it.out_d[tid] = od;
		}
		__syncthreads();
		__threadfence();
	}
}

static size_t stitch_count_smem_bytes(uint32_t m, uint32_t p2)
{
	uint32_t levels = 1;
	while ((2u << (levels - 1)) <= m) ++levels;
	return (size_t) p2 * 4 * (8 + 1 + levels) + (size_t) p2 * 2 * 3;
}

static cudaError_t launch_stitch_count(stitch_args const &A, cudaStream_t st)
{
	uint32_t p2 = 32;
	while (p2 < A.m) p2 <<= 1;
	size_t const smem = stitch_count_smem_bytes(A.m, p2);
	cudaError_t err = cudaFuncSetAttribute(stitch_count_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
	if (cudaSuccess != err) return err;
	stitch_count_kernel <<<A.n_ctas, p2, smem, st>>>(A);
	return cudaGetLastError();
}

size_t stitch_sort_smem_bytes(uint32_t m, uint32_t p2)
{
	uint32_t levels = 1;
	while ((2u << (levels - 1)) <= m) ++levels;
	return (size_t) p2 * 4 * (2 + levels) + (size_t) p2 * 2 * 2;
}

static cudaError_t launch_stitch_sort(stitch_args const &A, cudaStream_t st)
{
	uint32_t p2 = 32;
	while (p2 < A.m) p2 <<= 1;
	size_t const smem = stitch_sort_smem_bytes(A.m, p2);
	cudaError_t err = cudaFuncSetAttribute(stitch_sort_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
	if (cudaSuccess != err) return err;
	stitch_sort_kernel <<<A.n_ctas, p2, smem, st>>>(A);
	return cudaGetLastError();
}

size_t stitch_smem_bytes(uint32_t m, uint32_t n_warps)
{
	auto r16 = [](size_t b) { return (b + 15) & ~size_t(15); };
	return r16(2 * (size_t) m) * 2 + r16(4 * (size_t) m) + r16(4 * n_warps * FSQ_MAX_SYMS) * 2 + r16(4 * n_warps);
}

template <int TPW>
static cudaError_t launch_stitch_t(stitch_args const &A, uint32_t n_warps, cudaStream_t st)
{
	size_t const smem = stitch_smem_bytes(A.m, n_warps);
	auto kern = stitch_kernel <TPW>;
	cudaError_t err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
	if (cudaSuccess != err) return err;
	kern <<<A.n_ctas, n_warps * 32, smem, st>>>(A);
	return cudaGetLastError();
}

cudaError_t launch_stitch(stitch_args const &A, cudaStream_t st)
{
	if (2 == A.use_sort && A.m <= 1024) return launch_stitch_count(A, st);
	if (A.use_sort && A.m <= 1024) return launch_stitch_sort(A, st);
	uint32_t const n_warps = A.n_tiles < 32 ? A.n_tiles : 32;
	uint32_t const tpw = (A.n_tiles + n_warps - 1) / n_warps;
#define FSQ_CASE(T) if (tpw <= T) return launch_stitch_t <T>(A, n_warps, st);
	FSQ_CASE(1) FSQ_CASE(2) FSQ_CASE(4) FSQ_CASE(8) FSQ_CASE(16) FSQ_CASE(32) FSQ_CASE(64)
#undef FSQ_CASE
	return cudaErrorInvalidValue;
}

} // namespace fsq
