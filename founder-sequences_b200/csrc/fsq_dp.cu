// Min-max segmentation DP with the reference's exact tie-breaking and RMQ behaviour, the greedy reduction, and
// the segment-set / edge-weight kernels (sm_100a).
#include "fsq_internal.h"
#include "fsq_common.cuh"

namespace fsq {

// ---- founder_sequences::rmq <…, 64> emulated bit for bit (reference: include/founder_sequences/rmq.hh:61-118) ----
// lev[0][b]  = leftmost argmin of entries [64 b, 64 b + 64)
// lev[p][i]  = (M[lev[0][i + 2^p - 1]] < M[lev[p-1][i]]) ? lev[0][i + 2^p - 1] : lev[p-1][i]
//              -- rmq.hh:76-77 reads the same cell twice, so level p covers blocks {i, i+1, i+3, ..., i+2^p-1} only.

__device__ __forceinline__ uint32_t naive_min(uint32_t const *M, uint32_t first, uint32_t last)
{
	uint32_t best = first, bv = M[first];
	for (uint32_t i = first + 1; i < last; ++i)
	{
		uint32_t const v = M[i];
		if (v < bv) { bv = v; best = i; }
	}
	return best;
}

__device__ __forceinline__ uint32_t rmq_query(uint32_t const *M, uint32_t const *lev, uint32_t n_blk, uint32_t beg, uint32_t end)
{
	uint32_t const bb = beg / 64 + 1;
	uint32_t const eb = end / 64;
	if (bb >= eb)
		return naive_min(M, beg, end);
	uint32_t const p = 31u - __clz(eb - bb);
	uint32_t const s1 = lev[(size_t) p * n_blk + bb];
	uint32_t const s2 = lev[(size_t) p * n_blk + eb - (1u << p)];
	uint32_t smp = (M[s2] < M[s1]) ? s2 : s1;
	uint32_t const left = naive_min(M, beg, bb * 64);
	smp = (M[left] < M[smp]) ? left : smp;
	if (end == eb * 64)
		return smp;
	uint32_t const right = naive_min(M, eb * 64, end);
	return (M[right] < M[smp]) ? right : smp;
}

// One column of calculate_segmentation_lp_dp_arg (reference: founder-sequences/segmentation_lp_context.cc:393-481),
// evaluated by one warp over the column's top list.  Only ranges whose segment size S is <= cap are listed; any
// other range has a candidate value > cap, so the result is exact whenever it is <= cap (or the list is complete).
__device__ __forceinline__ void dp_column(dp_args const &A, uint64_t const k, uint64_t const t_out)
{
	uint32_t const lane = lane_id();
	uint32_t const L = (uint32_t) A.L;
	uint32_t const cw = A.cnt[k];
	uint32_t const E = cw & 0x7fffffffu;
	bool const complete = cw >> 31;
	uint2 const *list = A.toplist + k * (A.cap + 1);
	uint32_t const hi_clip = (uint32_t) (k + 2 - L);

	unsigned long long best = ~0ull;
	uint32_t best_idx = 0, best_S = 0;
	for (uint32_t t0 = 1; t0 < E; t0 += 32)
	{
		uint32_t const t = t0 + lane;
		unsigned long long key = ~0ull;
		uint32_t idx = 0, S = 0;
		if (t < E)
		{
			uint2 const en = list[t];
			uint32_t const up = list[t - 1].x;
			S = en.y;
			if (0 != en.x)         // the range starting at value 0 is the whole-range candidate, below
			{
				uint32_t const lo = max(en.x, L);
				uint32_t const hi = min(up, hi_clip);
				if (lo < hi)
				{
					idx = rmq_query(A.tb_max, A.lev, A.n_blk, lo - L, hi - L);
					uint32_t const val = max(A.tb_max[idx], S);
					key = ((unsigned long long) val << 32) | (0xffffffffu - t);   // ties: the lowest range (largest t) wins
				}
			}
		}
		unsigned long long kmin = key;
#pragma unroll
		for (int o = 16; o > 0; o >>= 1)
		{
			unsigned long long const other = __shfl_xor_sync(FSQ_FULL, kmin, o);
			kmin = other < kmin ? other : kmin;
		}
		if (kmin < best)
		{
			uint32_t const owner = __ffs(__ballot_sync(FSQ_FULL, key == kmin)) - 1;
			best = kmin;
			best_idx = __shfl_sync(FSQ_FULL, idx, owner);
			best_S = __shfl_sync(FSQ_FULL, S, owner);
		}
	}
	if (0 == lane)
	{
		uint32_t const h0 = A.h0[k];
		uint32_t const whole = A.m - h0;          // == m when value 0 is absent
		uint32_t const best_val = (uint32_t) (best >> 32);
		uint32_t lb = 0, mx = whole, sz = whole;
		if (~0ull != best && best_val < whole) { lb = best_idx + L; mx = best_val; sz = best_S; }
		if (!complete && mx > A.cap)
			atomicMax(A.overflow, (uint32_t) (k + 1));
		A.tb_max[t_out] = mx;
		A.tb_lb[t_out] = lb;
		A.tb_size[t_out] = sz;
	}
}

// First pass staging (reference: segmentation_lp_context.cc:26-188).  Column k reads only table entries
// <= k + 1 - 2L, so L consecutive columns are independent: one wavefront = up to 32 columns, one warp each.
__global__ void __launch_bounds__(1024) dp_kernel(dp_args const A)
{
	uint32_t const warp = threadIdx.x >> 5;
	uint32_t const lane = lane_id();
	uint64_t const n = A.n, L = A.L;
	uint64_t const k_first = L - 1, k_last = n - L - 1;
	uint32_t const WF = (uint32_t) (L < 32 ? L : 32);

	for (uint64_t k0 = k_first; k0 <= k_last; k0 += WF)
	{
		uint64_t const k = k0 + warp;
		if (warp < WF && k <= k_last)
			dp_column(A, k, k + 1 - L);
		__syncthreads();

		// rmq.update for every entry written in this wavefront (rmq.hh:61-84)
		uint64_t const t_lo = k0 + 1 - L;
		uint64_t const k_hi = (k0 + WF - 1 < k_last) ? k0 + WF - 1 : k_last;
		uint64_t const t_hi = k_hi + 1 - L;
		{
			// a block b completes when entry 64 b + 63 is written; at most one per wavefront (WF <= 32)
			uint64_t const b64 = (t_hi + 1) / 64;           // number of complete blocks now
			if (b64 > 0 && 64 * b64 - 1 >= t_lo)
			{
				uint32_t const b = (uint32_t) (b64 - 1);
				if (0 == warp)
				{
					uint32_t const i0 = 64 * b + lane, i1 = i0 + 32;
					uint32_t const v0 = A.tb_max[i0], v1 = A.tb_max[i1];
					unsigned long long key = (v1 < v0) ? (((unsigned long long) v1 << 32) | i1) : (((unsigned long long) v0 << 32) | i0);
#pragma unroll
					for (int o = 16; o > 0; o >>= 1)
					{
						unsigned long long const other = __shfl_xor_sync(FSQ_FULL, key, o);
						key = other < key ? other : key;
					}
					uint32_t const smp0 = (uint32_t) key;            // leftmost minimal entry of the block
					uint32_t const val0 = (uint32_t) (key >> 32);
					if (0 == lane) A.lev[b] = smp0;
					uint32_t const p = lane;
					if (p >= 1 && p < A.n_levels && (1ull << p) <= (uint64_t) b + 1)
					{
						uint32_t const i = b + 1 - (1u << p);
						uint32_t const prev = A.lev[(size_t) (p - 1) * A.n_blk + i];
						A.lev[(size_t) p * A.n_blk + i] = (val0 < A.tb_max[prev]) ? smp0 : prev;
					}
				}
			}
		}
		__syncthreads();
	}
	// segmentation_lp_context.cc:165-185: the final column, from the final histogram
	if (0 == warp)
		dp_column(A, n - 1, n - L);
}

cudaError_t launch_dp(dp_args const &A, cudaStream_t st)
{
	dp_kernel <<<1, 1024, 0, st>>>(A);
	return cudaGetLastError();
}


// ---- greedy reduction (reference: segmentation_lp_context.cc:335-390; pbwt.hh:206-216) ------------------------------
// #{d_rb > lb} is the number of distinct rows over [lb, rb), which cannot decrease as rb grows, so the run that the
// reference's sequential loop would merge from a given start is a function of the start alone: one warp per start u
// (u = 0: left bound 0; u >= 1: left bound rb[u-1]) extends while the count stays <= K.  The host then follows
// next[] from u = 0.
__global__ void reduce_kernel(reduce_args const A)
{
	uint32_t const lane = lane_id();
	uint32_t const u = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	if (u >= A.T) return;
	uint32_t const cur_lb = (0 == u) ? 0u : (uint32_t) A.tb_rb[u - 1];
	uint32_t prev = u, size = A.tb_size[u];
	for (uint32_t t = u + 1; t < A.T; ++t)
	{
		uint32_t const *d = A.st_d + (size_t) t * A.m;
		uint32_t s = 0;
		for (uint32_t i = lane; i < A.m; i += 32)
			s += (cur_lb < d[i]);
		s = __reduce_add_sync(FSQ_FULL, s);
		if (s <= A.K) { size = s; prev = t; }
		else break;
	}
	if (0 == lane) { A.next[u] = prev; A.next_size[u] = size; }
}

cudaError_t launch_reduce(reduce_args const &A, cudaStream_t st)
{
	uint32_t const wpb = 8;
	reduce_kernel <<<(A.T + wpb - 1) / wpb, wpb * 32, 0, st>>>(A);
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

} // namespace fsq
