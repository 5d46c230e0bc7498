// fsq_pair_counts_noalloc: the class-pair counting of fsq_pair_counts (reference: greedy_matcher.cc:293-343; fsq_dp.cu,
// pair_weights_kernel / pair_counts_kernel) for callers that must not allocate: code under CUDA graph capture, tight
// serving loops, harnesses that forbid cudaMalloc.  The contingency table the counting goes through -- scratch that
// fsq_pair_counts takes from cudaMalloc -- lives in a file-scope device array here, of a fixed maximum size, which every
// GPU that loads the library holds until the process ends.  Inputs and outputs are the caller's device memory.
// No counterpart in the reference.
//
// This entry point calls no cudaMalloc / cudaFree and nothing that allocates: a memset and two kernels on the caller's
// stream.  All calls on a device share the one table, which is cleared at the start of every call: calls that overlap
// on the GPU (two streams, two threads) race on it and must be serialised by the caller.
#include "../../include/fsq.h"
#include "fsq_internal.h"
#include "fsq_common.cuh"
#include <cstdarg>
#include <cstdio>
#include <mutex>

namespace {

constexpr uint32_t HEADER_WORDS = 64;            // 256 bytes in front of the table; word 0: a class index was out of range
constexpr size_t SCRATCH_WORDS = (size_t) HEADER_WORDS + FSQ_PAIR_COUNTS_NOALLOC_MAX_CELLS;

// In host code this name is a host-side stand-in, not a device address: kernels receive what cudaGetSymbolAddress returns.
static __device__ __align__(256) int32_t g_pair_scratch[SCRATCH_WORDS];

constexpr int MAX_DEVICES = 64;
std::mutex g_setup_mutex;
int32_t *g_scratch_address[MAX_DEVICES];          // per device, looked up once with that device current

int report(char *error, size_t error_size, int code, char const *fmt, ...)
{
	if (error && error_size)
	{
		va_list ap;
		va_start(ap, fmt);
		vsnprintf(error, error_size, fmt, ap);
		va_end(ap);
	}
	return code;
}

// The contingency table of the two class maps (the variant of pair_weights_kernel that checks its indices: the table
// has a fixed size here, and a row outside it must not be written).
__global__ void pair_weights_checked_kernel(uint32_t const *cls_l, uint32_t const *cls_r, uint32_t m, uint32_t runs_l, uint32_t runs_r, int32_t *scratch)
{
	uint32_t const x = blockIdx.x * blockDim.x + threadIdx.x;
	if (x >= m) return;
	uint32_t const l = cls_l[x], r = cls_r[x];
	if (l >= runs_l || r >= runs_r) { scratch[0] = 1; return; }
	atomicAdd(&scratch[HEADER_WORDS + (size_t) l * runs_r + r], 1);
}

// The non-zero cells in row-major order, i.e. the distinct pairs in ascending (lhs, rhs) order, with their counts: one
// CTA streams over the table and compacts (as pair_counts_kernel does).  *n_pairs = 0xffffffff when an index was bad.
__global__ void __launch_bounds__(1024) pair_compact_kernel(int32_t const *scratch, uint32_t runs_l, uint32_t runs_r, uint32_t *lhs, uint32_t *rhs, uint32_t *count, uint32_t *n_pairs)
{
	__shared__ uint32_t wsum[32];
	__shared__ uint32_t carry;
	if (scratch[0])
	{
		if (0 == threadIdx.x) *n_pairs = 0xffffffffu;
		return;
	}
	int32_t const *W = scratch + HEADER_WORDS;
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

} // namespace

extern "C" int fsq_pair_counts_noalloc(const uint32_t *class_lhs, const uint32_t *class_rhs, uint32_t m, uint32_t runs_lhs, uint32_t runs_rhs,
	uint32_t *lhs, uint32_t *rhs, uint32_t *count, uint32_t *n_pairs, void *stream, char *error, size_t error_size)
{
	if (error && error_size) error[0] = 0;
	if (!class_lhs || !class_rhs || !lhs || !rhs || !count || !n_pairs)
		return report(error, error_size, FSQ_ERR_ARG, "fsq_pair_counts_noalloc: null argument");
	if (0 == m || 0 == runs_lhs || 0 == runs_rhs)
		return report(error, error_size, FSQ_ERR_ARG, "fsq_pair_counts_noalloc: m, runs_lhs and runs_rhs must be positive");
	uint64_t const cells = (uint64_t) runs_lhs * runs_rhs;
	if (cells > FSQ_PAIR_COUNTS_NOALLOC_MAX_CELLS)
		return report(error, error_size, FSQ_ERR_UNSUPPORTED,
			"fsq_pair_counts_noalloc: runs_lhs * runs_rhs = %llu cells exceed the static table of %llu (FSQ_PAIR_COUNTS_NOALLOC_MAX_CELLS); use fsq_pair_counts",
			(unsigned long long) cells, (unsigned long long) FSQ_PAIR_COUNTS_NOALLOC_MAX_CELLS);

	int device = -1;
	cudaError_t e = cudaGetDevice(&device);
	if (cudaSuccess != e) return report(error, error_size, FSQ_ERR_CUDA, "fsq_pair_counts_noalloc: cudaGetDevice: %s", cudaGetErrorString(e));
	if (device < 0 || device >= MAX_DEVICES) return report(error, error_size, FSQ_ERR_UNSUPPORTED, "fsq_pair_counts_noalloc: device %d", device);
	int32_t *scratch = nullptr;
	{
		// set-up, once per device, with that device current: the array's address on this device
		std::lock_guard <std::mutex> lock(g_setup_mutex);
		if (!g_scratch_address[device])
		{
			void *p = nullptr;
			e = cudaGetSymbolAddress(&p, g_pair_scratch);
			if (cudaSuccess != e || !p)
				return report(error, error_size, FSQ_ERR_CUDA, "fsq_pair_counts_noalloc: cudaGetSymbolAddress on device %d: %s", device, cudaGetErrorString(e));
			g_scratch_address[device] = (int32_t *) p;
		}
		scratch = g_scratch_address[device];
	}

	cudaStream_t const st = (cudaStream_t) stream;
	// the table keeps the last call's contents: the header and the cells this call uses start from zero
	e = cudaMemsetAsync(scratch, 0, ((size_t) HEADER_WORDS + cells) * 4, st);
	if (cudaSuccess != e) return report(error, error_size, FSQ_ERR_CUDA, "fsq_pair_counts_noalloc: cudaMemsetAsync: %s", cudaGetErrorString(e));
	pair_weights_checked_kernel <<<(m + 255) / 256, 256, 0, st>>>(class_lhs, class_rhs, m, runs_lhs, runs_rhs, scratch);
	pair_compact_kernel <<<1, 1024, 0, st>>>(scratch, runs_lhs, runs_rhs, lhs, rhs, count, n_pairs);
	e = cudaGetLastError();
	if (cudaSuccess != e) return report(error, error_size, FSQ_ERR_CUDA, "fsq_pair_counts_noalloc: kernel launch: %s", cudaGetErrorString(e));
	return FSQ_OK;
}
