// Row-major bytes -> column-major 4-bit alphabet-packed alignment in HBM, and the device-side synthetic generator
// (sm_100a).  Replaces the reference's per-cell `inputs[a[i]][k]` + char_to_comp
// (lib/libbio/include/libbio/pbwt.hh:93-95, 119-121) by a layout in which a column is m/2 contiguous bytes.
#include "fsq_internal.h"
#include "fsq_common.cuh"

namespace fsq {

// Tile: 256 rows x 32 columns.  Reads: one 32-byte sector per row; writes: 128 contiguous bytes per column.
constexpr int PACK_ROWS = 256;
constexpr int PACK_COLS = 32;

__global__ void __launch_bounds__(PACK_ROWS) pack_kernel(
	uint8_t const *rows, uint64_t row_stride, uint32_t m, uint64_t n, uint8_t const *comp,
	uint8_t *cols, uint64_t col_stride, uint64_t col_first, uint32_t *sym_mask
)
{
	__shared__ uint8_t tile[PACK_COLS][PACK_ROWS + 4];
	__shared__ uint8_t scomp[256];
	uint32_t const tid = threadIdx.x;
	scomp[tid] = comp[tid];
	__syncthreads();

	uint64_t const c0 = (uint64_t) blockIdx.x * PACK_COLS;
	uint32_t const r0 = blockIdx.y * PACK_ROWS;
	uint32_t const r = r0 + tid;
	uint32_t const nc = (uint32_t) (n - c0 < PACK_COLS ? n - c0 : PACK_COLS);
	uint32_t present = 0;      // comp values seen by this thread
	if (r < m)
	{
		uint8_t const *src = rows + (uint64_t) r * row_stride + c0;
		if (PACK_COLS == nc && 0 == (((uintptr_t) src) & 15))
		{
			uint4 const v0 = ((uint4 const *) src)[0], v1 = ((uint4 const *) src)[1];
			uint32_t const w[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
#pragma unroll
			for (int q = 0; q < 8; ++q)
#pragma unroll
				for (int bb = 0; bb < 4; ++bb)
				{
					uint8_t const cv = scomp[(w[q] >> (8 * bb)) & 0xffu];
					tile[q * 4 + bb][tid] = cv;
					present |= 1u << (cv & 31u);
				}
		}
		else
		{
			for (uint32_t c = 0; c < nc; ++c)
			{
				uint8_t const cv = scomp[src[c]];
				tile[c][tid] = cv;
				present |= 1u << (cv & 31u);
			}
		}
	}
	else
	{
		for (uint32_t c = 0; c < PACK_COLS; ++c)
			tile[c][tid] = 0;
	}
	present = __reduce_or_sync(FSQ_FULL, present);
	if (0 == (tid & 31u) && present) atomicOr(sym_mask, present);
	__syncthreads();

	// 32 columns x 32 words; thread -> (column = tid / 8 + 0, 8, 16, 24; word = tid % 8 * 4 .. + 3)?  Keep it simple:
	// word index w = tid % 32, columns tid / 32 + 8 * pass.
	uint32_t const w = tid & 31u;
	uint32_t const words_per_col = (uint32_t) (col_stride / 4);
	uint32_t const gw = r0 / 8 + w;
	for (uint32_t c = tid >> 5; c < nc; c += PACK_ROWS / 32)
	{
		if (gw < words_per_col)
		{
			uint32_t word = 0;
#pragma unroll
			for (int q = 0; q < 8; ++q)
				word |= (uint32_t) (tile[c][w * 8 + q] & 15u) << (4 * q);
			((uint32_t *) (cols + (col_first + c0 + c) * col_stride))[gw] = word;
		}
	}
}

cudaError_t launch_pack(uint8_t const *d_rows, uint64_t row_stride, uint32_t m, uint64_t n, uint8_t const *d_comp, uint8_t *cols, uint64_t col_stride, uint64_t col_first, uint32_t *sym_mask, cudaStream_t st)
{
	// grid.y covers every word of a column, including the padding words beyond m (written as zero).
	uint32_t const words_per_col = (uint32_t) (col_stride / 4);
	dim3 const grid((unsigned) ((n + PACK_COLS - 1) / PACK_COLS), (words_per_col * 8 + PACK_ROWS - 1) / PACK_ROWS);
	pack_kernel <<<grid, PACK_ROWS, 0, st>>>(d_rows, row_stride, m, n, d_comp, cols, col_stride, col_first, sym_mask);
	return cudaGetLastError();
}


// Synthetic cells straight into the packed layout: thread = one 32-bit word (8 rows) of one column.
// Comps follow the reference alphabet for the characters "-ACGT" in byte order: without gaps A,C,G,T = 1..4;
// with gaps '-' = 1 and A,C,G,T = 2..5.
__global__ void synth_packed_kernel(uint8_t *cols, uint64_t col_stride, uint32_t m, uint64_t n, uint64_t col0, synth_params p)
{
	uint32_t const words_per_col = (uint32_t) (col_stride / 4);
	uint64_t const idx = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
	uint64_t const c = idx / words_per_col;
	uint32_t const w = (uint32_t) (idx % words_per_col);
	if (c >= n) return;
	uint32_t word = 0;
#pragma unroll
	for (int q = 0; q < 8; ++q)
	{
		uint32_t const r = w * 8 + q;
		if (r < m)
		{
			uint32_t const sym = synth_symbol(p, r, col0 + c);
			uint32_t const comp = p.gap32 ? (4 == sym ? 1u : sym + 2u) : sym + 1u;
			word |= comp << (4 * q);
		}
	}
	((uint32_t *) (cols + c * col_stride))[w] = word;
}

static synth_params make_params(uint64_t seed, uint32_t founders, int gaps)
{
	synth_params p;
	p.seed = seed;
	p.founders = founders;
	p.recomb_block = 1000;
	p.mu32 = (uint32_t) (1e-3 * 4294967296.0);
	p.gap32 = gaps ? (uint32_t) (2e-3 * 4294967296.0) : 0u;
	return p;
}

cudaError_t launch_synth_packed(uint8_t *cols, uint64_t col_stride, uint32_t m, uint64_t n, uint64_t col0, uint64_t seed, uint32_t founders, int gaps, cudaStream_t st)
{
	uint64_t const total = n * (col_stride / 4);
	synth_packed_kernel <<<(unsigned) ((total + 255) / 256), 256, 0, st>>>(cols, col_stride, m, n, col0, make_params(seed, founders, gaps));
	return cudaGetLastError();
}

__global__ void synth_rows_kernel(uint8_t *rows, uint64_t row_stride, uint32_t m, uint64_t n, uint64_t col0, synth_params p)
{
	uint64_t const c = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
	uint32_t const r = blockIdx.y;
	if (c >= n) return;
	char const alphabet[5] = {'A', 'C', 'G', 'T', '-'};
	rows[(uint64_t) r * row_stride + c] = (uint8_t) alphabet[synth_symbol(p, r, col0 + c)];
}

cudaError_t launch_synth_rows(uint8_t *d_rows, uint64_t row_stride, uint32_t m, uint64_t n, uint64_t col0, uint64_t seed, uint32_t founders, int gaps, cudaStream_t st)
{
	dim3 const grid((unsigned) ((n + 255) / 256), m);
	synth_rows_kernel <<<grid, 256, 0, st>>>(d_rows, row_stride, m, n, col0, make_params(seed, founders, gaps));
	return cudaGetLastError();
}

} // namespace fsq
