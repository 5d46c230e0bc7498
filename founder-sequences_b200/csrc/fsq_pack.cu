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


// ---- identity columns (reference: remove-identity-columns/main.cc:107-160) --------------------------------------------
// A column all of whose rows carry the same character.  On the packed, column-major alignment a column is col_stride
// contiguous bytes: one warp compares every nibble of it with the first row's.
__global__ void identity_columns_kernel(uint8_t const *cols, uint64_t col_stride, uint32_t m, uint64_t n, uint8_t *identity)
{
	uint64_t const c = (uint64_t) blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	if (c >= n) return;
	uint32_t const lane = threadIdx.x & 31u;
	uint32_t const *col = (uint32_t const *) (cols + c * col_stride);
	uint32_t const first = col[0] & 15u;
	uint32_t const want = first * 0x11111111u;
	uint32_t const full = m >> 3, rest = m & 7u;
	bool same = true;
	for (uint32_t w = lane; w < full; w += 32) same &= col[w] == want;
	if (rest && 0 == lane)
	{
		uint32_t const mask = (1u << (4 * rest)) - 1u;
		same &= ((col[full] ^ want) & mask) == 0;
	}
	same = __all_sync(FSQ_FULL, same);
	if (0 == lane) identity[c] = same ? 1 : 0;
}

// exclusive prefix sums of (1 - identity) in three steps: per-block totals, their scan by one block, the final positions
constexpr int SCAN_BLOCK = 1024;

__global__ void __launch_bounds__(SCAN_BLOCK) kept_block_totals_kernel(uint8_t const *identity, uint64_t n, uint32_t *block_tot)
{
	__shared__ uint32_t wsum[32];
	uint64_t const i = (uint64_t) blockIdx.x * SCAN_BLOCK + threadIdx.x;
	uint32_t const v = (i < n && !identity[i]) ? 1u : 0u;
	uint32_t const b = __ballot_sync(FSQ_FULL, v);
	if (0 == (threadIdx.x & 31u)) wsum[threadIdx.x >> 5] = __popc(b);
	__syncthreads();
	if (threadIdx.x < 32)
	{
		uint32_t const s = __reduce_add_sync(FSQ_FULL, wsum[threadIdx.x]);
		if (0 == threadIdx.x) block_tot[blockIdx.x] = s;
	}
}

__global__ void __launch_bounds__(1024) scan_block_totals_kernel(uint32_t *block_tot, uint32_t n_blocks, uint64_t *total)
{
	__shared__ uint32_t wsum[32];
	__shared__ uint32_t carry;
	if (0 == threadIdx.x) carry = 0;
	__syncthreads();
	for (uint32_t b0 = 0; b0 < n_blocks; b0 += 1024)
	{
		uint32_t const b = b0 + threadIdx.x;
		uint32_t const v = b < n_blocks ? block_tot[b] : 0u;
		uint32_t const inc = warp_incl_sum(v);
		if (31 == (threadIdx.x & 31u)) wsum[threadIdx.x >> 5] = inc;
		__syncthreads();
		uint32_t before = carry;
		for (uint32_t w = 0; w < (threadIdx.x >> 5); ++w) before += wsum[w];
		if (b < n_blocks) block_tot[b] = before + inc - v;
		__syncthreads();
		if (1023 == threadIdx.x) carry = before + inc;
		__syncthreads();
	}
	if (0 == threadIdx.x) *total = carry;
}

// the kept columns move to their new places: one CTA copies the columns of one scan block
__global__ void __launch_bounds__(SCAN_BLOCK) compact_columns_kernel(uint8_t const *cols, uint64_t col_stride, uint64_t n, uint8_t const *identity, uint32_t const *block_off, uint8_t *out)
{
	__shared__ uint32_t wsum[32];
	__shared__ uint32_t dst_of[SCAN_BLOCK];
	uint64_t const i = (uint64_t) blockIdx.x * SCAN_BLOCK + threadIdx.x;
	uint32_t const v = (i < n && !identity[i]) ? 1u : 0u;
	uint32_t const lane = threadIdx.x & 31u;
	uint32_t const b = __ballot_sync(FSQ_FULL, v);
	if (0 == lane) wsum[threadIdx.x >> 5] = __popc(b);
	__syncthreads();
	uint32_t before = block_off[blockIdx.x];
	for (uint32_t w = 0; w < (threadIdx.x >> 5); ++w) before += wsum[w];
	dst_of[threadIdx.x] = v ? before + __popc(b & ((1u << lane) - 1u)) : 0xffffffffu;
	__syncthreads();
	uint32_t const words = (uint32_t) (col_stride / 16);
	for (uint32_t c = threadIdx.x >> 5; c < SCAN_BLOCK; c += SCAN_BLOCK / 32)
	{
		uint32_t const d = dst_of[c];
		if (0xffffffffu == d) continue;
		uint4 const *src = (uint4 const *) (cols + ((uint64_t) blockIdx.x * SCAN_BLOCK + c) * col_stride);
		uint4 *dst = (uint4 *) (out + (uint64_t) d * col_stride);
		for (uint32_t w = lane; w < words; w += 32) dst[w] = src[w];
	}
}

cudaError_t launch_identity_columns(uint8_t const *cols, uint64_t col_stride, uint32_t m, uint64_t n, uint8_t *identity, cudaStream_t st)
{
	identity_columns_kernel <<<(unsigned) ((n + 7) / 8), 256, 0, st>>>(cols, col_stride, m, n, identity);
	return cudaGetLastError();
}

cudaError_t launch_kept_offsets(uint8_t const *identity, uint64_t n, uint32_t *block_off, uint64_t *total, cudaStream_t st)
{
	uint32_t const n_blocks = (uint32_t) ((n + SCAN_BLOCK - 1) / SCAN_BLOCK);
	kept_block_totals_kernel <<<n_blocks, SCAN_BLOCK, 0, st>>>(identity, n, block_off);
	scan_block_totals_kernel <<<1, 1024, 0, st>>>(block_off, n_blocks, total);
	return cudaGetLastError();
}

cudaError_t launch_compact_columns(uint8_t const *cols, uint64_t col_stride, uint64_t n, uint8_t const *identity, uint32_t const *block_off, uint8_t *out, cudaStream_t st)
{
	uint32_t const n_blocks = (uint32_t) ((n + SCAN_BLOCK - 1) / SCAN_BLOCK);
	compact_columns_kernel <<<n_blocks, SCAN_BLOCK, 0, st>>>(cols, col_stride, n, identity, block_off, out);
	return cudaGetLastError();
}

} // namespace fsq
