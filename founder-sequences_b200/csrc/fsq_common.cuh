// Shared device helpers for libfsq_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define FSQ_FULL 0xffffffffu
#ifndef FSQ_MAX_SYMS
#define FSQ_MAX_SYMS 16            // 4-bit packed symbols
#endif
#define FSQ_SENTINEL16 0xffffu

namespace fsq {

// ---- counter-based hash shared with synth.py -------------------------------------------------------------
__host__ __device__ __forceinline__ uint64_t mix64(uint64_t x)
{
	x ^= x >> 33; x *= 0xFF51AFD7ED558CCDull;
	x ^= x >> 33; x *= 0xC4CEB9FE1A85EC53ull;
	x ^= x >> 33;
	return x;
}

__host__ __device__ __forceinline__ uint32_t hash32(uint64_t seed, uint64_t stream, uint64_t i, uint64_t j)
{
	uint64_t const s = seed ^ (stream * 0x9E3779B97F4A7C15ull);
	uint64_t x = mix64(s + i * 0xD6E8FEB86659FD93ull);
	x = mix64(x ^ j);
	return (uint32_t) (x >> 32);
}

struct synth_params
{
	uint64_t seed;
	uint32_t founders;
	uint32_t recomb_block;   // 1000
	uint32_t mu32;           // mutation threshold, mu * 2^32
	uint32_t gap32;          // gap threshold (0 = no gaps)
};

// Symbol index 0..4 into "ACGT-" of cell (row, col).
__host__ __device__ __forceinline__ uint32_t synth_symbol(synth_params const &p, uint32_t row, uint64_t col)
{
	uint32_t const phi = hash32(p.seed, 2, row, 0) % p.recomb_block;
	uint32_t const f = hash32(p.seed, 1, row, (col + phi) / p.recomb_block) % p.founders;
	uint32_t sym = hash32(p.seed, 0, f, col) & 3u;
	if (hash32(p.seed, 3, row, col) < p.mu32)
		sym = hash32(p.seed, 5, row, col) & 3u;
	if (p.gap32 && hash32(p.seed, 4, row, col / 20) < p.gap32)
		sym = 4;
	return sym;
}

// ---- mbarrier + TMA bulk copy (1-D cp.async.bulk) ----------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(void const *p) { return (uint32_t) __cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count)
{
	asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(count) : "memory");
}

__device__ __forceinline__ void mbar_fence_init()
{
	asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}

__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes)
{
	asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(bar)), "r"(bytes) : "memory");
}

__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity)
{
	asm volatile(
		"{\n"
		".reg .pred p;\n"
		"WAIT_%=:\n"
		"mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
		"@p bra DONE_%=;\n"
		"bra WAIT_%=;\n"
		"DONE_%=:\n"
		"}\n" :: "r"(smem_u32(bar)), "r"(parity) : "memory");
}

// global → shared bulk copy through the TMA unit; bytes and both addresses must be multiples of 16.
__device__ __forceinline__ void tma_load_1d(void *smem_dst, void const *gmem_src, uint32_t bytes, uint64_t *bar)
{
	asm volatile(
		"cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
		:: "r"(smem_u32(smem_dst)), "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// ---- warp helpers ----------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t lane_id() { return threadIdx.x & 31u; }

__device__ __forceinline__ uint32_t warp_incl_sum(uint32_t v)
{
	uint32_t const lane = lane_id();
#pragma unroll
	for (int o = 1; o < 32; o <<= 1)
	{
		uint32_t const t = __shfl_up_sync(FSQ_FULL, v, o);
		if (lane >= (uint32_t) o) v += t;
	}
	return v;
}

} // namespace fsq
