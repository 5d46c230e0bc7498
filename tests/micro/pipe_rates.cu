// Micro-benchmark (development aid, not part of the product): issue rate of the instructions the column step is made
// of, in warp instructions per cycle per SM sub-partition, on the device it runs on.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipe_rates pipe_rates.cu && ./pipe_rates
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define OPS(X) \
	X(ffma,    "fma.rn.f32 %0, %0, %1, %2;") \
	X(fmnmx,   "max.f32 %0, %0, %1;") \
	X(fset,    "set.eq.f32.f32 %0, %0, %1;") \
	X(hfma2,   "fma.rn.f16x2 %0, %0, %1, %2;") \
	X(hmnmx2,  "max.f16x2 %0, %0, %1;") \
	X(hset2,   "set.eq.f16x2.f16x2 %0, %0, %1;") \
	X(hadd2,   "add.f16x2 %0, %0, %1;") \
	X(prmt,    "prmt.b32 %0, %0, %1, %2;") \
	X(lop3,    "lop3.b32 %0, %0, %1, %2, 0x96;") \
	X(imad,    "mad.lo.u32 %0, %0, %1, %2;") \
	X(iadd,    "add.u32 %0, %0, %1;") \
	X(umax,    "max.u32 %0, %0, %1;") \
	X(shf,     "shf.r.wrap.b32 %0, %0, %1, %2;")

template <int OP> __device__ __forceinline__ void op(uint32_t &x, uint32_t a, uint32_t b);
enum { 
#define X(n, s) OP_##n,
OPS(X)
#undef X
OP_COUNT };
#define X(n, s) template <> __device__ __forceinline__ void op <OP_##n>(uint32_t &x, uint32_t a, uint32_t b) { asm volatile(s : "+r"(x) : "r"(a), "r"(b)); }
OPS(X)
#undef X

template <int OP>
__global__ void __launch_bounds__(256) rate_kernel(uint32_t *out, uint32_t a, uint32_t b, int iters)
{
	uint32_t x[8];
#pragma unroll
	for (int i = 0; i < 8; ++i) x[i] = threadIdx.x + i;
	for (int it = 0; it < iters; ++it)
	{
#pragma unroll
		for (int r = 0; r < 8; ++r)
#pragma unroll
			for (int i = 0; i < 8; ++i) op <OP>(x[i], a, b);
	}
	uint32_t s = 0;
#pragma unroll
	for (int i = 0; i < 8; ++i) s ^= x[i];
	out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int OP> static void run(char const *name, uint32_t *d_out, int sms, double ghz)
{
	int const iters = 2000, ctas = sms * 8;
	cudaEvent_t e0, e1;
	cudaEventCreate(&e0); cudaEventCreate(&e1);
	rate_kernel <OP> <<<ctas, 256>>>(d_out, 0x3c003c00u, 0x00010001u, 10);
	cudaEventRecord(e0);
	rate_kernel <OP> <<<ctas, 256>>>(d_out, 0x3c003c00u, 0x00010001u, iters);
	cudaEventRecord(e1);
	cudaEventSynchronize(e1);
	float ms = 0;
	cudaEventElapsedTime(&ms, e0, e1);
	double const warp_instr = (double) ctas * 8 /* warps */ * iters * 64;
	double const per_smsp_cycle = warp_instr / (sms * 4) / (ms * 1e-3 * ghz * 1e9);
	printf("%-8s %8.3f ms  %.3f warp instr / cycle / SMSP (at %.3f GHz)\n", name, ms, per_smsp_cycle, ghz);
}

int main()
{
	cudaDeviceProp p;
	cudaGetDeviceProperties(&p, 0);
	int khz = 0;
	cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
	double const ghz = khz * 1e-6;
	uint32_t *d_out;
	cudaMalloc(&d_out, (size_t) p.multiProcessorCount * 8 * 256 * 4);
#define X(n, s) run <OP_##n>(#n, d_out, p.multiProcessorCount, ghz);
	OPS(X)
#undef X
	return 0;
}
