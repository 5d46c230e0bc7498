// C ABI of libfsq_b200.so (include/fsq.h): contexts, device memory, and the orchestration of the kernels.
// There is no CPU fallback anywhere in this file: every compute entry point launches the CUDA kernels or fails.
#include "../../include/fsq.h"
#include "fsq_internal.h"
#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <chrono>
#include <map>
#include <new>
#include <string>
#include <vector>

using namespace fsq;

namespace {

thread_local std::string g_create_error;

struct phase_timer;

} // namespace

struct fsq_ctx
{
	int device{0};
	cudaStream_t stream{nullptr};
	cudaDeviceProp prop{};
	std::string error;
	std::map <std::string, int64_t> opts;
	std::map <std::string, double> stats;
	std::vector <std::pair <std::string, std::pair <cudaEvent_t, cudaEvent_t>>> pending;   // phase timers
	cudaEvent_t timer_ev[2]{nullptr, nullptr};   // option "timer": 1 records the start, 0 the stop, on the context's stream ("ms_timer")
	cudaEvent_t call_start{nullptr};       // recorded at the entry of fsq_segment: phases report their offset from it ("at_<phase>")

	int64_t opt(char const *k, int64_t dflt) const { auto it = opts.find(k); return opts.end() == it ? dflt : it->second; }

	// Workspace cache: device and pinned-host blocks released by one call are reused by the next instead of going
	// through cudaMalloc / cudaFree (which synchronise the device).  Everything runs on the context's one stream, so a
	// block handed out again is only touched after the work that used it before.
	struct block { void *p; size_t bytes; };
	std::vector <block> dev_cache, host_cache;

	cudaError_t take(bool host, size_t bytes, void **out, size_t *got)
	{
		std::vector <block> &cache = host ? host_cache : dev_cache;
		bytes = (std::max <size_t>(bytes, 16) + 255) & ~size_t(255);
		size_t best = cache.size();
		for (size_t i = 0; i < cache.size(); ++i)
			if (cache[i].bytes >= bytes && cache[i].bytes <= bytes + bytes / 2 && (best == cache.size() || cache[i].bytes < cache[best].bytes))
				best = i;
		if (best != cache.size())
		{
			*out = cache[best].p; *got = cache[best].bytes;
			cache[best] = cache.back(); cache.pop_back();
			return cudaSuccess;
		}
		cudaError_t e = host ? cudaMallocHost(out, bytes) : cudaMalloc(out, bytes);
		if (cudaErrorMemoryAllocation == e)
		{
			cudaGetLastError();
			trim(host);
			e = host ? cudaMallocHost(out, bytes) : cudaMalloc(out, bytes);
		}
		*got = bytes;
		return e;
	}
	void give(bool host, void *p, size_t bytes) { if (p) (host ? host_cache : dev_cache).push_back({p, bytes}); }
	void trim(bool host)
	{
		std::vector <block> &cache = host ? host_cache : dev_cache;
		if (!cache.empty()) cudaStreamSynchronize(stream);
		for (auto &b : cache) { if (host) cudaFreeHost(b.p); else cudaFree(b.p); }
		cache.clear();
	}
};

struct fsq_msa
{
	fsq_ctx *ctx{};
	uint8_t *cols{};
	size_t cols_bytes{};       // 0: from cudaMalloc; else the size of the block taken from the context's cache
	uint64_t col_stride{};
	uint32_t m{}, nsym{};
	uint32_t sym_mask{};       // bit c set: comp value c occurs in some row
	uint64_t n{}, col0{};
};

struct fsq_segmentation
{
	fsq_ctx *ctx{};
	uint32_t m{};
	uint64_t n{}, L{};
	uint32_t K{};
	std::vector <uint64_t> lb, rb;            // reduced
	std::vector <uint32_t> size;
	uint32_t *d_a{}, *d_d{};                  // reduced states on the device, count x m
	uint32_t *h_a{}, *h_d{};                  // pinned host copies, made on first request
	size_t d_bytes[2]{}, h_bytes[2]{};        // sizes of those blocks in the context's workspace cache
	void *h_tb{};                             // traceback before reduction, pinned: rb u64[T], then lb, max, size u32[T] each
	size_t h_tb_bytes{};
	uint32_t T{};
	std::vector <uint32_t> dp_lb, dp_max, dp_size;   // keep_dp
};

namespace {

int fail(fsq_ctx *ctx, int code, char const *fmt, ...)
{
	char buf[512];
	va_list ap;
	va_start(ap, fmt);
	vsnprintf(buf, sizeof buf, fmt, ap);
	va_end(ap);
	if (ctx) ctx->error = buf; else g_create_error = buf;
	return code;
}

#define CK_(ctx, call) do { cudaError_t e_ = (call); if (cudaSuccess != e_) return fail(ctx, FSQ_ERR_CUDA, "%s:%d: %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); } while (0)
#define CK(call) do { cudaError_t e_ = (call); if (cudaSuccess != e_) return fail(ctx, FSQ_ERR_CUDA, "%s:%d: %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); } while (0)

// RAII workspace buffers (device / pinned host) out of the context's cache
template <bool HOST>
struct wbuf
{
	void *p{nullptr};
	size_t bytes{0};
	fsq_ctx *owner{nullptr};
	wbuf() = default;
	wbuf(wbuf const &) = delete;
	wbuf &operator=(wbuf const &) = delete;
	~wbuf() { release(); }
	void release() { if (p) owner->give(HOST, p, bytes); p = nullptr; }
	cudaError_t alloc(fsq_ctx *ctx, size_t want) { release(); owner = ctx; return ctx->take(HOST, want, &p, &bytes); }
	void *detach() { void *q = p; p = nullptr; return q; }
	template <typename T> T *as() const { return static_cast <T *>(p); }
};
typedef wbuf <false> dbuf;
typedef wbuf <true> hbuf;

struct phase
{
	fsq_ctx *ctx; std::string name; cudaEvent_t e0{}, e1{};
	phase(fsq_ctx *c, char const *n): ctx(c), name(n)
	{
		cudaEventCreate(&e0); cudaEventCreate(&e1);
		cudaEventRecord(e0, ctx->stream);
	}
	void stop()
	{
		cudaEventRecord(e1, ctx->stream);
		ctx->pending.push_back({name, {e0, e1}});
	}
};

// Host-side stopwatch for the stages whose time is not kernels (statistic "hms_<name>", milliseconds, accumulated).
struct host_watch
{
	fsq_ctx *ctx; char const *name; std::chrono::steady_clock::time_point t0;
	host_watch(fsq_ctx *c, char const *n): ctx(c), name(n), t0(std::chrono::steady_clock::now()) {}
	~host_watch() { ctx->stats[std::string("hms_") + name] += std::chrono::duration <double, std::milli>(std::chrono::steady_clock::now() - t0).count(); }
};

void flush_phases(fsq_ctx *ctx)
{
	for (auto &p : ctx->pending)
	{
		float ms = 0;
		cudaEventSynchronize(p.second.second);
		cudaEventElapsedTime(&ms, p.second.first, p.second.second);
		ctx->stats["ms_" + p.first] += ms;
		if (ctx->call_start)
		{
			float at = 0;
			if (cudaSuccess == cudaEventElapsedTime(&at, ctx->call_start, p.second.first)) ctx->stats["at_" + p.first] = at;
			if (cudaSuccess == cudaEventElapsedTime(&at, ctx->call_start, p.second.second)) ctx->stats["end_" + p.first] = at;
		}
		cudaEventDestroy(p.second.first);
		cudaEventDestroy(p.second.second);
	}
	ctx->pending.clear();
}

// The geometry of one scan: warps per block (column range), tiles per warp, columns per block.
struct geometry
{
	uint32_t n_tiles, n_warps, tpw, C, cg, B;
	// blocked column step (fsq_bstep.cuh): E elements per thread, T threads, symbols sym0 .. sym0 + ns - 1
	bool blocked;
	bool chunked;          // fsq_cstep.cuh (the default where it applies); otherwise blocked = fsq_bstep.cuh, else striped
	uint32_t E, T, ns, sym0;
};

geometry choose_geometry(fsq_ctx *ctx, fsq_msa const *msa)
{
	geometry g;
	memset(&g, 0, sizeof g);
	uint32_t const m = msa->m;
	g.n_tiles = (m + 31) / 32;
	int64_t w = ctx->opt("warps", 0);
	if (w <= 0)
	{
		w = 1;
		while (w < 32 && (g.n_tiles + w - 1) / w > 16) w *= 2;
	}
	g.n_warps = (uint32_t) std::min <int64_t>(32, std::max <int64_t>(1, w));
	g.tpw = (g.n_tiles + g.n_warps - 1) / g.n_warps;
	while (g.tpw > 32 && g.n_warps < 32) { g.n_warps *= 2; g.tpw = (g.n_tiles + g.n_warps - 1) / g.n_warps; }

	// The blocked step walks E consecutive positions per thread; it needs the symbols that occur to span at most 8 values.
	uint32_t const mask = msa->sym_mask ? msa->sym_mask : ((1u << msa->nsym) - 1u);
	uint32_t lo = 0, hi = 31;
	while (!((mask >> lo) & 1u)) ++lo;
	while (!((mask >> hi) & 1u)) --hi;
	g.sym0 = lo;
	g.ns = hi - lo + 1;
	int64_t const kern = ctx->opt("kernel", 0);        // 0 auto, 1 striped, 2 blocked, 3 chunked
	g.chunked = (0 == kern || 3 == kern) && g.ns <= 8 && m <= cblock_thread_limit(53) * 53u;
	g.blocked = !g.chunked && (1 != kern) && g.ns <= 8 && m <= 448u * 33u;
	uint32_t threads = g.n_warps * 32, regs = 64;
	if (g.chunked)
	{
		// Chunk length: a thread's column costs about (14 + E) element times (scans across the chunk summaries, unpacking and
		// the carry-in fix-up are per thread), and the padding up to whole warps is worked on like real rows.
		static uint32_t const lengths[] = {
#define FSQ_LIST(E) E,
			FSQ_CHUNK_LENGTHS(FSQ_LIST)
#undef FSQ_LIST
		};
		int64_t const want = ctx->opt("chunk", 0);
		uint64_t best = ~0ull;
		for (int pass = 0; pass < 2 && ~0ull == best; ++pass)
			for (uint32_t e : lengths)
			{
				if (0 == pass && 0 != want && want != (int64_t) e) continue;
				uint64_t const t = ((m + e - 1) / e + 31) / 32 * 32;
				if (t > cblock_thread_limit(e)) continue;
				// E = 53 exists for row counts the others cannot hold: 168 registers with a few spills and a loop body beyond the
				// instruction cache make it ~1.4 times slower per element than the model says (5,000 rows, measured)
				uint64_t const cost = t * (14 + e) * (e > 33 ? 3 : 2);
				if (cost < best) { best = cost; g.E = e; g.T = (uint32_t) t; }
			}
		threads = g.T;
		regs = cblock_registers(g.E);
	}
	else if (g.blocked)
	{
		int64_t e = ctx->opt("chunk", 0);
		if (9 != e && 17 != e && 33 != e)
			e = m <= 32 * 9 ? 9 : (m <= 32 * 17 ? 17 : 33);
		uint32_t const tmax = 9 == e ? 1024 : (17 == e ? 512 : 448);
		if ((uint64_t) tmax * e < m) e = 33;
		g.E = (uint32_t) e;
		g.T = ((m + g.E - 1) / g.E + 31) / 32 * 32;
		threads = g.T;
		regs = 9 == g.E ? 64 : 128;
	}

	g.cg = (uint32_t) std::max <uint64_t>(1, (g.chunked ? 512 : 2048) / msa->col_stride);
	int64_t C = ctx->opt("block_columns", 0);
	if (C <= 0)
	{
		// enough blocks to fill every SM once (the same blocks serve passes A and B, so the pass with the larger
		// footprint decides), but long enough to amortise the two extra passes' fixed costs
		uint64_t per_sm = std::min <uint64_t>(32, 2048 / threads);
		per_sm = std::min <uint64_t>(per_sm, 65536 / ((uint64_t) threads * regs));
		size_t smem = 2 * (size_t) g.cg * msa->col_stride + 4 * (size_t) threads * (g.blocked ? g.E + 2 * g.ns : 1) + 2 * (size_t) m + 1024;
		if (g.chunked)
		{
			// The histogram of pass B grows with the block length, so the footprint depends on the answer: take the largest
			// number of resident CTAs whose own block length fits (every CTA also reserves 1 KB of shared memory).
			per_sm = std::min <uint64_t>(per_sm, 65536 / ((uint64_t) threads * regs));
			for (; per_sm > 1; --per_sm)
			{
				block_args P;
				memset(&P, 0, sizeof P);
				P.cg = g.cg; P.col_stride = msa->col_stride; P.m = m;
				uint64_t const res = (uint64_t) ctx->prop.multiProcessorCount * per_sm;
				P.C = (uint32_t) std::min <uint64_t>(4096, std::max <uint64_t>(64, (msa->n + res - 1) / res));
				if ((cblock_smem_bytes(P, g.T, g.E, g.ns, true) + 1024) * per_sm <= (size_t) ctx->prop.sharedMemPerMultiprocessor) break;
			}
			smem = 0;
		}
		if (smem) per_sm = std::min <uint64_t>(per_sm, (227 * 1024) / smem);
		per_sm = std::max <uint64_t>(per_sm, 1);
		uint64_t const resident = (uint64_t) ctx->prop.multiProcessorCount * per_sm;
		C = (int64_t) ((msa->n + resident - 1) / resident);
		C = std::max <int64_t>(C, 64);
		int64_t c_max = std::min <int64_t>(4096, 65534 - (int64_t) m - 2);
		if (g.chunked)
		{
			// one CTA per SM and still too large (many rows): the histogram has one counter per row and per column of the block
			block_args P;
			memset(&P, 0, sizeof P);
			P.cg = g.cg; P.col_stride = msa->col_stride; P.m = m; P.C = 0;
			size_t const fixed = cblock_smem_bytes(P, g.T, g.E, g.ns, true) + 1024;
			size_t const room = (size_t) ctx->prop.sharedMemPerBlockOptin > fixed ? ((size_t) ctx->prop.sharedMemPerBlockOptin - fixed) / 4 : 0;
			c_max = std::min <int64_t>(c_max, (int64_t) std::max <size_t>(room, 1));
		}
		if (C > c_max)
		{
			// More columns than one resident wave of the longest blocks holds: whole waves of equal blocks, not blocks of
			// the maximal length whose last wave fills a part of the SMs (2,500 x 5M: 1,221 blocks of 4,096 columns on 740
			// resident CTAs = 1.65 waves; 1,480 blocks of 3,379 columns are two full ones).
			c_max = std::max <int64_t>(c_max, 1);
			uint64_t const waves = (msa->n + resident * (uint64_t) c_max - 1) / (resident * (uint64_t) c_max);
			C = (int64_t) ((msa->n + waves * resident - 1) / (waves * resident));
			if (0 != ctx->opt("whole_waves", 1)) C = std::min <int64_t>(C, c_max); else C = c_max;
		}
	}
	C = std::min <int64_t>(C, 65534 - (int64_t) m - 2);
	C = std::max <int64_t>(C, 1);
	g.C = (uint32_t) C;
	g.cg = std::min(g.cg, g.C);
	g.B = (uint32_t) ((msa->n + g.C - 1) / g.C);
	return g;
}

// One launch of the column-block scan in the geometry's flavour.
cudaError_t launch_scan(block_args const &A, uint32_t n_blocks, geometry const &g, bool hist, cudaStream_t st)
{
	if (g.chunked)
	{
		switch (g.E)
		{
#define FSQ_CASE(E) case E: return launch_cblock_e##E(A, n_blocks, g.T, g.ns, hist, st);
			FSQ_CHUNK_LENGTHS(FSQ_CASE)
#undef FSQ_CASE
			default: return cudaErrorInvalidConfiguration;
		}
	}
	if (!g.blocked) return launch_block(A, n_blocks, g.n_warps, g.tpw, hist, st);
	switch (g.E)
	{
		case 9: return launch_bblock_e9(A, n_blocks, g.T, g.ns, hist, st);
		case 17: return launch_bblock_e17(A, n_blocks, g.T, g.ns, hist, st);
		default: return launch_bblock_e33(A, n_blocks, g.T, g.ns, hist, st);
	}
}

void fill_common(block_args &A, fsq_msa const *msa, geometry const &g)
{
	memset(&A, 0, sizeof A);
	A.cols = msa->cols;
	A.col_stride = msa->col_stride;
	A.n_local = msa->n;
	A.col0 = msa->col0;
	A.m = msa->m;
	A.n_tiles = g.n_tiles;
	A.nsym = msa->nsym;
	A.sym0 = g.sym0;
	A.C = g.C;
	A.cg = g.cg;
	A.magic = 0x4B000000u;
}

__global__ void iota_kernel(uint32_t *a, uint32_t *d, uint32_t m, uint32_t dval)
{
	uint32_t const i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i < m) { a[i] = i; d[i] = dval; }
}

__global__ void gather_states_kernel(uint32_t const *src_a, uint32_t const *src_d, uint32_t const *index, uint32_t m, uint32_t *dst_a, uint32_t *dst_d)
{
	uint32_t const s = blockIdx.y;
	uint32_t const i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i < m)
	{
		size_t const from = (size_t) index[s] * m + i, to = (size_t) s * m + i;
		dst_a[to] = src_a[from];
		dst_d[to] = src_d[from];
	}
}

// States kept by pass B (packed row | code << 16, one row of `stride` words per column) -> (a, d) at the given (global) right
// bounds: the state before column rb is the one stored after column rb - 1; its codes are relative to that column's block.
__global__ void gather_stored_states_kernel(
	uint32_t const *states, uint64_t stride, uint64_t const *rbs, uint32_t m, uint32_t C, uint64_t col0,
	uint32_t const *oldvals, uint32_t const *R, uint32_t *dst_a, uint32_t *dst_d
)
{
	uint32_t const t = blockIdx.x;                 // grid.x: up to 2^31 - 1 positions
	uint32_t const i = blockIdx.y * blockDim.x + threadIdx.x;
	if (i >= m) return;
	uint64_t const k = rbs[t] - col0 - 1;          // local column whose output state this is (rbs are global right bounds)
	uint32_t const b = (uint32_t) (k / C);
	uint32_t const r = R[b];
	uint32_t const w = states[k * stride + i];
	uint32_t const code = w >> 16;
	uint32_t const v = code >= r ? (uint32_t) (col0 + (uint64_t) b * C + 1 + (code - r)) : oldvals[(size_t) b * m + code];
	dst_a[(size_t) t * m + i] = w & 0xffffu;
	dst_d[(size_t) t * m + i] = v;
}

// Everything the scans share: local summaries, true block start states, their rank compression.
struct scan_state
{
	geometry g;
	dbuf sum_a, sum_d;          // [B][m]: block summaries (pass A); after the up-sweep, summaries of aligned spans
	dbuf st_a, st_d;            // [B + 1][m]: state at every block start (slot B: at the end of the last block)
	dbuf code, oldvals, R, items, item_begin;
	hbuf h_items, h_begin;
	bool relative{false};      // st_* are still relative to the identity state at the first local column
};

// Levels of stitch items (one CTA per item, one launch per level), uploaded in one go.
struct stitch_plan
{
	std::vector <stitch_item> items;
	std::vector <size_t> level_begin{0};
	void end_level() { level_begin.push_back(items.size()); }
};

int run_stitch_plan(fsq_ctx *ctx, scan_state &S, stitch_plan const &plan, uint32_t m, std::vector <std::vector <std::pair <uint32_t *, uint32_t const *>>> const &copies)
{
	size_t const n = plan.items.size();
	if (n)
	{
		// pinned staging lives in the scan state: it is released only after the caller's final synchronisation
		CK(S.h_items.alloc(ctx, n * sizeof(stitch_item))); CK(S.h_begin.alloc(ctx, (n + 1) * 4));
		memcpy(S.h_items.p, plan.items.data(), n * sizeof(stitch_item));
		for (size_t i = 0; i <= n; ++i) S.h_begin.as <uint32_t>()[i] = (uint32_t) i;
		CK(S.items.alloc(ctx, n * sizeof(stitch_item))); CK(S.item_begin.alloc(ctx, (n + 1) * 4));
		CK(cudaMemcpyAsync(S.items.p, S.h_items.p, n * sizeof(stitch_item), cudaMemcpyHostToDevice, ctx->stream));
		CK(cudaMemcpyAsync(S.item_begin.p, S.h_begin.p, (n + 1) * 4, cudaMemcpyHostToDevice, ctx->stream));
	}
	for (size_t l = 0; l + 1 < plan.level_begin.size(); ++l)
	{
		if (l < copies.size())
			for (auto const &c : copies[l])
				CK(cudaMemcpyAsync(c.first, c.second, (size_t) m * 4, cudaMemcpyDeviceToDevice, ctx->stream));
		size_t const first = plan.level_begin[l], count = plan.level_begin[l + 1] - first;
		if (0 == count) continue;
		stitch_args SA;
		SA.items = S.items.as <stitch_item>();
		SA.item_begin = S.item_begin.as <uint32_t>() + first;
		SA.n_ctas = (uint32_t) count;
		SA.m = m;
		SA.n_tiles = S.g.n_tiles;
		SA.use_sort = ctx->opt("stitch_sort", 1) ? 1u : 0u;
		CK(launch_stitch(SA, ctx->stream));
		ctx->stats["kernel_launches"] += 1;
	}
	return FSQ_OK;
}

// Pass A (local scans) and the prefix composition of the block summaries (composing summaries is
// associative): an up-sweep turns the block summaries into summaries of aligned power-of-two spans in place, a down-sweep
// turns those into the state at every block start *relative to the identity state at the first local column* -- which
// for an alignment that starts at column 0 is the true state.  2 log2(B) levels of independent stitches.
int local_scan(fsq_ctx *ctx, fsq_msa const *msa, scan_state &S)
{
	S.g = choose_geometry(ctx, msa);
	geometry const &g = S.g;
	uint32_t const m = msa->m;
	uint32_t const B = g.B;
	size_t const bm = (size_t) B * m;
	CK(S.sum_a.alloc(ctx, bm * 4)); CK(S.sum_d.alloc(ctx, bm * 4));
	CK(S.st_a.alloc(ctx, (bm + m) * 4)); CK(S.st_d.alloc(ctx, (bm + m) * 4));
	CK(S.code.alloc(ctx, bm * 2)); CK(S.oldvals.alloc(ctx, bm * 4)); CK(S.R.alloc(ctx, (size_t) B * 4));
	uint32_t *sum_a = S.sum_a.as <uint32_t>(), *sum_d = S.sum_d.as <uint32_t>();
	uint32_t *st_a = S.st_a.as <uint32_t>(), *st_d = S.st_d.as <uint32_t>();

	iota_kernel <<<(m + 255) / 256, 256, 0, ctx->stream>>>(st_a, st_d, m, (uint32_t) msa->col0);
	CK(cudaGetLastError());
	S.relative = 0 != msa->col0;

	{
		phase pa(ctx, "pass_a");
		block_args A;
		fill_common(A, msa, g);
		A.mode = MODE_LOCAL;
		A.sum_a = sum_a;
		A.sum_d = sum_d;
		CK(launch_scan(A, B, g, false, ctx->stream));
		pa.stop();
		ctx->stats["kernel_launches"] += 1;
	}

	phase ps(ctx, "stitch");
	auto start_of = [&](uint64_t b) { return (uint32_t) (msa->col0 + b * g.C); };
	auto U = [&](uint64_t i, uint32_t *&a, uint32_t *&d) { a = sum_a + i * m; d = sum_d + i * m; };
	auto P = [&](uint64_t i, uint32_t *&a, uint32_t *&d) { a = st_a + (i + 1) * m; d = st_d + (i + 1) * m; };      // state at the end of block i
	uint32_t levels = 0;
	while ((1ull << (levels + 1)) <= B) ++levels;          // 2^levels <= B
	stitch_plan plan;
	std::vector <std::vector <std::pair <uint32_t *, uint32_t const *>>> copies;
	// up-sweep: U[i] <- U[i - 2^l] o U[i] for every i with (i + 1) a multiple of 2^(l + 1)
	for (uint32_t l = 0; l < levels; ++l)
	{
		uint64_t const half = 1ull << l, span = half << 1;
		for (uint64_t i = span - 1; i < B; i += span)
		{
			stitch_item it;
			memset(&it, 0, sizeof it);
			uint32_t *la, *ld, *ra, *rd;
			U(i - half, la, ld); U(i, ra, rd);
			it.in_a = la; it.in_d = ld;
			it.pi = ra; it.delta = rd;
			it.out_a = ra; it.out_d = rd;                   // in place: the kernel holds the item in shared memory before it writes
			it.s_mark = start_of(i + 1 - half);
			plan.items.push_back(it);
		}
		plan.end_level();
		copies.emplace_back();
	}
	// down-sweep: with 2^l the lowest set bit of i + 1, P[i] = U[i] when i + 1 == 2^l, else P[i - 2^l] o U[i]
	for (int l = (int) levels; l >= 0; --l)
	{
		uint64_t const half = 1ull << l, span = half << 1;
		copies.emplace_back();
		for (uint64_t i = half - 1; i < B; i += span)
		{
			uint32_t *ua, *ud, *pa_, *pd_;
			U(i, ua, ud); P(i, pa_, pd_);
			if (i + 1 == half)
			{
				copies.back().push_back({pa_, ua});
				copies.back().push_back({pd_, ud});
				continue;
			}
			stitch_item it;
			memset(&it, 0, sizeof it);
			uint32_t *qa, *qd;
			P(i - half, qa, qd);
			it.in_a = qa; it.in_d = qd;
			it.pi = ua; it.delta = ud;
			it.out_a = pa_; it.out_d = pd_;
			it.s_mark = start_of(i + 1 - half);
			plan.items.push_back(it);
		}
		plan.end_level();
	}
	{
		int const rc = run_stitch_plan(ctx, S, plan, m, copies);
		if (rc) return rc;
	}
	ps.stop();
	return FSQ_OK;
}

// Column-sharded operation: the states of local_scan() are relative to the identity state at the first local column;
// given the true state there (d_a, d_d on the device), every block start becomes (true state) o (relative state).
int apply_start_state(fsq_ctx *ctx, fsq_msa const *msa, scan_state &S, uint32_t const *d_a, uint32_t const *d_d)
{
	geometry const &g = S.g;
	uint32_t const m = msa->m;
	uint32_t *st_a = S.st_a.as <uint32_t>(), *st_d = S.st_d.as <uint32_t>();
	phase ps(ctx, "stitch");
	stitch_plan plan;
	for (uint64_t b = 1; b <= g.B; ++b)
	{
		stitch_item it;
		memset(&it, 0, sizeof it);
		it.in_a = d_a; it.in_d = d_d;
		it.pi = st_a + b * m; it.delta = st_d + b * m;
		it.out_a = st_a + b * m; it.out_d = st_d + b * m;
		it.s_mark = (uint32_t) msa->col0;
		plan.items.push_back(it);
	}
	plan.end_level();
	CK(cudaStreamSynchronize(ctx->stream));      // the staging buffers of the scan state are about to be rewritten
	int const rc = run_stitch_plan(ctx, S, plan, m, {});
	if (rc) return rc;
	CK(cudaMemcpyAsync(st_a, d_a, (size_t) m * 4, cudaMemcpyDeviceToDevice, ctx->stream));
	CK(cudaMemcpyAsync(st_d, d_d, (size_t) m * 4, cudaMemcpyDeviceToDevice, ctx->stream));
	ps.stop();
	S.relative = false;
	return FSQ_OK;
}

// Rank compression of the (true) block start divergences.
int compress_starts(fsq_ctx *ctx, fsq_msa const *msa, scan_state &S)
{
	phase pp(ctx, "prep");
	CK(launch_prep(S.st_d.as <uint32_t>(), msa->m, S.code.as <uint16_t>(), S.oldvals.as <uint32_t>(), S.R.as <uint32_t>(), 0, S.g.B, ctx->stream));
	pp.stop();
	ctx->stats["kernel_launches"] += 1;
	return FSQ_OK;
}

// Pass A, stitch, rank compression for an alignment that starts at column 0.  After this, st_[ad] hold the exact PBWT
// state at every block start.
int prepare_scan(fsq_ctx *ctx, fsq_msa const *msa, scan_state &S)
{
	int rc = local_scan(ctx, msa, S);
	if (rc) return rc;
	if (S.relative) return fail(ctx, FSQ_ERR_ARG, "this entry point needs an alignment that starts at column 0 (column shards go through fsq_shard_*)");
	return compress_starts(ctx, msa, S);
}

// Pass C: (a_c, d_c) at the given ascending local columns c (1 <= c <= n) into dump_a / dump_d (count x m).
int dump_states(fsq_ctx *ctx, fsq_msa const *msa, scan_state &S, std::vector <uint64_t> const &columns, uint32_t *dump_a, uint32_t *dump_d)
{
	geometry const &g = S.g;
	std::vector <uint32_t> begin(g.B + 1, 0), ids;
	{
		size_t i = 0;
		for (uint32_t b = 0; b < g.B; ++b)
		{
			begin[b] = (uint32_t) i;
			uint64_t const e = std::min <uint64_t>(msa->n, (uint64_t) (b + 1) * g.C);
			while (i < columns.size() && columns[i] <= e) ++i;     // block b owns right bounds in (s, e]
			if (i > begin[b]) ids.push_back(b);
		}
		begin[g.B] = (uint32_t) i;
		if (i != columns.size())
			return fail(ctx, FSQ_ERR_ARG, "state columns must be ascending and within [1, n]");
	}
	if (ids.empty()) return FSQ_OK;
	dbuf d_cols, d_begin, d_ids;
	CK(d_cols.alloc(ctx, columns.size() * 8)); CK(d_begin.alloc(ctx, begin.size() * 4)); CK(d_ids.alloc(ctx, ids.size() * 4));
	CK(cudaMemcpyAsync(d_cols.p, columns.data(), columns.size() * 8, cudaMemcpyHostToDevice, ctx->stream));
	CK(cudaMemcpyAsync(d_begin.p, begin.data(), begin.size() * 4, cudaMemcpyHostToDevice, ctx->stream));
	CK(cudaMemcpyAsync(d_ids.p, ids.data(), ids.size() * 4, cudaMemcpyHostToDevice, ctx->stream));
	phase pc(ctx, "pass_c");
	block_args A;
	fill_common(A, msa, g);
	A.mode = MODE_DUMP;
	A.st_a = S.st_a.as <uint32_t>();
	A.code = S.code.as <uint16_t>();
	A.oldvals = S.oldvals.as <uint32_t>();
	A.R = S.R.as <uint32_t>();
	A.blk_ids = d_ids.as <uint32_t>();
	A.dump_cols = d_cols.as <uint64_t>();
	A.dump_begin = d_begin.as <uint32_t>();
	A.dump_a = dump_a;
	A.dump_d = dump_d;
	CK(launch_scan(A, (uint32_t) ids.size(), g, false, ctx->stream));
	pc.stop();
	ctx->stats["kernel_launches"] += 1;
	CK(cudaStreamSynchronize(ctx->stream));   // the three staging buffers die here
	return FSQ_OK;
}

} // namespace

// One run of the long path over the columns one rank holds (all of them on a single GPU), cut into the stages between
// which column-sharded ranks exchange data.  fsq_segment() runs the stages back to back; fsq_shard_*() exposes them.
struct fsq_shard
{
	fsq_ctx *ctx{};
	fsq_msa const *msa{};
	uint32_t m{};
	uint64_t n{};               // columns of the whole alignment
	uint64_t L{};
	uint64_t c0{}, c1{};        // this rank's columns [c0, c1)
	scan_state S;
	bool started{false};        // block start states are true and compressed

	// DP: the table is global (every rank holds all of it), the top lists are local
	uint64_t dp_size{};
	uint32_t n_entries{}, n_blk{}, n_levels{}, t_first{}, t_last{};
	uint32_t chunk{}, n_chunks{}, cap{}, rounds{};
	uint32_t reach_known{};     // the farthest look-back any rank has seen in the rounds so far
	bool single_chain{false};
	dbuf tb_max, tb_lb, tb_size, mk, lev, flags, chunk_level, cnt, h0, toplist, states;
	uint64_t state_stride{};

	// traceback (replicated) and reduction
	uint32_t T{}, K{};
	dbuf tr_lb, tr_rb, tr_max, tr_size, lists, d_next, d_nsize;
	hbuf h_tb;                                 // the traceback on the host (pinned): rb u64[T], then lb, max, size u32[T] each
	uint32_t pos_first{}, pos_last{};          // traceback positions whose right bound this rank holds the state of
	std::vector <uint64_t> red_lb, red_rb;
	std::vector <uint32_t> red_size, red_pos;  // reduced segments; red_pos: index into the traceback
	uint32_t seg_first{}, seg_last{};          // reduced segments whose state this rank holds
	uint32_t n_reduced{};
	dbuf red_a, red_d;                         // [seg_last - seg_first][m]
	dbuf d_red_lb, d_red_rb, d_red_size, walk_jump, walk_info;      // the reduced segments on the device (launch_walk)
	hbuf h_red, h_info;                        // their host staging; red_pending: copies in flight, vectors not yet filled
	bool red_pending{false}, red_on_host{false};
	bool final_done{false};                    // the last DP entry was computed behind the final round already
	uint32_t final_overflow{};

	dp_args dp() const
	{
		dp_args D;
		memset(&D, 0, sizeof D);
		D.m = m; D.n = n; D.L = L; D.col_base = c0; D.t_first = t_first; D.t_last = t_last;
		D.toplist = toplist.as <uint2>(); D.cap = cap; D.cnt = cnt.as <uint32_t>(); D.h0 = h0.as <uint32_t>();
		D.tb_max = tb_max.as <uint32_t>(); D.tb_lb = tb_lb.as <uint32_t>(); D.tb_size = tb_size.as <uint32_t>();
		D.mk = mk.as <uint64_t>();
		D.lev = lev.as <uint32_t>(); D.n_blk = n_blk; D.n_levels = n_levels;
		D.n_entries = n_entries;
		D.chunk = single_chain ? (n_entries + 63) / 64 * 64 : chunk;
		D.n_chunks = single_chain ? 1 : n_chunks;
		D.chunk_level = chunk_level.as <uint32_t>();
		D.flags = flags.as <uint32_t>();
		D.reach_bound = reach_known;
		{
			// what this rank's chains can read: its own entries and the look-back before them (twice the bound, for slack)
			uint64_t const back = 2 * (uint64_t) reach_known + 128;
			D.w_first = (uint32_t) ((t_first > back ? t_first - back : 0) & ~63ull);
			D.w_last = n_entries;
			if (c1 != n) D.w_last = (uint32_t) std::min <uint64_t>(n_entries, t_last);      // nothing this rank reads lies past its own entries
		}
		return D;
	}
};

namespace {

int shard_init(fsq_shard *P, fsq_ctx *ctx, fsq_msa const *msa, uint64_t n_total, uint64_t L)
{
	P->ctx = ctx; P->msa = msa; P->m = msa->m; P->n = n_total; P->L = L;
	P->c0 = msa->col0; P->c1 = msa->col0 + msa->n;
	uint64_t const n = n_total;
	P->dp_size = n - L + 1;
	P->n_entries = (uint32_t) (n - 2 * L + 1);
	P->n_blk = (uint32_t) (P->dp_size / 64 + 1);
	P->n_levels = 1;
	while ((1ull << P->n_levels) <= P->n_blk) ++P->n_levels;
	// entry t is computed at column t + L - 1
	uint64_t const lo = P->c0 + 1 > L ? P->c0 + 1 - L : 0, hi = P->c1 + 1 > L ? P->c1 + 1 - L : 0;
	P->t_first = (uint32_t) std::min <uint64_t>(lo, P->n_entries);
	P->t_last = (uint32_t) std::min <uint64_t>(hi, P->n_entries);
	// chains: enough to fill the device a few times over, each a whole number of 64-entry rmq blocks, on a global grid
	{
		int64_t want = ctx->opt("dp_chunks", 0);
		if (want <= 0) want = (int64_t) ctx->prop.multiProcessorCount * (2048 / (32 * (int64_t) std::min <uint64_t>(L, 16)) > 8 ? 8 : 2048 / (32 * (int64_t) std::min <uint64_t>(L, 16)));      // one resident wave of chains
		uint64_t const own = std::max <uint64_t>(1, P->t_last - P->t_first);
		uint64_t c = (own + (uint64_t) want - 1) / (uint64_t) want;
		c = std::max <uint64_t>(c, std::max <uint64_t>(256, 8 * L));
		if (1 == want) c = P->n_entries;
		c = (c + 63) / 64 * 64;
		P->chunk = (uint32_t) std::min <uint64_t>(c, 0xffffffc0ull);
		P->n_chunks = (P->n_entries + P->chunk - 1) / P->chunk;
	}
	P->cap = (uint32_t) std::min <int64_t>(P->m, std::max <int64_t>(1, ctx->opt("cap", 127)));
	P->state_stride = ((uint64_t) P->m + 3) & ~3ull;
	return FSQ_OK;
}

// Pass B over this rank's columns with the current top-list bound; resets the DP table.
int shard_scan(fsq_shard *P)
{
	fsq_ctx *ctx = P->ctx;
	fsq_msa const *msa = P->msa;
	geometry const &g = P->S.g;
	uint32_t const m = P->m;
	uint64_t const nl = msa->n;
	if (!P->tb_max.p)
	{
		CK(P->tb_max.alloc(ctx, P->dp_size * 4)); CK(P->tb_lb.alloc(ctx, P->dp_size * 4)); CK(P->tb_size.alloc(ctx, P->dp_size * 4));
		CK(P->mk.alloc(ctx, P->dp_size * 8)); CK(P->lev.alloc(ctx, (size_t) P->n_levels * P->n_blk * 4)); CK(P->flags.alloc(ctx, 32));
		CK(P->cnt.alloc(ctx, nl * 4)); CK(P->h0.alloc(ctx, nl * 4));
		CK(P->chunk_level.alloc(ctx, (size_t) P->n_chunks * 8));
		// Pass B keeps the state of every column when 4 m n bytes fit comfortably (option store_states: -1 auto, 0 never,
		// 1 always); the states at the segment boundaries are then a gather instead of a third scan.
		int64_t const want = ctx->opt("store_states", -1);
		bool keep = false;
		if (g.chunked && 0 != want)
		{
			size_t free_b = 0, total_b = 0;
			CK(cudaMemGetInfo(&free_b, &total_b));      // (a driver round trip per call; the answer decides only the first time)
			size_t cached = 0;
			for (auto const &blk : ctx->dev_cache) cached += blk.bytes;
			size_t const need = (size_t) (P->state_stride * 4 * nl);
			size_t const other = (size_t) nl * (P->cap + 1) * 8 + (size_t) nl * 64 + ((size_t) 1 << 30);
			keep = (1 == want) || need + other < free_b + cached;
			if (keep && cudaSuccess != P->states.alloc(ctx, need)) { cudaGetLastError(); keep = false; }
		}
		ctx->stats["store_states"] = keep ? 1 : 0;
	}
	CK(P->toplist.alloc(ctx, (size_t) nl * (P->cap + 1) * sizeof(uint2)));
	// entries the per-column DP writes start from 0 (the first round's stand-in for "nothing known"); the ones the
	// reference never writes keep its defaults; masks and rmq cells start from values that index inside the table
	CK(cudaMemsetAsync(P->tb_max.p, 0xff, P->dp_size * 4, ctx->stream));
	CK(cudaMemsetAsync(P->tb_max.p, 0, (size_t) P->n_entries * 4, ctx->stream));
	CK(cudaMemsetAsync(P->tb_size.p, 0xff, P->dp_size * 4, ctx->stream));
	CK(cudaMemsetAsync(P->tb_lb.p, 0, P->dp_size * 4, ctx->stream));
	CK(cudaMemsetAsync(P->mk.p, 0xff, P->dp_size * 8, ctx->stream));
	CK(cudaMemsetAsync(P->lev.p, 0, (size_t) P->n_levels * P->n_blk * 4, ctx->stream));
	CK(cudaMemsetAsync(P->flags.p, 0, 32, ctx->stream));
	P->rounds = 0;
	P->final_done = false;
	P->reach_known = 0;
	P->single_chain = false;
	{
		phase pb(ctx, "pass_b");
		block_args A;
		fill_common(A, msa, g);
		A.mode = MODE_EXACT;
		A.st_a = P->S.st_a.as <uint32_t>();
		A.code = P->S.code.as <uint16_t>();
		A.oldvals = P->S.oldvals.as <uint32_t>();
		A.R = P->S.R.as <uint32_t>();
		A.toplist = P->toplist.as <uint2>();
		A.cap = P->cap;
		A.cnt = P->cnt.as <uint32_t>();
		A.h0 = P->h0.as <uint32_t>();
		A.states = P->states.as <uint32_t>();
		A.state_stride = P->state_stride;
		CK(launch_scan(A, g.B, g, true, ctx->stream));
		pb.stop();
		ctx->stats["kernel_launches"] += 1;
	}
	return FSQ_OK;
}

// One DP round: every own chain once against the table as it stands.  changed: another round is needed -- in the first
// round, some stored value changed; later, some value changed that a chain after its owner may have read while it was
// still the old one (dp_chain_kernel: the last reach_known entries of a chain, or any rmq cell), or a chain read
// further back than that.  A round without such a change has computed every entry from final values.
// reach: the farthest any column has looked back so far (entries).
int shard_dp_round(fsq_shard *P, uint32_t *changed, uint32_t *reach)
{
	fsq_ctx *ctx = P->ctx;
	phase pd(ctx, "dp");
	dp_args const D = P->dp();
	CK(cudaMemsetAsync(P->flags.p, 0, 4, ctx->stream));
	CK(cudaMemsetAsync(P->flags.as <uint32_t>() + 3, 0, 4, ctx->stream));
	CK(launch_dp_chains(D, ctx->stream));
	ctx->stats["kernel_launches"] += 1;
	++P->rounds;
	uint32_t h_flags[4] = {0, 0, 0, 0};
	CK(cudaMemcpyAsync(h_flags, P->flags.p, 16, cudaMemcpyDeviceToHost, ctx->stream));
	pd.stop();
	CK(cudaStreamSynchronize(ctx->stream));
	*changed = (h_flags[0] && (1 == P->rounds || h_flags[3])) ? 1u : 0u;
	*reach = h_flags[2];
	P->reach_known = std::max(P->reach_known, h_flags[2]);
	ctx->stats["dp_reach"] = h_flags[2];
	return FSQ_OK;
}

// The first two rounds queued back to back, with the guess between them, and one read-back at the end: round 1 starts
// from nothing and can never be the last, and the guess takes the look-back from the device (flags[2]) anyway, so the
// host has no decision to make before round 2.  Single rank only (column shards exchange their slices after every round).
int shard_dp_first_two_rounds(fsq_shard *P, uint32_t *changed, uint32_t *reach)
{
	fsq_ctx *ctx = P->ctx;
	phase pd(ctx, "dp");
	dp_args D = P->dp();
	CK(cudaMemsetAsync(P->flags.p, 0, 4, ctx->stream));
	CK(cudaMemsetAsync(P->flags.as <uint32_t>() + 3, 0, 4, ctx->stream));
	CK(launch_dp_chains(D, ctx->stream));
	CK(launch_dp_guess(D, ctx->stream));
	CK(cudaMemsetAsync(P->flags.p, 0, 4, ctx->stream));
	CK(cudaMemsetAsync(P->flags.as <uint32_t>() + 3, 0, 4, ctx->stream));
	D.bound_on_device = 1;
	CK(launch_dp_chains(D, ctx->stream));
	// the last entry too, behind the same read-back: it stands if round 2 turns out to have been the last one
	CK(launch_dp_final(D, ctx->stream));
	ctx->stats["kernel_launches"] += 3 + 5;
	P->rounds += 2;
	uint32_t h_flags[4] = {0, 0, 0, 0};
	CK(cudaMemcpyAsync(h_flags, P->flags.p, 16, cudaMemcpyDeviceToHost, ctx->stream));
	pd.stop();
	CK(cudaStreamSynchronize(ctx->stream));
	*changed = (h_flags[0] && h_flags[3]) ? 1u : 0u;
	*reach = h_flags[2];
	P->reach_known = std::max(P->reach_known, h_flags[2]);
	ctx->stats["dp_reach"] = h_flags[2];
	if (0 == *changed)
	{
		P->final_done = true;
		P->final_overflow = h_flags[1];
		ctx->stats["dp_rounds"] += P->rounds;
		ctx->stats["dp_chunks"] = P->n_chunks;
	}
	return FSQ_OK;
}

// Between rounds: with guess, level the table (fsq_dp.cu) using the given look-back; rebuild masks and rmq tables.
int shard_dp_sync(fsq_shard *P, int guess, uint32_t reach)
{
	fsq_ctx *ctx = P->ctx;
	phase pd(ctx, "dp");
	dp_args const D = P->dp();
	P->reach_known = std::max(P->reach_known, reach);
	if (guess)
	{
		CK(cudaMemcpyAsync(P->flags.as <uint32_t>() + 2, &reach, 4, cudaMemcpyHostToDevice, ctx->stream));
		CK(launch_dp_guess(D, ctx->stream));
		ctx->stats["kernel_launches"] += 5;
	}
	else
	{
		CK(launch_dp_rebuild(D, ctx->stream));
		ctx->stats["kernel_launches"] += 2;
	}
	pd.stop();
	return FSQ_OK;        // (a copy from pageable host memory has left the source when cudaMemcpyAsync returns)
}

// The last entry (segmentation_lp_context.cc:165-185), on the rank that holds the last column; and this rank's
// "top list too short" flag (1 + column, 0 = none).
int shard_dp_final(fsq_shard *P, uint32_t *overflow)
{
	fsq_ctx *ctx = P->ctx;
	phase pd(ctx, "dp");
	if (P->c1 == P->n)
	{
		CK(launch_dp_final(P->dp(), ctx->stream));
		ctx->stats["kernel_launches"] += 1;
	}
	uint32_t h_flags[2] = {0, 0};
	CK(cudaMemcpyAsync(h_flags, P->flags.p, 8, cudaMemcpyDeviceToHost, ctx->stream));
	pd.stop();
	CK(cudaStreamSynchronize(ctx->stream));
	*overflow = h_flags[1];
	ctx->stats["dp_rounds"] += P->rounds;
	ctx->stats["dp_chunks"] = P->n_chunks;
	return FSQ_OK;
}

// Traceback over the (complete, global) table -- segmentation_lp_context.cc:191-224 -- on the device by pointer doubling;
// every rank does the same.  Leaves the traceback on the device (text order) and on the host.
int shard_traceback(fsq_shard *P)
{
	fsq_ctx *ctx = P->ctx;
	host_watch hw(ctx, "traceback");
	phase pt(ctx, "traceback");
	trace_args TA;
	memset(&TA, 0, sizeof TA);
	TA.tb_lb = P->tb_lb.as <uint32_t>(); TA.tb_max = P->tb_max.as <uint32_t>(); TA.tb_size = P->tb_size.as <uint32_t>();
	TA.L = (uint32_t) P->L; TA.dp_size = (uint32_t) P->dp_size;
	TA.levels = 1;
	while ((1ull << TA.levels) <= P->dp_size / P->L + 1) ++TA.levels;
	dbuf jump, count;
	CK(jump.alloc(ctx, (size_t) TA.levels * P->dp_size * 4)); CK(count.alloc(ctx, 4));
	TA.jump = jump.as <uint32_t>(); TA.count = count.as <uint32_t>();
	CK(launch_trace_tables(TA, ctx->stream));
	ctx->stats["kernel_launches"] += 1;
	uint32_t T = 0;
	CK(cudaMemcpyAsync(&T, count.p, 4, cudaMemcpyDeviceToHost, ctx->stream));
	CK(cudaStreamSynchronize(ctx->stream));
	P->T = T;
	CK(P->tr_lb.alloc(ctx, (size_t) T * 4)); CK(P->tr_rb.alloc(ctx, (size_t) T * 8)); CK(P->tr_max.alloc(ctx, (size_t) T * 4)); CK(P->tr_size.alloc(ctx, (size_t) T * 4));
	TA.out_lb = P->tr_lb.as <uint32_t>(); TA.out_rb = P->tr_rb.as <uint64_t>(); TA.out_max = P->tr_max.as <uint32_t>(); TA.out_size = P->tr_size.as <uint32_t>();
	CK(launch_trace_gather(TA, T, ctx->stream));
	ctx->stats["kernel_launches"] += 1;
	CK(P->h_tb.alloc(ctx, (size_t) T * 20));
	uint64_t *h_rb = P->h_tb.as <uint64_t>();
	uint32_t *h_lb = (uint32_t *) (h_rb + T), *h_max = h_lb + T, *h_size = h_max + T;
	CK(cudaMemcpyAsync(h_rb, P->tr_rb.p, (size_t) T * 8, cudaMemcpyDeviceToHost, ctx->stream));
	CK(cudaMemcpyAsync(h_lb, P->tr_lb.p, (size_t) T * 4, cudaMemcpyDeviceToHost, ctx->stream));
	CK(cudaMemcpyAsync(h_max, P->tr_max.p, (size_t) T * 4, cudaMemcpyDeviceToHost, ctx->stream));
	CK(cudaMemcpyAsync(h_size, P->tr_size.p, (size_t) T * 4, cudaMemcpyDeviceToHost, ctx->stream));
	pt.stop();
	CK(cudaStreamSynchronize(ctx->stream));
	(void) h_size;
	P->K = h_max[T - 1];
	// the traceback positions this rank holds the state of: right bounds in (c0, c1]
	P->pos_first = (uint32_t) (std::upper_bound(h_rb, h_rb + T, P->c0) - h_rb);
	P->pos_last = (uint32_t) (std::upper_bound(h_rb, h_rb + T, P->c1) - h_rb);
	return FSQ_OK;
}

// The rows of the reduction's count lists for the traceback positions this rank owns (the others arrive by exchange).
int shard_reduce_lists(fsq_shard *P)
{
	fsq_ctx *ctx = P->ctx;
	uint32_t const width = P->K + 1;
	CK(P->lists.alloc(ctx, (size_t) P->T * width * sizeof(uint2)));
	phase pr(ctx, "reduce");
	reduce_args R;
	memset(&R, 0, sizeof R);
	R.T = P->T; R.K = P->K; R.width = width;
	R.tb_rb = P->tr_rb.as <uint64_t>(); R.tb_size = P->tr_size.as <uint32_t>();
	R.lists = P->lists.as <uint2>();
	R.toplist = P->toplist.as <uint2>(); R.cap = P->cap; R.cnt = P->cnt.as <uint32_t>(); R.col_base = P->c0;
	R.pos_first = P->pos_first; R.pos_last = P->pos_last;
	CK(launch_reduce_lists(R, ctx->stream));
	ctx->stats["kernel_launches"] += 1;
	pr.stop();
	return FSQ_OK;
}

int start_reduced_copy(fsq_shard *P);

// Greedy reduction (segmentation_lp_context.cc:335-390) from the complete lists; every rank does the same.
int shard_reduce(fsq_shard *P, bool want_host_bounds)
{
	fsq_ctx *ctx = P->ctx;
	uint32_t const T = P->T;
	host_watch hw(ctx, "reduce");
	phase pr(ctx, "reduce");
	CK(P->d_next.alloc(ctx, (size_t) T * 4)); CK(P->d_nsize.alloc(ctx, (size_t) T * 4));
	reduce_args R;
	memset(&R, 0, sizeof R);
	R.T = T; R.K = P->K; R.width = P->K + 1;
	R.tb_rb = P->tr_rb.as <uint64_t>(); R.tb_size = P->tr_size.as <uint32_t>();
	R.lists = P->lists.as <uint2>();
	R.next = P->d_next.as <uint32_t>(); R.next_size = P->d_nsize.as <uint32_t>();
	CK(launch_reduce(R, ctx->stream));
	ctx->stats["kernel_launches"] += 1;
	// the walk u -> next[u] + 1 from 0, on the device; the host learns the segment count and which segments are its own
	walk_args W;
	memset(&W, 0, sizeof W);
	W.T = T;
	W.levels = 1;
	while ((1ull << W.levels) <= T) ++W.levels;
	W.next = P->d_next.as <uint32_t>(); W.next_size = P->d_nsize.as <uint32_t>(); W.tb_rb = P->tr_rb.as <uint64_t>();
	CK(P->walk_jump.alloc(ctx, (size_t) W.levels * T * 4)); CK(P->walk_info.alloc(ctx, 16)); CK(P->h_info.alloc(ctx, 16));
	CK(P->d_red_lb.alloc(ctx, (size_t) T * 8)); CK(P->d_red_rb.alloc(ctx, (size_t) T * 8)); CK(P->d_red_size.alloc(ctx, (size_t) T * 4));
	W.jump = P->walk_jump.as <uint32_t>(); W.info = P->walk_info.as <uint32_t>();
	W.out_lb = P->d_red_lb.as <uint64_t>(); W.out_rb = P->d_red_rb.as <uint64_t>(); W.out_size = P->d_red_size.as <uint32_t>();
	W.c0 = P->c0; W.c1 = P->c1;
	CK(launch_walk(W, ctx->stream));
	ctx->stats["kernel_launches"] += 1;
	CK(cudaMemcpyAsync(P->h_info.p, P->walk_info.p, 12, cudaMemcpyDeviceToHost, ctx->stream));
	pr.stop();
	CK(cudaStreamSynchronize(ctx->stream));
	uint32_t const n_red = P->h_info.as <uint32_t>()[0];
	P->seg_first = P->h_info.as <uint32_t>()[1];
	P->seg_last = P->h_info.as <uint32_t>()[2];
	P->n_reduced = n_red;
	P->red_lb.clear(); P->red_rb.clear(); P->red_size.clear();
	P->red_on_host = false;
	if (!want_host_bounds) return FSQ_OK;      // column shards fetch them on request (fsq_shard_bounds): they are replicated
	return start_reduced_copy(P);
}

// Host copies of the reduced bounds (what segmentation_container holds): started here, completed by finish_reduced().
int start_reduced_copy(fsq_shard *P)
{
	fsq_ctx *ctx = P->ctx;
	uint32_t const n_red = P->n_reduced;
	CK(P->h_red.alloc(ctx, (size_t) std::max <uint32_t>(n_red, 1) * 20));
	{
		uint64_t *h_lb = P->h_red.as <uint64_t>(), *h_rb = h_lb + n_red;
		uint32_t *h_sz = (uint32_t *) (h_rb + n_red);
		CK(cudaMemcpyAsync(h_lb, P->d_red_lb.p, (size_t) n_red * 8, cudaMemcpyDeviceToHost, ctx->stream));
		CK(cudaMemcpyAsync(h_rb, P->d_red_rb.p, (size_t) n_red * 8, cudaMemcpyDeviceToHost, ctx->stream));
		CK(cudaMemcpyAsync(h_sz, P->d_red_size.p, (size_t) n_red * 4, cudaMemcpyDeviceToHost, ctx->stream));
	}
	P->red_pending = true;
	return FSQ_OK;
}

// After the stream has drained: the reduced bounds as host vectors.
int finish_reduced(fsq_shard *P)
{
	fsq_ctx *ctx = P->ctx;
	if (P->red_on_host) return FSQ_OK;
	if (!P->red_pending)
	{
		int const rc = start_reduced_copy(P);
		if (rc) return rc;
	}
	CK(cudaStreamSynchronize(ctx->stream));
	uint32_t const R = P->n_reduced;
	uint64_t const *h_lb = P->h_red.as <uint64_t>(), *h_rb = h_lb + R;
	uint32_t const *h_sz = (uint32_t const *) (h_rb + R);
	P->red_lb.assign(h_lb, h_lb + R);
	P->red_rb.assign(h_rb, h_rb + R);
	P->red_size.assign(h_sz, h_sz + R);
	P->red_pending = false;
	P->red_on_host = true;
	return FSQ_OK;
}

// (a, d) at the right bounds of this rank's reduced segments (segmentation_lp_context.cc:229-332, update_pbwt_task.cc:13-35):
// a gather out of the states pass B kept, else a re-scan of the blocks that hold them (pass C).
int shard_reduced_states(fsq_shard *P)
{
	fsq_ctx *ctx = P->ctx;
	uint32_t const m = P->m;
	uint32_t const cnt = P->seg_last - P->seg_first;
	host_watch hw(ctx, "reduced_states");
	CK(P->red_a.alloc(ctx, (size_t) std::max <uint32_t>(cnt, 1) * m * 4)); CK(P->red_d.alloc(ctx, (size_t) std::max <uint32_t>(cnt, 1) * m * 4));
	if (cnt && P->states.p)
	{
		phase pc(ctx, "pass_c");
		dim3 const grid(cnt, (m + 255) / 256);
		gather_stored_states_kernel <<<grid, 256, 0, ctx->stream>>>(P->states.as <uint32_t>(), P->state_stride, P->d_red_rb.as <uint64_t>() + P->seg_first, m, P->S.g.C, P->c0,
			P->S.oldvals.as <uint32_t>(), P->S.R.as <uint32_t>(), P->red_a.as <uint32_t>(), P->red_d.as <uint32_t>());
		CK(cudaGetLastError());
		pc.stop();
		ctx->stats["kernel_launches"] += 1;
	}
	int rc = FSQ_OK;
	if (P->red_pending || (cnt && !P->states.p))
	{
		if ((rc = finish_reduced(P))) return rc;      // drains the stream
	}
	else
		CK(cudaStreamSynchronize(ctx->stream));
	if (cnt && P->states.p)
		P->states.release();
	else if (cnt)
	{
		std::vector <uint64_t> cols(cnt);
		for (uint32_t i = 0; i < cnt; ++i) cols[i] = P->red_rb[P->seg_first + i] - P->c0;
		rc = dump_states(ctx, P->msa, P->S, cols, P->red_a.as <uint32_t>(), P->red_d.as <uint32_t>());
		if (rc) return rc;
	}
	return FSQ_OK;
}

int check_shape(fsq_ctx *ctx, fsq_msa const *msa)
{
	if (!ctx || !msa) return fail(ctx, FSQ_ERR_ARG, "null context or alignment");
	if (msa->ctx != ctx) return fail(ctx, FSQ_ERR_ARG, "alignment belongs to another context");
	if (msa->m > 65000) return fail(ctx, FSQ_ERR_UNSUPPORTED, "m = %u rows: the scan kernels keep row ids and divergence codes in 16 bits (m <= 65000)", msa->m);
	if (msa->col0 + msa->n >= (1ull << 31)) return fail(ctx, FSQ_ERR_UNSUPPORTED, "n = %llu columns: positions are kept in 32 bits (n < 2^31)", (unsigned long long) msa->n);
	return FSQ_OK;
}

constexpr unsigned long long FSQ_MSA_CACHE_LIMIT = 4ull << 30;

int new_msa(fsq_ctx *ctx, uint32_t m, uint64_t n, uint64_t col0, uint32_t nsym, fsq_msa **out)
{
	if (!ctx || !out) return fail(ctx, FSQ_ERR_ARG, "null argument");
	if (0 == m || 0 == n) return fail(ctx, FSQ_ERR_ARG, "empty alignment (m = %u, n = %llu)", m, (unsigned long long) n);
	fsq_msa *msa = new (std::nothrow) fsq_msa;
	if (!msa) return fail(ctx, FSQ_ERR_NOMEM, "out of host memory");
	msa->ctx = ctx; msa->m = m; msa->n = n; msa->col0 = col0; msa->nsym = nsym;
	msa->col_stride = (((uint64_t) m + 1) / 2 + 15) & ~15ull;
	unsigned long long const bytes = msa->col_stride * n;
	// Alignments of a few GB come from (and return to) the context's cache, so that a caller that uploads one alignment
	// after another pays no cudaMalloc / cudaFree (each synchronises the device) per call; larger ones are not kept.
	cudaError_t e;
	if (bytes <= FSQ_MSA_CACHE_LIMIT)
	{
		void *p = nullptr;
		e = ctx->take(false, bytes, &p, &msa->cols_bytes);        // trims the cache and retries when memory is short
		msa->cols = static_cast <uint8_t *>(p);
	}
	else
	{
		e = cudaMalloc(&msa->cols, bytes);
		if (cudaErrorMemoryAllocation == e)
		{
			cudaGetLastError();
			ctx->trim(false);        // give the workspace cache back and try once more
			e = cudaMalloc(&msa->cols, bytes);
		}
	}
	if (cudaSuccess != e)
	{
		delete msa;
		return fail(ctx, cudaErrorMemoryAllocation == e ? FSQ_ERR_NOMEM : FSQ_ERR_CUDA, "cudaMalloc of %llu bytes for the packed alignment: %s", bytes, cudaGetErrorString(e));
	}
	*out = msa;
	return FSQ_OK;
}

} // namespace


extern "C" {

const char *fsq_version(void) { return "fsq_b200 0.1 (sm_100a)"; }

int fsq_create(int device, fsq_ctx **out)
{
	if (!out) return fail(nullptr, FSQ_ERR_ARG, "null out pointer");
	int count = 0;
	cudaError_t e = cudaGetDeviceCount(&count);
	if (cudaSuccess != e || 0 == count)
		return fail(nullptr, FSQ_ERR_CUDA, "no CUDA device (%s); libfsq_b200 has no CPU fallback", cudaSuccess != e ? cudaGetErrorString(e) : "device count is 0");
	if (device < 0 || device >= count) return fail(nullptr, FSQ_ERR_ARG, "device %d out of range [0, %d)", device, count);
	fsq_ctx *ctx = new (std::nothrow) fsq_ctx;
	if (!ctx) return fail(nullptr, FSQ_ERR_NOMEM, "out of host memory");
	ctx->device = device;
	if (cudaSuccess != (e = cudaSetDevice(device)) || cudaSuccess != (e = cudaGetDeviceProperties(&ctx->prop, device))
		|| cudaSuccess != (e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking)))
	{
		delete ctx;
		return fail(nullptr, FSQ_ERR_CUDA, "device %d: %s", device, cudaGetErrorString(e));
	}
	if (ctx->prop.major < 10)
	{
		int const major = ctx->prop.major, minor = ctx->prop.minor;
		cudaStreamDestroy(ctx->stream);
		delete ctx;
		return fail(nullptr, FSQ_ERR_CUDA, "device %d is sm_%d%d; this library is built for sm_100a only", device, major, minor);
	}
	*out = ctx;
	return FSQ_OK;
}

void fsq_destroy(fsq_ctx *ctx)
{
	if (!ctx) return;
	cudaSetDevice(ctx->device);
	flush_phases(ctx);
	ctx->trim(false);
	ctx->trim(true);
	cudaStreamDestroy(ctx->stream);
	delete ctx;
}

const char *fsq_last_error(const fsq_ctx *ctx) { return ctx ? ctx->error.c_str() : g_create_error.c_str(); }

int fsq_set_option(fsq_ctx *ctx, const char *key, int64_t value)
{
	if (!ctx || !key) return fail(ctx, FSQ_ERR_ARG, "null argument");
	if (0 == strcmp("timer", key))
	{
		// device-side timing of a region of calls, on the stream the kernels are launched on
		CK(cudaSetDevice(ctx->device));
		int const which = value ? 0 : 1;
		if (!ctx->timer_ev[which]) CK(cudaEventCreate(&ctx->timer_ev[which]));
		CK(cudaEventRecord(ctx->timer_ev[which], ctx->stream));
		if (1 == which && ctx->timer_ev[0])
		{
			float ms = 0;
			CK(cudaEventSynchronize(ctx->timer_ev[1]));
			CK(cudaEventElapsedTime(&ms, ctx->timer_ev[0], ctx->timer_ev[1]));
			ctx->stats["ms_timer"] = ms;
		}
		return FSQ_OK;
	}
	static char const *known[] = {"block_columns", "cap", "keep_dp", "warps", "kernel", "chunk", "stitch_group", "stitch_sort", "dp_chunks", "dp_max_rounds", "dp_guess_rounds", "store_states", "whole_waves", nullptr};
	for (char const **k = known; *k; ++k)
		if (0 == strcmp(*k, key)) { ctx->opts[key] = value; return FSQ_OK; }
	return fail(ctx, FSQ_ERR_ARG, "unknown option '%s'", key);
}

int fsq_get_stat(const fsq_ctx *cctx, const char *key, double *value)
{
	fsq_ctx *ctx = const_cast <fsq_ctx *>(cctx);
	if (!ctx || !key || !value) return fail(ctx, FSQ_ERR_ARG, "null argument");
	cudaSetDevice(ctx->device);
	flush_phases(ctx);
	if (0 == strcmp(key, "reset")) { ctx->stats.clear(); *value = 0; return FSQ_OK; }
	auto it = ctx->stats.find(key);
	*value = ctx->stats.end() == it ? 0.0 : it->second;
	return FSQ_OK;
}

// ---- alignment ---------------------------------------------------------------------------------------------------

static int comp_nsym(fsq_ctx *ctx, uint8_t const comp[256], uint32_t *nsym)
{
	uint32_t mx = 0;
	for (int c = 0; c < 256; ++c) mx = std::max <uint32_t>(mx, comp[c]);
	if (mx >= FSQ_MAX_SYMS)
		return fail(ctx, FSQ_ERR_ARG, "alphabet has %u comps; at most 15 distinct characters (plus NUL) fit the 4-bit packing", mx + 1);
	*nsym = mx + 1;
	return FSQ_OK;
}

int fsq_msa_from_device_rows(fsq_ctx *ctx, const uint8_t *d_rows, uint64_t row_stride, uint32_t m, uint64_t n, const uint8_t comp[256], fsq_msa **out)
{
	if (!ctx || !d_rows || !comp || !out) return fail(ctx, FSQ_ERR_ARG, "null argument");
	CK(cudaSetDevice(ctx->device));
	uint32_t nsym = 0;
	int rc = comp_nsym(ctx, comp, &nsym);
	if (rc) return rc;
	fsq_msa *msa = nullptr;
	if ((rc = new_msa(ctx, m, n, 0, nsym, &msa))) return rc;
	dbuf d_comp, d_mask;
	cudaError_t e = d_comp.alloc(ctx, 256);
	if (cudaSuccess == e) e = d_mask.alloc(ctx, 4);
	if (cudaSuccess == e) e = cudaMemsetAsync(d_mask.p, 0, 4, ctx->stream);
	if (cudaSuccess == e) e = cudaMemcpyAsync(d_comp.p, comp, 256, cudaMemcpyHostToDevice, ctx->stream);
	if (cudaSuccess == e)
	{
		phase pk(ctx, "pack");
		e = launch_pack(d_rows, row_stride, m, n, d_comp.as <uint8_t>(), msa->cols, msa->col_stride, 0, d_mask.as <uint32_t>(), ctx->stream);
		pk.stop();
		ctx->stats["kernel_launches"] += 1;
	}
	if (cudaSuccess == e) e = cudaMemcpyAsync(&msa->sym_mask, d_mask.p, 4, cudaMemcpyDeviceToHost, ctx->stream);
	if (cudaSuccess == e) e = cudaStreamSynchronize(ctx->stream);
	if (cudaSuccess != e) { fsq_msa_free(msa); return fail(ctx, FSQ_ERR_CUDA, "pack: %s", cudaGetErrorString(e)); }
	*out = msa;
	return FSQ_OK;
}

int fsq_msa_from_rows(fsq_ctx *ctx, const uint8_t *const *rows, uint32_t m, uint64_t n, const uint8_t comp_in[256], fsq_msa **out)
{
	if (!ctx || !rows || !out) return fail(ctx, FSQ_ERR_ARG, "null argument");
	if (0 == m || 0 == n) return fail(ctx, FSQ_ERR_ARG, "empty alignment (m = %u, n = %llu)", m, (unsigned long long) n);
	CK(cudaSetDevice(ctx->device));
	uint8_t comp[256];
	if (comp_in)
		memcpy(comp, comp_in, 256);
	else
	{
		// lib/libbio/include/libbio/consecutive_alphabet.hh:259-264, 309-326: NUL = 0, bytes present get 1.. in byte order
		bool seen[256] = {false};
		for (uint32_t r = 0; r < m; ++r)
			for (uint64_t c = 0; c < n; ++c) seen[rows[r][c]] = true;
		seen[0] = true;
		uint32_t next = 0;
		for (int c = 0; c < 256; ++c) comp[c] = seen[c] ? (uint8_t) std::min <uint32_t>(255, next++) : 0;
	}
	uint32_t nsym = 0;
	int rc = comp_nsym(ctx, comp, &nsym);
	if (rc) return rc;
	fsq_msa *msa = nullptr;
	if ((rc = new_msa(ctx, m, n, 0, nsym, &msa))) return rc;

	// Rows go up in column chunks through a device staging matrix (row-major, m x chunk), then get packed.
	uint64_t const chunk = std::max <uint64_t>(32, std::min <uint64_t>(n, ((256ull << 20) / m) & ~31ull));
	uint64_t const row_stride = (chunk + 15) & ~15ull;
	dbuf stage, d_comp, d_mask;
	cudaError_t e = stage.alloc(ctx, (size_t) m * row_stride);
	if (cudaSuccess == e) e = d_comp.alloc(ctx, 256);
	if (cudaSuccess == e) e = d_mask.alloc(ctx, 4);
	if (cudaSuccess == e) e = cudaMemsetAsync(d_mask.p, 0, 4, ctx->stream);
	if (cudaSuccess == e) e = cudaMemcpyAsync(d_comp.p, comp, 256, cudaMemcpyHostToDevice, ctx->stream);
	phase pk(ctx, "h2d_pack");
	for (uint64_t c0 = 0; c0 < n && cudaSuccess == e; c0 += chunk)
	{
		uint64_t const nc = std::min(chunk, n - c0);
		// one strided copy when the rows are equally spaced in host memory, else one copy per row
		bool strided = m > 1;
		ptrdiff_t const pitch = m > 1 ? rows[1] - rows[0] : 0;
		for (uint32_t r = 1; r + 1 < m && strided; ++r) strided = (rows[r + 1] - rows[r]) == pitch;
		if (strided && pitch >= (ptrdiff_t) n)
			e = cudaMemcpy2DAsync(stage.p, row_stride, rows[0] + c0, (size_t) pitch, nc, m, cudaMemcpyHostToDevice, ctx->stream);
		else
			for (uint32_t r = 0; r < m && cudaSuccess == e; ++r)
				e = cudaMemcpyAsync(stage.as <uint8_t>() + (size_t) r * row_stride, rows[r] + c0, nc, cudaMemcpyHostToDevice, ctx->stream);
		if (cudaSuccess == e) e = launch_pack(stage.as <uint8_t>(), row_stride, m, nc, d_comp.as <uint8_t>(), msa->cols, msa->col_stride, c0, d_mask.as <uint32_t>(), ctx->stream);
		ctx->stats["kernel_launches"] += 1;
	}
	pk.stop();
	if (cudaSuccess == e) e = cudaMemcpyAsync(&msa->sym_mask, d_mask.p, 4, cudaMemcpyDeviceToHost, ctx->stream);
	if (cudaSuccess == e) e = cudaStreamSynchronize(ctx->stream);
	if (cudaSuccess != e) { fsq_msa_free(msa); return fail(ctx, FSQ_ERR_CUDA, "upload/pack: %s", cudaGetErrorString(e)); }
	*out = msa;
	return FSQ_OK;
}

int fsq_msa_synth(fsq_ctx *ctx, uint32_t m, uint64_t n, uint64_t col0, uint64_t seed, uint32_t founders, int gaps, fsq_msa **out)
{
	if (!ctx || !out || 0 == founders) return fail(ctx, FSQ_ERR_ARG, "null argument or zero founders");
	CK(cudaSetDevice(ctx->device));
	fsq_msa *msa = nullptr;
	int rc = new_msa(ctx, m, n, col0, gaps ? 6 : 5, &msa);
	if (rc) return rc;
	cudaError_t e = launch_synth_packed(msa->cols, msa->col_stride, m, n, col0, seed, founders, gaps, ctx->stream);
	if (cudaSuccess == e) e = cudaStreamSynchronize(ctx->stream);
	if (cudaSuccess != e) { fsq_msa_free(msa); return fail(ctx, FSQ_ERR_CUDA, "synth: %s", cudaGetErrorString(e)); }
	ctx->stats["kernel_launches"] += 1;
	msa->sym_mask = gaps ? 0x3eu : 0x1eu;      // comps 1..5 ('-', A, C, G, T) or 1..4
	*out = msa;
	return FSQ_OK;
}

int fsq_synth_rows_device(fsq_ctx *ctx, uint8_t *d_rows, uint64_t row_stride, uint32_t m, uint64_t n, uint64_t col0, uint64_t seed, uint32_t founders, int gaps)
{
	if (!ctx || !d_rows || 0 == founders || 0 == m || 0 == n) return fail(ctx, FSQ_ERR_ARG, "bad argument");
	CK(cudaSetDevice(ctx->device));
	CK(launch_synth_rows(d_rows, row_stride, m, n, col0, seed, founders, gaps, ctx->stream));
	CK(cudaStreamSynchronize(ctx->stream));
	return FSQ_OK;
}

void fsq_msa_free(fsq_msa *msa)
{
	if (!msa) return;
	if (msa->ctx) cudaSetDevice(msa->ctx->device);
	if (msa->cols_bytes && msa->ctx)
		msa->ctx->give(false, msa->cols, msa->cols_bytes);      // work that reads it is ordered before any reuse: one stream
	else
		cudaFree(msa->cols);
	delete msa;
}

int fsq_msa_set_first_column(fsq_msa *msa, uint64_t first_column)
{
	if (!msa) return FSQ_ERR_ARG;
	msa->col0 = first_column;
	return FSQ_OK;
}

uint32_t fsq_msa_rows(const fsq_msa *msa) { return msa ? msa->m : 0; }
uint64_t fsq_msa_columns(const fsq_msa *msa) { return msa ? msa->n : 0; }

// ---- PBWT states ---------------------------------------------------------------------------------------------------

int fsq_pbwt_states(fsq_ctx *ctx, const fsq_msa *msa, const uint64_t *columns, uint32_t count, uint32_t *a, uint32_t *d)
{
	int rc = check_shape(ctx, msa);
	if (rc) return rc;
	if (!columns || !a || !d) return fail(ctx, FSQ_ERR_ARG, "null argument");
	if (0 == count) return FSQ_OK;
	CK(cudaSetDevice(ctx->device));
	std::vector <uint64_t> cols(columns, columns + count);
	for (uint32_t i = 0; i < count; ++i)
		if (cols[i] < 1 || cols[i] > msa->n || (i && cols[i] <= cols[i - 1]))
			return fail(ctx, FSQ_ERR_ARG, "state columns must be strictly ascending and within [1, n]");
	scan_state S;
	if ((rc = prepare_scan(ctx, msa, S))) return rc;
	dbuf da, dd;
	size_t const bytes = (size_t) count * msa->m * 4;
	CK(da.alloc(ctx, bytes)); CK(dd.alloc(ctx, bytes));
	if ((rc = dump_states(ctx, msa, S, cols, da.as <uint32_t>(), dd.as <uint32_t>()))) return rc;
	CK(cudaMemcpyAsync(a, da.p, bytes, cudaMemcpyDeviceToHost, ctx->stream));
	CK(cudaMemcpyAsync(d, dd.p, bytes, cudaMemcpyDeviceToHost, ctx->stream));
	CK(cudaStreamSynchronize(ctx->stream));
	return FSQ_OK;
}

// ---- long path -----------------------------------------------------------------------------------------------------

int fsq_segment(fsq_ctx *ctx, const fsq_msa *msa, uint64_t L, fsq_segmentation **out)
{
	int rc = check_shape(ctx, msa);
	if (rc) return rc;
	if (!out) return fail(ctx, FSQ_ERR_ARG, "null out pointer");
	if (0 == L) return fail(ctx, FSQ_ERR_ARG, "segment length bound must be positive");
	if (0 != msa->col0) return fail(ctx, FSQ_ERR_ARG, "fsq_segment needs an alignment that starts at column 0");
	uint64_t const n = msa->n;
	uint32_t const m = msa->m;
	if (n < 2 * L) return fail(ctx, FSQ_ERR_SHORT, "n = %llu < 2L = %llu: short path", (unsigned long long) n, (unsigned long long) (2 * L));
	CK(cudaSetDevice(ctx->device));
	flush_phases(ctx);
	if (!ctx->call_start) CK(cudaEventCreate(&ctx->call_start));
	CK(cudaEventRecord(ctx->call_start, ctx->stream));

	fsq_shard run;
	fsq_shard *P = &run;
	if ((rc = shard_init(P, ctx, msa, n, L))) return rc;
	if ((rc = prepare_scan(ctx, msa, P->S))) return rc;
	geometry const &g = P->S.g;

	// ---- pass B + DP; the top-list bound grows until the DP proves the result exact ----
	// A round costs one chain's length of sequential work, the single-chain fallback as much as n_chunks rounds (and round r
	// leaves the first r chains exact, so n_chunks rounds always suffice): give up on the rounds only after half of that.
	// Slowly rising levels (thousands of distinct rows per segment) take 15-20 rounds where config 2 takes two.
	int64_t max_rounds = ctx->opt("dp_max_rounds", 0);
	if (max_rounds <= 0) max_rounds = std::max <int64_t>(16, P->n_chunks / 2);
	max_rounds = std::max <int64_t>(2, max_rounds);
	int64_t guess_rounds = ctx->opt("dp_guess_rounds", -1);           // after these the table is left to settle by itself
	if (guess_rounds < 0) guess_rounds = max_rounds;
	while (true)
	{
		if ((rc = shard_scan(P))) return rc;
		while (true)
		{
			uint32_t changed = 0, reach = 0;
			if (0 == P->rounds && P->n_chunks > 1 && !P->single_chain && max_rounds > 2 && guess_rounds >= 1)
			{
				if ((rc = shard_dp_first_two_rounds(P, &changed, &reach))) return rc;
			}
			else if ((rc = shard_dp_round(P, &changed, &reach))) return rc;
			if (1 == P->n_chunks || P->single_chain) break;      // a single chain is the reference's own order: exact as it stands
			if (P->rounds > 1 && 0 == changed) break;            // nothing changed: every entry was computed from the table it is part of
			if (P->rounds >= max_rounds)
			{
				// the guesses keep failing: one chain over everything, exact in a single round
				P->single_chain = true;
				ctx->stats["dp_fallbacks"] += 1;
			}
			else if (P->rounds <= guess_rounds)
			{
				if ((rc = shard_dp_sync(P, 1, reach))) return rc;
			}
		}
		uint32_t ovf = 0;
		if (P->final_done) ovf = P->final_overflow;
		else if ((rc = shard_dp_final(P, &ovf))) return rc;
		if (0 == ovf) break;
		if (P->cap >= m) return fail(ctx, FSQ_ERR_INTERNAL, "DP reported an incomplete top list at column %u with cap = m", ovf - 1);
		uint64_t next = std::min <uint64_t>(m, (uint64_t) P->cap * 4);
		{
			// four times the refused bound unless the lists of that length do not fit: then twice (config 3 whole on one
			// GPU: 127 -> 508 -> 2,032 would be 163 GB of lists next to 25 GB of alignment; 1,016 is enough and fits)
			size_t free_b = 0, total_b = 0;
			CK(cudaMemGetInfo(&free_b, &total_b));
			size_t avail = free_b + P->toplist.bytes;
			for (auto const &blk : ctx->dev_cache) avail += blk.bytes;
			while (next / 2 > P->cap && (size_t) n * (next + 1) * sizeof(uint2) + ((size_t) 1 << 30) > avail) next /= 2;
		}
		P->cap = (uint32_t) next;
		ctx->opts["cap"] = P->cap;          // later calls on this context start from the bound that worked
		ctx->stats["cap_retries"] += 1;
	}
	ctx->stats["cap"] = P->cap;
	ctx->stats["block_columns"] = g.C;
	ctx->stats["blocks"] = g.B;
	ctx->stats["warps"] = (g.blocked || g.chunked) ? g.T / 32 : g.n_warps;
	ctx->stats["chunk"] = (g.blocked || g.chunked) ? g.E : 0;
	ctx->stats["kernel"] = g.chunked ? 3 : (g.blocked ? 2 : 1);

	// ---- traceback (reference: segmentation_lp_context.cc:191-224) ----
	fsq_segmentation *seg = new (std::nothrow) fsq_segmentation;
	if (!seg) return fail(ctx, FSQ_ERR_NOMEM, "out of host memory");
	struct guard { fsq_segmentation *&s; bool ok{false}; ~guard() { if (!ok) { fsq_segmentation_free(s); s = nullptr; } } } gd{seg};
	seg->ctx = ctx; seg->m = m; seg->n = n; seg->L = L;
	if ((rc = shard_traceback(P))) return rc;
	seg->T = P->T; seg->h_tb_bytes = P->h_tb.bytes; seg->h_tb = P->h_tb.detach();
	seg->K = P->K;
	if (ctx->opt("keep_dp", 0))
	{
		uint64_t const dp_size = P->dp_size;
		seg->dp_lb.resize(dp_size); seg->dp_max.resize(dp_size); seg->dp_size.resize(dp_size);
		CK(cudaMemcpyAsync(seg->dp_lb.data(), P->tb_lb.p, dp_size * 4, cudaMemcpyDeviceToHost, ctx->stream));
		CK(cudaMemcpyAsync(seg->dp_max.data(), P->tb_max.p, dp_size * 4, cudaMemcpyDeviceToHost, ctx->stream));
		CK(cudaMemcpyAsync(seg->dp_size.data(), P->tb_size.p, dp_size * 4, cudaMemcpyDeviceToHost, ctx->stream));
		CK(cudaStreamSynchronize(ctx->stream));
	}
	if (!(seg->K < m))
	{
		// generate_context.cc:192-200; the traceback stays available through the handle for inspection
		gd.ok = true;
		*out = seg;
		return fail(ctx, FSQ_ERR_NOT_REDUCIBLE, "the maximum segment size %u is equal to the number of input sequences", seg->K);
	}

	// ---- greedy reduction (segmentation_lp_context.cc:335-390), then the states at the reduced right bounds
	//      (segmentation_lp_context.cc:229-332, update_pbwt_task.cc:13-35) ----
	if ((rc = shard_reduce_lists(P))) return rc;
	if ((rc = shard_reduce(P, true))) return rc;
	if ((rc = shard_reduced_states(P))) return rc;
	seg->lb = P->red_lb; seg->rb = P->red_rb; seg->size = P->red_size;
	seg->d_bytes[0] = P->red_a.bytes; seg->d_bytes[1] = P->red_d.bytes;
	seg->d_a = (uint32_t *) P->red_a.detach(); seg->d_d = (uint32_t *) P->red_d.detach();
	CK(cudaStreamSynchronize(ctx->stream));
	gd.ok = true;
	*out = seg;
	return FSQ_OK;
}

void fsq_segmentation_free(fsq_segmentation *seg)
{
	if (!seg) return;
	// the buffers go back to the context's workspace cache (the context must outlive its segmentations)
	fsq_ctx *ctx = seg->ctx;
	ctx->give(false, seg->d_a, seg->d_bytes[0]);
	ctx->give(false, seg->d_d, seg->d_bytes[1]);
	ctx->give(true, seg->h_tb, seg->h_tb_bytes);
	ctx->give(true, seg->h_a, seg->h_bytes[0]);
	ctx->give(true, seg->h_d, seg->h_bytes[1]);
	delete seg;
}

uint32_t fsq_segmentation_count(const fsq_segmentation *seg) { return seg ? (uint32_t) seg->lb.size() : 0; }
uint32_t fsq_segmentation_max_segment_size(const fsq_segmentation *seg) { return seg ? seg->K : 0; }

int fsq_segmentation_bounds(const fsq_segmentation *seg, uint64_t *lb, uint64_t *rb, uint32_t *size)
{
	if (!seg || !lb || !rb || !size) return FSQ_ERR_ARG;
	std::copy(seg->lb.begin(), seg->lb.end(), lb);
	std::copy(seg->rb.begin(), seg->rb.end(), rb);
	std::copy(seg->size.begin(), seg->size.end(), size);
	return FSQ_OK;
}

static int fetch_states(fsq_segmentation *seg)
{
	fsq_ctx *ctx = seg->ctx;
	if (seg->h_a) return FSQ_OK;
	size_t const bytes = seg->lb.size() * (size_t) seg->m * 4;
	CK(cudaSetDevice(ctx->device));
	phase pd(ctx, "states_d2h");
	{
		hbuf ha, hd;
		CK(ha.alloc(ctx, bytes)); CK(hd.alloc(ctx, bytes));
		seg->h_bytes[0] = ha.bytes; seg->h_bytes[1] = hd.bytes;
		seg->h_a = (uint32_t *) ha.detach(); seg->h_d = (uint32_t *) hd.detach();
	}
	CK(cudaMemcpyAsync(seg->h_a, seg->d_a, bytes, cudaMemcpyDeviceToHost, ctx->stream));
	CK(cudaMemcpyAsync(seg->h_d, seg->d_d, bytes, cudaMemcpyDeviceToHost, ctx->stream));
	pd.stop();
	CK(cudaStreamSynchronize(ctx->stream));
	return FSQ_OK;
}

int fsq_segmentation_states(const fsq_segmentation *cseg, uint32_t *a, uint32_t *d)
{
	fsq_segmentation *seg = const_cast <fsq_segmentation *>(cseg);
	if (!seg || !a || !d) return FSQ_ERR_ARG;
	int const rc = fetch_states(seg);
	if (rc) return rc;
	size_t const bytes = seg->lb.size() * (size_t) seg->m * 4;
	memcpy(a, seg->h_a, bytes);
	memcpy(d, seg->h_d, bytes);
	return FSQ_OK;
}

int fsq_segmentation_states_view(const fsq_segmentation *cseg, const uint32_t **a, const uint32_t **d)
{
	fsq_segmentation *seg = const_cast <fsq_segmentation *>(cseg);
	if (!seg || !a || !d) return FSQ_ERR_ARG;
	int const rc = fetch_states(seg);
	if (rc) return rc;
	*a = seg->h_a;
	*d = seg->h_d;
	return FSQ_OK;
}

uint32_t fsq_segmentation_traceback_count(const fsq_segmentation *seg) { return seg ? seg->T : 0; }

// rb u64[T], then lb, max, size u32[T] each (shard_traceback)
static void copy_traceback(void const *h_tb, uint32_t T, uint64_t *lb, uint64_t *rb, uint32_t *max_size, uint32_t *size)
{
	uint64_t const *h_rb = static_cast <uint64_t const *>(h_tb);
	uint32_t const *h_lb = reinterpret_cast <uint32_t const *>(h_rb + T), *h_max = h_lb + T, *h_size = h_max + T;
	std::copy(h_lb, h_lb + T, lb);
	std::copy(h_rb, h_rb + T, rb);
	std::copy(h_max, h_max + T, max_size);
	std::copy(h_size, h_size + T, size);
}

int fsq_segmentation_traceback(const fsq_segmentation *seg, uint64_t *lb, uint64_t *rb, uint32_t *max_size, uint32_t *size)
{
	if (!seg || !lb || !rb || !max_size || !size) return FSQ_ERR_ARG;
	if (seg->h_tb) copy_traceback(seg->h_tb, seg->T, lb, rb, max_size, size);
	return FSQ_OK;
}

int fsq_segmentation_dp_table(const fsq_segmentation *seg, uint64_t *lb, uint64_t *rb, uint32_t *max_size, uint32_t *size)
{
	if (!seg || !lb || !rb || !max_size || !size) return FSQ_ERR_ARG;
	if (seg->dp_lb.empty())
		return fail(seg->ctx, FSQ_ERR_ARG, "the DP table was not kept; set option keep_dp = 1 before fsq_segment");
	for (size_t t = 0; t < seg->dp_lb.size(); ++t)
	{
		bool const written = 0xffffffffu != seg->dp_max[t];
		lb[t] = seg->dp_lb[t];
		rb[t] = written ? t + seg->L : 0;
		max_size[t] = seg->dp_max[t];
		size[t] = seg->dp_size[t];
	}
	return FSQ_OK;
}

// ---- short path ----------------------------------------------------------------------------------------------------

static int sets_on_device(fsq_ctx *ctx, uint32_t const *d_a, uint32_t const *d_d, uint32_t m, uint64_t lb, dbuf &off, dbuf &ids, dbuf &cls, dbuf &nruns)
{
	CK(off.alloc(ctx, ((size_t) m + 1) * 4)); CK(ids.alloc(ctx, (size_t) m * 4)); CK(cls.alloc(ctx, (size_t) m * 4)); CK(nruns.alloc(ctx, 4));
	CK(launch_segment_sets(d_a, d_d, m, lb, off.as <uint32_t>(), ids.as <uint32_t>(), cls.as <uint32_t>(), nruns.as <uint32_t>(), ctx->stream));
	ctx->stats["kernel_launches"] += 1;
	return FSQ_OK;
}

int fsq_distinct_rows(fsq_ctx *ctx, const fsq_msa *msa, uint32_t *row_of_run, uint32_t *run_len, uint32_t *count)
{
	int rc = check_shape(ctx, msa);
	if (rc) return rc;
	if (!row_of_run || !run_len || !count) return fail(ctx, FSQ_ERR_ARG, "null argument");
	CK(cudaSetDevice(ctx->device));
	uint32_t const m = msa->m;
	scan_state S;
	if ((rc = prepare_scan(ctx, msa, S))) return rc;
	dbuf da, dd, off, ids, cls, nruns;
	CK(da.alloc(ctx, (size_t) m * 4)); CK(dd.alloc(ctx, (size_t) m * 4));
	std::vector <uint64_t> cols(1, msa->n);
	if ((rc = dump_states(ctx, msa, S, cols, da.as <uint32_t>(), dd.as <uint32_t>()))) return rc;
	if ((rc = sets_on_device(ctx, da.as <uint32_t>(), dd.as <uint32_t>(), m, 0, off, ids, cls, nruns))) return rc;
	std::vector <uint32_t> h_off(m + 1), h_a(m);
	uint32_t runs = 0;
	CK(cudaMemcpyAsync(h_off.data(), off.p, ((size_t) m + 1) * 4, cudaMemcpyDeviceToHost, ctx->stream));
	CK(cudaMemcpyAsync(h_a.data(), da.p, (size_t) m * 4, cudaMemcpyDeviceToHost, ctx->stream));
	CK(cudaMemcpyAsync(&runs, nruns.p, 4, cudaMemcpyDeviceToHost, ctx->stream));
	CK(cudaStreamSynchronize(ctx->stream));
	for (uint32_t r = 0; r < runs; ++r)
	{
		row_of_run[r] = h_a[h_off[r]];
		run_len[r] = h_off[r + 1] - h_off[r];
	}
	*count = runs;
	return FSQ_OK;
}

// ---- segment sets, edge weights ---------------------------------------------------------------------------------------

int fsq_segment_sets(fsq_ctx *ctx, const fsq_segmentation *seg, uint32_t s, uint32_t *run_off, uint32_t *ids_out, uint32_t *class_of_row, uint32_t *n_runs)
{
	if (!ctx || !seg || !run_off || !ids_out || !class_of_row || !n_runs) return fail(ctx, FSQ_ERR_ARG, "null argument");
	if (s >= seg->lb.size()) return fail(ctx, FSQ_ERR_ARG, "segment %u out of range", s);
	CK(cudaSetDevice(ctx->device));
	uint32_t const m = seg->m;
	dbuf off, ids, cls, nruns;
	int rc = sets_on_device(ctx, seg->d_a + (size_t) s * m, seg->d_d + (size_t) s * m, m, seg->lb[s], off, ids, cls, nruns);
	if (rc) return rc;
	uint32_t runs = 0;
	CK(cudaMemcpyAsync(&runs, nruns.p, 4, cudaMemcpyDeviceToHost, ctx->stream));
	CK(cudaMemcpyAsync(ids_out, ids.p, (size_t) m * 4, cudaMemcpyDeviceToHost, ctx->stream));
	CK(cudaMemcpyAsync(class_of_row, cls.p, (size_t) m * 4, cudaMemcpyDeviceToHost, ctx->stream));
	CK(cudaStreamSynchronize(ctx->stream));
	CK(cudaMemcpyAsync(run_off, off.p, ((size_t) runs + 1) * 4, cudaMemcpyDeviceToHost, ctx->stream));
	CK(cudaStreamSynchronize(ctx->stream));
	*n_runs = runs;
	return FSQ_OK;
}

int fsq_pair_weights(fsq_ctx *ctx, const fsq_segmentation *seg, uint32_t s, int32_t *W, uint32_t *runs_lhs, uint32_t *runs_rhs)
{
	if (!ctx || !seg || !W || !runs_lhs || !runs_rhs) return fail(ctx, FSQ_ERR_ARG, "null argument");
	if ((size_t) s + 1 >= seg->lb.size()) return fail(ctx, FSQ_ERR_ARG, "segment pair %u out of range", s);
	CK(cudaSetDevice(ctx->device));
	uint32_t const m = seg->m;
	dbuf off_l, ids_l, cls_l, n_l, off_r, ids_r, cls_r, n_r, dW;
	int rc;
	if ((rc = sets_on_device(ctx, seg->d_a + (size_t) s * m, seg->d_d + (size_t) s * m, m, seg->lb[s], off_l, ids_l, cls_l, n_l))) return rc;
	if ((rc = sets_on_device(ctx, seg->d_a + (size_t) (s + 1) * m, seg->d_d + (size_t) (s + 1) * m, m, seg->lb[s + 1], off_r, ids_r, cls_r, n_r))) return rc;
	uint32_t rl = 0, rr = 0;
	CK(cudaMemcpyAsync(&rl, n_l.p, 4, cudaMemcpyDeviceToHost, ctx->stream));
	CK(cudaMemcpyAsync(&rr, n_r.p, 4, cudaMemcpyDeviceToHost, ctx->stream));
	CK(cudaStreamSynchronize(ctx->stream));
	size_t const bytes = (size_t) rl * rr * 4;
	CK(dW.alloc(ctx, bytes));
	CK(cudaMemsetAsync(dW.p, 0, bytes, ctx->stream));
	CK(launch_pair_weights(cls_l.as <uint32_t>(), cls_r.as <uint32_t>(), m, rr, dW.as <int32_t>(), ctx->stream));
	ctx->stats["kernel_launches"] += 1;
	CK(cudaMemcpyAsync(W, dW.p, bytes, cudaMemcpyDeviceToHost, ctx->stream));
	CK(cudaStreamSynchronize(ctx->stream));
	*runs_lhs = rl;
	*runs_rhs = rr;
	return FSQ_OK;
}

// ---- column-sharded operation -------------------------------------------------------------------------------------------
// One process per GPU; rank r holds the columns [c0_r, c1_r) of one alignment.  The stages below are the long path of
// fsq_segment() cut where ranks exchange data; the exchange itself (NCCL through torch.distributed, see
// founder-sequences_b200/sharded.py) is the caller's.

int fsq_shard_create(fsq_ctx *ctx, const fsq_msa *msa, uint64_t n_total, uint64_t L, fsq_shard **out)
{
	int rc = check_shape(ctx, msa);
	if (rc) return rc;
	if (!out) return fail(ctx, FSQ_ERR_ARG, "null out pointer");
	if (0 == L) return fail(ctx, FSQ_ERR_ARG, "segment length bound must be positive");
	if (msa->col0 + msa->n > n_total) return fail(ctx, FSQ_ERR_ARG, "the shard's columns end beyond the alignment");
	if (n_total >= (1ull << 31)) return fail(ctx, FSQ_ERR_UNSUPPORTED, "n = %llu columns: positions are kept in 32 bits (n < 2^31)", (unsigned long long) n_total);
	if (n_total < 2 * L) return fail(ctx, FSQ_ERR_SHORT, "n = %llu < 2L = %llu: short path", (unsigned long long) n_total, (unsigned long long) (2 * L));
	CK(cudaSetDevice(ctx->device));
	fsq_shard *P = new (std::nothrow) fsq_shard;
	if (!P) return fail(ctx, FSQ_ERR_NOMEM, "out of host memory");
	if ((rc = shard_init(P, ctx, msa, n_total, L))) { delete P; return rc; }
	*out = P;
	return FSQ_OK;
}

void fsq_shard_free(fsq_shard *P)
{
	if (!P) return;
	cudaSetDevice(P->ctx->device);
	cudaStreamSynchronize(P->ctx->stream);
	delete P;
}

int fsq_shard_local_summary(fsq_shard *P, const uint32_t **d_pi, const uint32_t **d_delta)
{
	if (!P || !d_pi || !d_delta) return FSQ_ERR_ARG;
	fsq_ctx *ctx = P->ctx;
	CK(cudaSetDevice(ctx->device));
	flush_phases(ctx);
	if (!ctx->call_start) CK(cudaEventCreate(&ctx->call_start));
	CK(cudaEventRecord(ctx->call_start, ctx->stream));
	int const rc = local_scan(ctx, P->msa, P->S);
	if (rc) return rc;
	CK(cudaStreamSynchronize(ctx->stream));
	// the state at the end of the last block, relative to the identity state at c0: this rank's summary
	*d_pi = P->S.st_a.as <uint32_t>() + (size_t) P->S.g.B * P->m;
	*d_delta = P->S.st_d.as <uint32_t>() + (size_t) P->S.g.B * P->m;
	return FSQ_OK;
}

int fsq_shard_set_start(fsq_shard *P, const uint32_t *d_pis, const uint32_t *d_deltas, const uint64_t *first_columns, uint32_t count)
{
	if (!P) return FSQ_ERR_ARG;
	fsq_ctx *ctx = P->ctx;
	CK(cudaSetDevice(ctx->device));
	if (!P->S.sum_a.p) return fail(ctx, FSQ_ERR_ARG, "fsq_shard_local_summary has not run");
	uint32_t const m = P->m;
	if (0 == count)
	{
		if (0 != P->c0) return fail(ctx, FSQ_ERR_ARG, "a shard that does not start at column 0 needs the summaries of the ranks before it");
	}
	else
	{
		if (!d_pis || !d_deltas || !first_columns) return fail(ctx, FSQ_ERR_ARG, "null argument");
		if (0 != first_columns[0]) return fail(ctx, FSQ_ERR_ARG, "the first summary must start at column 0");
		// true state at c0 = summary_0 o summary_1 o ... (summary_0 starts from the true initial state); one CTA, in order
		dbuf tmp_a, tmp_d;
		CK(tmp_a.alloc(ctx, (size_t) 2 * m * 4)); CK(tmp_d.alloc(ctx, (size_t) 2 * m * 4));
		uint32_t const *cur_a = d_pis, *cur_d = d_deltas;
		if (count > 1)
		{
			stitch_plan plan;
			for (uint32_t i = 1; i < count; ++i)
			{
				stitch_item it;
				memset(&it, 0, sizeof it);
				it.in_a = cur_a; it.in_d = cur_d;
				it.pi = d_pis + (size_t) i * m; it.delta = d_deltas + (size_t) i * m;
				it.out_a = tmp_a.as <uint32_t>() + (size_t) (i & 1) * m; it.out_d = tmp_d.as <uint32_t>() + (size_t) (i & 1) * m;
				it.s_mark = (uint32_t) first_columns[i];
				plan.items.push_back(it);
				plan.end_level();              // one launch per step: each reads what the previous one wrote
				cur_a = it.out_a; cur_d = it.out_d;
			}
			phase ps(ctx, "stitch");
			CK(cudaStreamSynchronize(ctx->stream));
			int const rc = run_stitch_plan(ctx, P->S, plan, m, {});
			if (rc) return rc;
			ps.stop();
			CK(cudaStreamSynchronize(ctx->stream));
		}
		int const rc = apply_start_state(ctx, P->msa, P->S, cur_a, cur_d);
		if (rc) return rc;
		CK(cudaStreamSynchronize(ctx->stream));
	}
	int const rc = compress_starts(ctx, P->msa, P->S);
	if (rc) return rc;
	P->started = true;
	return FSQ_OK;
}

int fsq_shard_scan(fsq_shard *P)
{
	if (!P) return FSQ_ERR_ARG;
	fsq_ctx *ctx = P->ctx;
	CK(cudaSetDevice(ctx->device));
	if (!P->started) return fail(ctx, FSQ_ERR_ARG, "fsq_shard_set_start has not run");
	int const rc = shard_scan(P);
	if (rc) return rc;
	CK(cudaStreamSynchronize(ctx->stream));
	return FSQ_OK;
}

int fsq_shard_dp_table(fsq_shard *P, uint32_t **d_max, uint32_t **d_lb, uint32_t **d_size, uint64_t *entries, uint64_t *own_first, uint64_t *own_last)
{
	if (!P || !d_max || !d_lb || !d_size || !entries || !own_first || !own_last) return FSQ_ERR_ARG;
	if (!P->tb_max.p) return fail(P->ctx, FSQ_ERR_ARG, "fsq_shard_scan has not run");
	*d_max = P->tb_max.as <uint32_t>(); *d_lb = P->tb_lb.as <uint32_t>(); *d_size = P->tb_size.as <uint32_t>();
	*entries = P->dp_size;
	*own_first = P->t_first;
	*own_last = (P->c1 == P->n) ? P->dp_size : P->t_last;       // the rank with the last column also owns the final entry
	return FSQ_OK;
}

int fsq_shard_dp_round(fsq_shard *P, uint32_t *changed, uint32_t *reach)
{
	if (!P || !changed || !reach) return FSQ_ERR_ARG;
	CK_(P->ctx, cudaSetDevice(P->ctx->device));
	return shard_dp_round(P, changed, reach);
}

int fsq_shard_dp_sync(fsq_shard *P, int guess, uint32_t reach)
{
	if (!P) return FSQ_ERR_ARG;
	CK_(P->ctx, cudaSetDevice(P->ctx->device));
	return shard_dp_sync(P, guess, reach);
}

int fsq_shard_dp_single_chain(fsq_shard *P)
{
	if (!P) return FSQ_ERR_ARG;
	P->single_chain = true;
	P->ctx->stats["dp_fallbacks"] += 1;
	return FSQ_OK;
}

int fsq_shard_dp_final(fsq_shard *P, uint32_t *overflow)
{
	if (!P || !overflow) return FSQ_ERR_ARG;
	CK_(P->ctx, cudaSetDevice(P->ctx->device));
	return shard_dp_final(P, overflow);
}

int fsq_shard_grow_cap(fsq_shard *P)
{
	if (!P) return FSQ_ERR_ARG;
	fsq_ctx *ctx = P->ctx;
	if (P->cap >= P->m) return fail(ctx, FSQ_ERR_INTERNAL, "DP reported an incomplete top list with cap = m");
	P->cap = (uint32_t) std::min <uint64_t>(P->m, (uint64_t) P->cap * 4);
	ctx->opts["cap"] = P->cap;
	ctx->stats["cap_retries"] += 1;
	return FSQ_OK;
}

int fsq_shard_traceback(fsq_shard *P, uint32_t *count, uint32_t *K)
{
	if (!P || !count || !K) return FSQ_ERR_ARG;
	CK_(P->ctx, cudaSetDevice(P->ctx->device));
	int const rc = shard_traceback(P);
	if (rc) return rc;
	*count = P->T;
	*K = P->K;
	return FSQ_OK;
}

int fsq_shard_get_traceback(const fsq_shard *P, uint64_t *lb, uint64_t *rb, uint32_t *max_size, uint32_t *size)
{
	if (!P || !lb || !rb || !max_size || !size) return FSQ_ERR_ARG;
	if (P->h_tb.p) copy_traceback(P->h_tb.p, P->T, lb, rb, max_size, size);
	return FSQ_OK;
}

int fsq_shard_position_range(const fsq_shard *P, uint64_t c0, uint64_t c1, uint32_t *first, uint32_t *last)
{
	if (!P || !first || !last) return FSQ_ERR_ARG;
	if (!P->h_tb.p) return fail(P->ctx, FSQ_ERR_ARG, "fsq_shard_traceback has not run");
	uint64_t const *h_rb = P->h_tb.as <uint64_t>();
	*first = (uint32_t) (std::upper_bound(h_rb, h_rb + P->T, c0) - h_rb);
	*last = (uint32_t) (std::upper_bound(h_rb, h_rb + P->T, c1) - h_rb);
	return FSQ_OK;
}

int fsq_shard_reduce_lists(fsq_shard *P, uint32_t **d_lists, uint32_t *width, uint32_t *own_first, uint32_t *own_last)
{
	if (!P || !d_lists || !width || !own_first || !own_last) return FSQ_ERR_ARG;
	fsq_ctx *ctx = P->ctx;
	CK(cudaSetDevice(ctx->device));
	if (!(P->K < P->m)) return fail(ctx, FSQ_ERR_NOT_REDUCIBLE, "the maximum segment size %u is equal to the number of input sequences", P->K);
	int const rc = shard_reduce_lists(P);
	if (rc) return rc;
	CK(cudaStreamSynchronize(ctx->stream));
	*d_lists = P->lists.as <uint32_t>();
	*width = P->K + 1;
	*own_first = P->pos_first;
	*own_last = P->pos_last;
	return FSQ_OK;
}

int fsq_shard_reduce(fsq_shard *P, uint32_t *count, uint32_t *own_first, uint32_t *own_last)
{
	if (!P || !count || !own_first || !own_last) return FSQ_ERR_ARG;
	fsq_ctx *ctx = P->ctx;
	CK(cudaSetDevice(ctx->device));
	int rc = shard_reduce(P, false);
	if (rc) return rc;
	if ((rc = shard_reduced_states(P))) return rc;
	CK(cudaStreamSynchronize(ctx->stream));
	*count = P->n_reduced;
	*own_first = P->seg_first;
	*own_last = P->seg_last;
	ctx->stats["cap"] = P->cap;
	ctx->stats["block_columns"] = P->S.g.C;
	ctx->stats["blocks"] = P->S.g.B;
	ctx->stats["warps"] = (P->S.g.blocked || P->S.g.chunked) ? P->S.g.T / 32 : P->S.g.n_warps;
	ctx->stats["chunk"] = (P->S.g.blocked || P->S.g.chunked) ? P->S.g.E : 0;
	ctx->stats["kernel"] = P->S.g.chunked ? 3 : (P->S.g.blocked ? 2 : 1);
	return FSQ_OK;
}

int fsq_shard_bounds(const fsq_shard *cP, uint64_t *lb, uint64_t *rb, uint32_t *size)
{
	fsq_shard *P = const_cast <fsq_shard *>(cP);
	if (!P || !lb || !rb || !size) return FSQ_ERR_ARG;
	CK_(P->ctx, cudaSetDevice(P->ctx->device));
	{
		int const rc = finish_reduced(P);
		if (rc) return rc;
	}
	std::copy(P->red_lb.begin(), P->red_lb.end(), lb);
	std::copy(P->red_rb.begin(), P->red_rb.end(), rb);
	std::copy(P->red_size.begin(), P->red_size.end(), size);
	return FSQ_OK;
}

int fsq_shard_states(fsq_shard *P, uint32_t *a, uint32_t *d)
{
	if (!P || !a || !d) return FSQ_ERR_ARG;
	fsq_ctx *ctx = P->ctx;
	CK(cudaSetDevice(ctx->device));
	size_t const bytes = (size_t) (P->seg_last - P->seg_first) * P->m * 4;
	if (0 == bytes) return FSQ_OK;
	CK(cudaMemcpyAsync(a, P->red_a.p, bytes, cudaMemcpyDeviceToHost, ctx->stream));
	CK(cudaMemcpyAsync(d, P->red_d.p, bytes, cudaMemcpyDeviceToHost, ctx->stream));
	CK(cudaStreamSynchronize(ctx->stream));
	return FSQ_OK;
}

} // extern "C"
