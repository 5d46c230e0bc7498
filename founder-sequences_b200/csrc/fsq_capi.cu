// C ABI of libfsq_b200.so (include/fsq.h): contexts, device memory, and the orchestration of the kernels.
// There is no CPU fallback anywhere in this file: every compute entry point launches the CUDA kernels or fails.
#include "../../include/fsq.h"
#include "fsq_internal.h"
#include "fsq_comm.h"
#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <chrono>
#include <map>
#include <memory>
#include <new>
#include <string>
#include <tuple>
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
	size_t call_first_pending{0};         // the phase timers from here on belong to the call whose start call_start marks
	std::vector <cudaEvent_t> event_pool;  // phase timers take their events here and give them back (no create / destroy per phase)
	cudaEvent_t take_event() { if (event_pool.empty()) { cudaEvent_t e{}; cudaEventCreate(&e); return e; } cudaEvent_t e = event_pool.back(); event_pool.pop_back(); return e; }

	// what calls on a shape (m, n, L) have learned: the top-list bound that proved exact.  The option "cap" stays the caller's.
	std::map <std::tuple <uint32_t, uint64_t, uint64_t>, uint32_t> learned_cap;

	// ... and the memory plan agreed for (m, n, L, bound, ranks, local columns, blocks, block length, option store_states):
	// {one window, states kept, blocks per window}
	struct mem_plan { bool whole, keep; uint32_t per_window; };
	std::map <std::tuple <uint32_t, uint64_t, uint64_t, uint32_t, int, uint64_t, uint32_t, uint32_t, int64_t>, mem_plan> plans;

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
	void adopt(wbuf &o) { release(); p = o.p; bytes = o.bytes; owner = o.owner; o.p = nullptr; }
	template <typename T> T *as() const { return static_cast <T *>(p); }
};
typedef wbuf <false> dbuf;
typedef wbuf <true> hbuf;


struct fsq_segmentation
{
	fsq_ctx *ctx{};
	uint32_t m{};
	uint64_t n{}, L{};
	uint32_t K{};
	uint32_t n_reduced{};
	std::vector <uint64_t> lb, rb;            // reduced (every rank holds all of them); on the host once asked for (or copied for free)
	std::vector <uint32_t> size;
	dbuf d_lb, d_rb, d_size;                  // ... on the device: n_reduced entries each
	bool bounds_on_host{false};
	hbuf h_bounds;                            // the bounds as they rode along with the control words: lb u64[stride], rb u64[stride], size u32[stride]
	uint32_t h_bounds_stride{0};
	uint32_t own_first{}, own_last{};         // the reduced segments whose state this rank holds (all of them on one GPU)
	dbuf red_a, red_d;                        // their states on the device, (own_last - own_first) x m
	hbuf h_a, h_d;                            // pinned host copies, made on first request
	dbuf tr_lb, tr_rb, tr_max, tr_size;       // traceback before reduction, on the device: T entries
	hbuf h_tb;                                // ... and on the host once asked for: rb u64[T], then lb, max, size u32[T] each
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

struct phase
{
	fsq_ctx *ctx; std::string name; cudaEvent_t e0{}, e1{};
	phase(fsq_ctx *c, char const *n): ctx(c), name(n)
	{
		e0 = ctx->take_event(); e1 = ctx->take_event();
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
		if (ctx->call_start && (size_t) (&p - ctx->pending.data()) >= ctx->call_first_pending)
		{
			float at = 0;
			if (cudaSuccess == cudaEventElapsedTime(&at, ctx->call_start, p.second.first)) ctx->stats["at_" + p.first] = at;
			if (cudaSuccess == cudaEventElapsedTime(&at, ctx->call_start, p.second.second)) ctx->stats["end_" + p.first] = at;
			if (ctx->opt("timeline", 0))
			{
				// every occurrence of every phase of the last call, in order: "tl_count", "tl_id_<i>" (index into the list below),
				// "tl_at_<i>", "tl_end_<i>" (ms from the call's start); tests/gpu_timeline_sharded.py prints them
				static char const *names[] = {"pass_a", "stitch", "prep", "pass_b", "dp", "exchange", "traceback", "reduce", "pass_c", nullptr};
				int id = 0;
				while (names[id] && p.first != names[id]) ++id;
				size_t const i = (size_t) (&p - ctx->pending.data()) - ctx->call_first_pending;
				char key[48];
				snprintf(key, sizeof key, "tl_id_%zu", i); ctx->stats[key] = id;
				snprintf(key, sizeof key, "tl_end_%zu", i); ctx->stats[key] = at;
				if (cudaSuccess == cudaEventElapsedTime(&at, ctx->call_start, p.second.first)) { snprintf(key, sizeof key, "tl_at_%zu", i); ctx->stats[key] = at; }
				ctx->stats["tl_count"] = (double) (i + 1);
			}
		}
		ctx->event_pool.push_back(p.second.first);
		ctx->event_pool.push_back(p.second.second);
	}
	ctx->pending.clear();
	ctx->call_first_pending = 0;
}

// The geometry of one scan: warps per block (column range), tiles per warp, columns per block.
struct geometry
{
	uint32_t n_tiles, n_warps, tpw, C, cg, B;
	// chunked column step (fsq_cstep.cuh): E elements per thread, T threads, symbols sym0 .. sym0 + ns - 1
	bool chunked;          // the default where it applies; otherwise the striped kernel (fsq_step.cuh)
	uint32_t E, T, ns, sym0;
	uint32_t digits;       // chunked: 2 when the comps that occur span more than 8 values (two base-4 digit steps per column)
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

	// The chunked step walks E consecutive positions per thread; it needs the symbols that occur to span at most 8 values.
	uint32_t const mask = msa->sym_mask ? msa->sym_mask : ((1u << msa->nsym) - 1u);
	uint32_t lo = 0, hi = 31;
	while (!((mask >> lo) & 1u)) ++lo;
	while (!((mask >> hi) & 1u)) --hi;
	g.sym0 = lo;
	g.ns = hi - lo + 1;
	int64_t const kern = ctx->opt("kernel", 0);        // 0 auto, 1 striped, 3 chunked
	g.chunked = (0 == kern || 3 == kern) && m <= cblock_thread_limit(53) * 53u;
	g.digits = 1;
	if (g.chunked && g.ns > 8)
	{
		// 9 .. 16 comps: each column as two steps over the base-4 digits of the comps as they stand (no offset)
		g.digits = 2;
		g.sym0 = 0;
		g.ns = 4;
	}
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

	g.cg = (uint32_t) std::max <uint64_t>(1, (g.chunked ? 512 : 2048) / msa->col_stride);
	int64_t C = ctx->opt("block_columns", 0);
	if (C <= 0)
	{
		// enough blocks to fill every SM once (the same blocks serve passes A and B, so the pass with the larger
		// footprint decides), but long enough to amortise the two extra passes' fixed costs
		uint64_t per_sm = std::min <uint64_t>(32, 2048 / threads);
		per_sm = std::min <uint64_t>(per_sm, 65536 / ((uint64_t) threads * regs));
		size_t smem = 2 * (size_t) g.cg * msa->col_stride + 4 * (size_t) threads + 2 * (size_t) m + 1024;
		if (g.chunked)
		{
			// The histogram of pass B grows with the block length, so the footprint depends on the answer: take the largest
			// number of resident CTAs whose own block length fits (every CTA also reserves 1 KB of shared memory).
			per_sm = std::min <uint64_t>(per_sm, 65536 / ((uint64_t) threads * regs));
			for (; per_sm > 1; --per_sm)
			{
				block_args P;
				memset(&P, 0, sizeof P);
				P.cg = g.cg; P.col_stride = msa->col_stride; P.m = m; P.digits = g.digits;
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
		int64_t c_max = std::min <int64_t>(4096, (g.chunked ? (int64_t) FSQ_CODE_LIMIT : 65534) - (int64_t) m - 2);
		if (g.chunked)
		{
			// one CTA per SM and still too large (many rows): the histogram has one counter per row and per column of the block
			block_args P;
			memset(&P, 0, sizeof P);
			P.cg = g.cg; P.col_stride = msa->col_stride; P.m = m; P.C = 0; P.digits = g.digits;
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
	C = std::min <int64_t>(C, (g.chunked ? (int64_t) FSQ_CODE_LIMIT : 65534) - (int64_t) m - 2);
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
	return launch_block(A, n_blocks, g.n_warps, g.tpw, hist, st);
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
	A.digits = g.digits;
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
	uint32_t const *states, uint64_t stride, uint64_t const *rbs, uint32_t const *ctl, uint32_t m, uint32_t C, uint64_t col0,
	uint32_t const *oldvals, uint32_t const *R, uint32_t *dst_a, uint32_t *dst_d
)
{
	// rbs: the right bounds of all reduced segments; this rank's are [ctl[I_SEG_FIRST], ctl[I_SEG_LAST]) (the grid may be larger)
	uint32_t const t = blockIdx.x;                 // grid.x: up to 2^31 - 1 positions; one CTA copies one state
	if (ctl[I_SEG_FIRST] + t >= ctl[I_SEG_LAST]) return;
	uint64_t const k = rbs[ctl[I_SEG_FIRST] + t] - col0 - 1;          // local column whose output state this is (rbs are global right bounds)
	uint32_t const b = (uint32_t) (k / C);
	uint32_t const r = R[b];
	uint32_t const base = (uint32_t) (col0 + (uint64_t) b * C + 1) - r;
	uint32_t const *src = states + k * stride;
	uint32_t const *old = oldvals + (size_t) b * m;
	uint32_t *da = dst_a + (size_t) t * m, *dd = dst_d + (size_t) t * m;
	auto value = [&](uint32_t w) { uint32_t const code = w >> 16; return code >= r ? base + code : old[code]; };
	if (0 == (m & 3u))
	{
		// four rows per thread and access (the stored rows start at multiples of 16 bytes, the outputs too when m % 4 == 0)
		uint4 const *src4 = reinterpret_cast <uint4 const *>(src);
		uint4 *da4 = reinterpret_cast <uint4 *>(da), *dd4 = reinterpret_cast <uint4 *>(dd);
		for (uint32_t i = threadIdx.x; i < m / 4; i += blockDim.x)
		{
			uint4 const w = src4[i];
			da4[i] = make_uint4(w.x & 0xffffu, w.y & 0xffffu, w.z & 0xffffu, w.w & 0xffffu);
			dd4[i] = make_uint4(value(w.x), value(w.y), value(w.z), value(w.w));
		}
		return;
	}
	for (uint32_t i = threadIdx.x; i < m; i += blockDim.x)
	{
		uint32_t const w = src[i];
		da[i] = w & 0xffffu;
		dd[i] = value(w);
	}
}

// Everything the scans share: local summaries, true block start states, their rank compression.
struct scan_state
{
	geometry g;
	dbuf sum_a, sum_d;          // [B][m]: block summaries (pass A); after the up-sweep, summaries of aligned spans
	dbuf st_a, st_d;            // [B + 1][m]: state at every block start (slot B: at the end of the last block)
	dbuf code, oldvals, R;
	// stitch plans: device copies and their pinned staging, one pair per plan, alive until the run ends (no plan waits
	// for the copies of the one before it)
	std::vector <std::unique_ptr <dbuf>> plan_dev;
	std::vector <std::unique_ptr <hbuf>> plan_host;
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
	stitch_item *d_items = nullptr;
	uint32_t *d_begin = nullptr;
	if (n)
	{
		// pinned staging lives in the scan state: it is released only after the caller's final synchronisation
		size_t const ib = n * sizeof(stitch_item), bb = (n + 1) * 4, ib_al = (ib + 15) & ~size_t(15);
		S.plan_host.emplace_back(new hbuf); S.plan_dev.emplace_back(new dbuf);
		hbuf &h = *S.plan_host.back(); dbuf &d = *S.plan_dev.back();
		CK(h.alloc(ctx, ib_al + bb)); CK(d.alloc(ctx, ib_al + bb));
		memcpy(h.p, plan.items.data(), ib);
		uint32_t *hb = (uint32_t *) (h.as <uint8_t>() + ib_al);
		for (size_t i = 0; i <= n; ++i) hb[i] = (uint32_t) i;
		CK(cudaMemcpyAsync(d.p, h.p, ib_al + bb, cudaMemcpyHostToDevice, ctx->stream));
		d_items = d.as <stitch_item>();
		d_begin = (uint32_t *) (d.as <uint8_t>() + ib_al);
	}
	for (size_t l = 0; l + 1 < plan.level_begin.size(); ++l)
	{
		if (l < copies.size())
			for (auto const &c : copies[l])
				CK(cudaMemcpyAsync(c.first, c.second, (size_t) m * 4, cudaMemcpyDeviceToDevice, ctx->stream));
		size_t const first = plan.level_begin[l], count = plan.level_begin[l + 1] - first;
		if (0 == count) continue;
		stitch_args SA;
		SA.items = d_items;
		SA.item_begin = d_begin + first;
		SA.n_ctas = (uint32_t) count;
		SA.m = m;
		SA.n_tiles = S.g.n_tiles;
		SA.use_sort = (uint32_t) std::min <int64_t>(2, std::max <int64_t>(0, ctx->opt("stitch_sort", 2)));      // 0 radix passes, 1 sorting network, 2 counting (m <= 1024)
		CK(launch_stitch(SA, ctx->stream));
		ctx->stats["kernel_launches"] += 1;
	}
	return FSQ_OK;
}

// Pass A (local scans) and the prefix composition of the block summaries (composing summaries is
// associative): an up-sweep turns the block summaries into summaries of aligned power-of-two spans in place, a down-sweep
// turns those into the state at every block start *relative to the state at the first local column* -- the identity
// state for an alignment that starts at column 0, which makes them the true states.  2 log2(B) levels of independent stitches.
// A column shard behind the first does not know its start state before the ranks have exchanged the summaries of their
// whole shards: it runs the up-sweep with a side chain that yields that summary, and seeds the down-sweep with the true
// start state once it has arrived.
typedef std::vector <std::vector <std::pair <uint32_t *, uint32_t const *>>> stitch_copies;

int scan_pass_a(fsq_ctx *ctx, fsq_msa const *msa, scan_state &S)
{
	S.g = choose_geometry(ctx, msa);
	geometry const &g = S.g;
	uint32_t const m = msa->m;
	uint32_t const B = g.B;
	size_t const bm = (size_t) B * m;
	CK(S.sum_a.alloc(ctx, bm * 4)); CK(S.sum_d.alloc(ctx, bm * 4));
	CK(S.st_a.alloc(ctx, (bm + m) * 4)); CK(S.st_d.alloc(ctx, (bm + m) * 4));
	CK(S.code.alloc(ctx, bm * 2)); CK(S.oldvals.alloc(ctx, bm * 4)); CK(S.R.alloc(ctx, (size_t) B * 4));
	S.relative = 0 != msa->col0;
	phase pa(ctx, "pass_a");
	block_args A;
	fill_common(A, msa, g);
	A.mode = MODE_LOCAL;
	A.sum_a = S.sum_a.as <uint32_t>();
	A.sum_d = S.sum_d.as <uint32_t>();
	CK(launch_scan(A, B, g, false, ctx->stream));
	pa.stop();
	ctx->stats["kernel_launches"] += 1;
	return FSQ_OK;
}

static uint32_t sweep_levels(uint32_t B)
{
	uint32_t levels = 0;
	while ((1ull << (levels + 1)) <= B) ++levels;          // 2^levels <= B
	return levels;
}

// up-sweep: U[i] <- U[i - 2^l] o U[i] for every i with (i + 1) a multiple of 2^(l + 1).
// total (optional, 2 m words: a then d): the summary of all B blocks.  B is a sum of powers of two, highest first; the spans
// they stand for are complete nodes of the up-sweep, and the side chain composes them from the right as they complete --
// (node of bit l) o (what lies behind it) in the launch of level l -- without touching U, which the down-sweep still needs.
void plan_up_sweep(fsq_msa const *msa, scan_state &S, stitch_plan &plan, stitch_copies &copies, uint32_t *total)
{
	geometry const &g = S.g;
	uint32_t const m = msa->m, B = g.B;
	uint32_t *sum_a = S.sum_a.as <uint32_t>(), *sum_d = S.sum_d.as <uint32_t>();
	auto start_of = [&](uint64_t b) { return (uint32_t) (msa->col0 + b * g.C); };
	auto U = [&](uint64_t i, uint32_t *&a, uint32_t *&d) { a = sum_a + i * m; d = sum_d + i * m; };
	uint32_t const levels = sweep_levels(B);
	uint32_t const *tail_a = nullptr, *tail_d = nullptr;      // the summary of the blocks [B - rem, B), rem = B mod 2^l
	for (uint32_t l = 0; l <= levels; ++l)
	{
		uint64_t const half = 1ull << l, span = half << 1;
		copies.emplace_back();
		if (l < levels)
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
		if (total && (B & half))
		{
			uint64_t const rem = B & (half - 1);              // blocks behind this node
			uint32_t *na, *nd;
			U(B - rem - 1, na, nd);                           // complete since the launch before this one (a block summary when l == 0)
			if (!tail_a) { tail_a = na; tail_d = nd; }
			else
			{
				stitch_item it;
				memset(&it, 0, sizeof it);
				it.in_a = na; it.in_d = nd;
				it.pi = tail_a; it.delta = tail_d;
				it.out_a = total; it.out_d = total + m;
				it.s_mark = start_of(B - rem);
				plan.items.push_back(it);
				tail_a = total; tail_d = total + m;
			}
		}
		plan.end_level();
	}
	if (total && tail_a != total)
	{
		// B is a power of two: the root is the total
		copies.emplace_back();
		copies.back().push_back({total, tail_a});
		copies.back().push_back({total + m, tail_d});
		plan.end_level();
	}
}

// down-sweep: with 2^l the lowest set bit of i + 1, P[i] = U[i] when i + 1 == 2^l, else P[i - 2^l] o U[i]; P[i] is the state
// at the end of block i.  With a seed (the true state at the first local column): P[i] = seed o U[i] in the first case.
void plan_down_sweep(fsq_msa const *msa, scan_state &S, stitch_plan &plan, stitch_copies &copies, uint32_t const *seed_a, uint32_t const *seed_d)
{
	geometry const &g = S.g;
	uint32_t const m = msa->m, B = g.B;
	uint32_t *sum_a = S.sum_a.as <uint32_t>(), *sum_d = S.sum_d.as <uint32_t>();
	uint32_t *st_a = S.st_a.as <uint32_t>(), *st_d = S.st_d.as <uint32_t>();
	auto start_of = [&](uint64_t b) { return (uint32_t) (msa->col0 + b * g.C); };
	auto U = [&](uint64_t i, uint32_t *&a, uint32_t *&d) { a = sum_a + i * m; d = sum_d + i * m; };
	auto P = [&](uint64_t i, uint32_t *&a, uint32_t *&d) { a = st_a + (i + 1) * m; d = st_d + (i + 1) * m; };
	for (int l = (int) sweep_levels(B); l >= 0; --l)
	{
		uint64_t const half = 1ull << l, span = half << 1;
		copies.emplace_back();
		for (uint64_t i = half - 1; i < B; i += span)
		{
			uint32_t *ua, *ud, *pa_, *pd_;
			U(i, ua, ud); P(i, pa_, pd_);
			stitch_item it;
			memset(&it, 0, sizeof it);
			if (i + 1 == half)
			{
				if (!seed_a)
				{
					copies.back().push_back({pa_, ua});
					copies.back().push_back({pd_, ud});
					continue;
				}
				it.in_a = seed_a; it.in_d = seed_d;
				it.s_mark = start_of(0);
			}
			else
			{
				uint32_t *qa, *qd;
				P(i - half, qa, qd);
				it.in_a = qa; it.in_d = qd;
				it.s_mark = start_of(i + 1 - half);
			}
			it.pi = ua; it.delta = ud;
			it.out_a = pa_; it.out_d = pd_;
			plan.items.push_back(it);
		}
		plan.end_level();
	}
}

int local_scan(fsq_ctx *ctx, fsq_msa const *msa, scan_state &S)
{
	int rc = scan_pass_a(ctx, msa, S);
	if (rc) return rc;
	uint32_t const m = msa->m;
	iota_kernel <<<(m + 255) / 256, 256, 0, ctx->stream>>>(S.st_a.as <uint32_t>(), S.st_d.as <uint32_t>(), m, (uint32_t) msa->col0);
	CK(cudaGetLastError());
	phase ps(ctx, "stitch");
	stitch_plan plan;
	stitch_copies copies;
	plan_up_sweep(msa, S, plan, copies, nullptr);
	plan_down_sweep(msa, S, plan, copies, nullptr, nullptr);
	if ((rc = run_stitch_plan(ctx, S, plan, m, copies))) return rc;
	ps.stop();
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

// One run of the long path over the columns one rank holds (all of them on a single GPU).  Every stage is enqueued on
// the context's stream behind the one before it; what a stage decides (another DP round? lists long enough? how many
// traceback entries?) stays in the control words on the device, where the next stage reads it, and the host looks at
// them once per batch of stages.  With several ranks the exchanges are all-gathers enqueued on the same stream.
struct fsq_run
{
	fsq_ctx *ctx{};
	fsq_msa const *msa{};
	fsq_comm *comm{};
	int rank{0}, size{1};
	uint32_t m{};
	uint64_t n{};               // columns of the whole alignment
	uint64_t L{};
	uint64_t c0{}, c1{};        // this rank's columns [c0, c1)
	uint64_t n_per{};           // columns per rank (the last rank may hold fewer)
	scan_state S;

	// DP.  The table is global (every rank holds all of it) and indexed by the column an entry is computed at: entry t
	// lives at tbx[t + L - 1], so that rank r's slice is tbx[r n_per, (r + 1) n_per) and the exchange is an in-place all-gather.
	uint64_t dp_size{}, tb_len{};
	uint32_t n_entries{}, n_blk{}, n_levels{}, t_first{}, t_last{};
	uint32_t chunk{}, n_chunks{}, cap{};
	uint32_t back_bound{1u << 16};     // masks and rmq cells are rebuilt this far below the rank's first entry
	bool single_chain{false};
	dbuf tbx_max, tbx_lb, tbx_size, mk, lev, ctl, chunk_level, cnt, h0, toplist, states, xflags, sums;
	// Column shards exchange, round by round, only what another rank can read of a slice: its last `tails` entries (the
	// look-back window of the rank behind) and the window minima at the ends of its chains.  0: whole slices (one rank; or
	// the look-back bound was given up, back_bound == 0xffffffff).
	uint32_t tails{0};
	dbuf xtails, xlevels;
	uint64_t state_stride{};
	struct window { uint32_t b0, b1; uint64_t col0, col1; };      // blocks [b0, b1) = local columns [col0, col1)
	std::vector <window> windows;
	bool whole{true};           // every rank keeps the top lists of all its columns at once (one window each)
	hbuf h_x;

	// traceback (replicated), reduction (own positions + halo), reduced segments (replicated) and their states (own)
	uint32_t t_bound{}, rows_bound{}, width{}, halo{16};
	dbuf tr_lb, tr_rb, tr_max, tr_size, jump, rows, rows_a, halo_rows, pos_all, next_own, nsize_own, next_all, nsize_all, next, nsize;
	dbuf walk_jump, red_lb, red_rb, red_size, red_pos, red_a, red_d;
	dbuf lt_jump, lt_hops, lt_xport, lt_xport_all, lt_entry, lt_own, lt_own_all;      // traceback over column shards
	std::chrono::steady_clock::time_point t_mark{std::chrono::steady_clock::now()};
	bool trace_replicated{false};       // fall back to the traceback over the whole table on every rank (needs all of it gathered)
	bool tables_gathered{false};
	hbuf h_ctl, h_red;
	uint32_t h[N_CONTROL]{};    // the control words as last read

	uint32_t *tb_max() const { return tbx_max.as <uint32_t>() + (L - 1); }
	uint32_t *tb_lb() const { return tbx_lb.as <uint32_t>() + (L - 1); }
	uint32_t *tb_size() const { return tbx_size.as <uint32_t>() + (L - 1); }
	uint32_t *control() const { return ctl.as <uint32_t>(); }
};

namespace {

uint64_t columns_per_rank(uint64_t n, int size)
{
	if (size <= 1) return n;
	uint64_t const per = (n + size - 1) / size;
	return (per + 31) / 32 * 32;
}

// Device memory a run may plan with: what is free plus what the context's cache holds (option mem_limit, a test knob,
// stands in for the answer).
size_t plannable_bytes(fsq_ctx *ctx)
{
	int64_t const lim = ctx->opt("mem_limit", 0);
	if (lim > 0) return (size_t) lim;
	size_t free_b = 0, total_b = 0;
	if (cudaSuccess != cudaMemGetInfo(&free_b, &total_b)) { cudaGetLastError(); return 0; }
	size_t cached = 0;
	for (auto const &blk : ctx->dev_cache) cached += blk.bytes;
	return free_b + cached;
}

int run_init(fsq_run *R, fsq_ctx *ctx, fsq_msa const *msa, fsq_comm *comm, uint64_t n_total, uint64_t L)
{
	R->ctx = ctx; R->msa = msa; R->comm = comm; R->m = msa->m; R->n = n_total; R->L = L;
	R->rank = comm ? comm->rank : 0;
	R->size = comm ? comm->size : 1;
	R->n_per = columns_per_rank(n_total, R->size);
	R->c0 = msa->col0; R->c1 = msa->col0 + msa->n;
	if (R->size > 1)
	{
		uint64_t const want0 = (uint64_t) R->rank * R->n_per, want1 = std::min <uint64_t>(n_total, want0 + R->n_per);
		if (want0 >= n_total)
			return fail(ctx, FSQ_ERR_ARG, "%d ranks are too many for %llu columns: rank %d would hold none (fsq_shard_columns)", R->size, (unsigned long long) n_total, R->rank);
		if (R->c0 != want0 || R->c1 != want1)
			return fail(ctx, FSQ_ERR_ARG, "rank %d of %d must hold columns [%llu, %llu) (fsq_shard_columns), the alignment handed in holds [%llu, %llu)", R->rank, R->size,
				(unsigned long long) want0, (unsigned long long) want1, (unsigned long long) R->c0, (unsigned long long) R->c1);
	}
	else if (0 != R->c0 || R->c1 != n_total)
		return fail(ctx, FSQ_ERR_ARG, "a single rank holds the whole alignment: columns [0, %llu)", (unsigned long long) n_total);
	uint64_t const n = n_total;
	R->dp_size = n - L + 1;
	R->tb_len = std::max <uint64_t>(n, (uint64_t) R->size * R->n_per) + 64;
	R->n_entries = (uint32_t) (n - 2 * L + 1);
	R->n_blk = (uint32_t) (R->dp_size / 64 + 1);
	R->n_levels = 1;
	while ((1ull << R->n_levels) <= R->n_blk) ++R->n_levels;
	// entry t is computed at column t + L - 1
	uint64_t const lo = R->c0 + 1 > L ? R->c0 + 1 - L : 0, hi = R->c1 + 1 > L ? R->c1 + 1 - L : 0;
	R->t_first = (uint32_t) std::min <uint64_t>(lo, R->n_entries);
	R->t_last = (uint32_t) std::min <uint64_t>(hi, R->n_entries);
	// chains: enough to fill the device a few times over, each a whole number of 64-entry rmq blocks, on a global grid
	{
		int64_t want = ctx->opt("dp_chunks", 0);
		if (want <= 0)
		{
			int64_t const per_sm = 2048 / (32 * (int64_t) std::min <uint64_t>(L, 16));
			want = (int64_t) ctx->prop.multiProcessorCount * std::min <int64_t>(per_sm, 8);      // one resident wave of chains
		}
		uint64_t const own = std::max <uint64_t>(1, R->t_last - R->t_first);
		uint64_t c = (own + (uint64_t) want - 1) / (uint64_t) want;
		c = std::max <uint64_t>(c, std::max <uint64_t>(256, 8 * L));
		if (1 == want) c = R->n_entries;
		c = (c + 63) / 64 * 64;
		R->chunk = (uint32_t) std::min <uint64_t>(c, 0xffffffc0ull);
		R->n_chunks = (R->n_entries + R->chunk - 1) / R->chunk;
	}
	// the bound a call on this shape has worked with before, else the option, else 127
	{
		auto it = ctx->learned_cap.find(std::make_tuple(R->m, R->n, R->L));
		int64_t const start = ctx->learned_cap.end() != it ? (int64_t) it->second : ctx->opt("cap", 127);
		R->cap = (uint32_t) std::min <int64_t>(R->m, std::max <int64_t>(1, start));
	}
	R->state_stride = ((uint64_t) R->m + 3) & ~3ull;
	R->t_bound = (uint32_t) (R->dp_size / L + 2);
	R->rows_bound = 1 == R->size ? R->t_bound : (uint32_t) (R->n_per / L + 2);
	R->halo = (uint32_t) std::max <int64_t>(1, ctx->opt("halo", 16));
	R->back_bound = (uint32_t) std::min <int64_t>(0xffffffffll, std::max <int64_t>(64, ctx->opt("dp_back_bound", 1 << 16)));
	return FSQ_OK;
}

// ---- pass A and the true start states ---------------------------------------------------------------------------------
int run_prepare(fsq_run *R)
{
	fsq_ctx *ctx = R->ctx;
	uint32_t const m = R->m;
	int rc;
	if (R->size <= 1)
	{
		if ((rc = local_scan(ctx, R->msa, R->S))) return rc;
		if (R->S.relative) return fail(ctx, FSQ_ERR_ARG, "this entry point needs an alignment that starts at column 0");
		return compress_starts(ctx, R->msa, R->S);
	}
	// Column shards.  This rank's summary = the state at the end of its last block relative to the identity state at c0: it
	// comes out of the up-sweep (plan_up_sweep's side chain); every rank receives all of them and composes the ones before
	// it (composition is associative: SURVEY.md A.10) into the true state at c0, with which the down-sweep starts.
	if ((rc = scan_pass_a(ctx, R->msa, R->S))) return rc;
	CK(R->sums.alloc(ctx, ((size_t) R->size + 3) * 2 * m * 4));
	uint32_t *all = R->sums.as <uint32_t>();
	uint32_t *mine = all + (size_t) R->rank * 2 * m;
	{
		phase ps(ctx, "stitch");
		stitch_plan plan;
		stitch_copies copies;
		plan_up_sweep(R->msa, R->S, plan, copies, mine);
		if ((rc = run_stitch_plan(ctx, R->S, plan, m, copies))) return rc;
		ps.stop();
	}
	{
		phase px(ctx, "exchange");
		CK(R->comm->all_gather(mine, all, (size_t) 2 * m * 4, ctx->stream));
		px.stop();
	}
	{
		phase ps(ctx, "stitch");
		stitch_plan plan;
		stitch_copies copies;
		uint32_t *tmp = all + (size_t) R->size * 2 * m;          // two scratch states behind the summaries
		uint32_t const *seed = R->rank > 0 ? all : nullptr;
		for (int i = 1; i < R->rank; ++i)
		{
			stitch_item it;
			memset(&it, 0, sizeof it);
			uint32_t *out = tmp + (size_t) (i & 1) * 2 * m;
			it.in_a = seed; it.in_d = seed + m;
			it.pi = all + (size_t) i * 2 * m; it.delta = it.pi + m;
			it.out_a = out; it.out_d = out + m;
			it.s_mark = (uint32_t) ((uint64_t) i * R->n_per);
			plan.items.push_back(it);
			plan.end_level();
			copies.emplace_back();
			seed = out;
		}
		plan_down_sweep(R->msa, R->S, plan, copies, seed, seed ? seed + m : nullptr);
		if ((rc = run_stitch_plan(ctx, R->S, plan, m, copies))) return rc;
		uint32_t *st_a = R->S.st_a.as <uint32_t>(), *st_d = R->S.st_d.as <uint32_t>();
		if (seed)
		{
			CK(cudaMemcpyAsync(st_a, seed, (size_t) m * 4, cudaMemcpyDeviceToDevice, ctx->stream));
			CK(cudaMemcpyAsync(st_d, seed + m, (size_t) m * 4, cudaMemcpyDeviceToDevice, ctx->stream));
		}
		else
		{
			iota_kernel <<<(m + 255) / 256, 256, 0, ctx->stream>>>(st_a, st_d, m, (uint32_t) R->msa->col0);
			CK(cudaGetLastError());
		}
		ps.stop();
	}
	R->S.relative = false;
	return compress_starts(ctx, R->msa, R->S);
}

// ---- memory plan: the top lists of all columns at once, or windows of blocks over one reused buffer ---------------------
// Bitwise AND of a word over the ranks (a host round trip: used once per shape, the answer is cached in the context).
int run_agree_and(fsq_run *R, uint32_t *bits)
{
	fsq_ctx *ctx = R->ctx;
	if (R->size <= 1) return FSQ_OK;
	CK(R->xflags.alloc(ctx, ((size_t) R->size + 1) * 16));
	CK(R->h_x.alloc(ctx, ((size_t) R->size + 1) * 16));
	uint32_t *xs = R->xflags.as <uint32_t>(), *hx = R->h_x.as <uint32_t>();
	hx[0] = *bits; hx[1] = hx[2] = hx[3] = 0;
	CK(cudaMemcpyAsync(xs + 4 * (size_t) R->rank, hx, 16, cudaMemcpyHostToDevice, ctx->stream));
	if (cudaSuccess != R->comm->all_gather(xs + 4 * (size_t) R->rank, xs, 16, ctx->stream))
		return fail(ctx, FSQ_ERR_CUDA, "exchange of the memory plans failed: %s", R->comm->error.c_str());
	CK(cudaMemcpyAsync(hx, xs, (size_t) R->size * 16, cudaMemcpyDeviceToHost, ctx->stream));
	CK(cudaStreamSynchronize(ctx->stream));
	ctx->stats["host_syncs"] += 1;
	uint32_t v = 0xffffffffu;
	for (int r = 0; r < R->size; ++r) v &= hx[4 * r];
	*bits = v;
	return FSQ_OK;
}

int run_plan(fsq_run *R)
{
	fsq_ctx *ctx = R->ctx;
	geometry const &g = R->S.g;
	uint64_t const nl = R->msa->n;
	size_t const per_col = ((size_t) R->cap + 1) * sizeof(uint2);
	auto const key = std::make_tuple(R->m, R->n, R->L, R->cap, R->size, nl, g.B, g.C, ctx->opt("store_states", -1));
	auto it = ctx->plans.find(key);
	fsq_ctx::mem_plan plan;
	if (ctx->plans.end() != it && 0 == ctx->opt("mem_limit", 0) && 0 == ctx->opt("window_blocks", 0)) plan = it->second;
	else
	{
		size_t const avail = plannable_bytes(ctx);
		// beside the lists: per-column words, the tables, and -- should the lists not fit at once -- what the tail then needs
		// (divergence and prefix rows at the own traceback positions, the reduced states)
		size_t const other = (size_t) nl * 64 + (size_t) R->tb_len * 32 + ((size_t) 1 << 30) + (size_t) R->rows_bound * R->m * 16;
		size_t const states = (size_t) (R->state_stride * 4 * nl);
		int64_t const want_states = ctx->opt("store_states", -1);
		int64_t const forced = ctx->opt("window_blocks", 0);
		plan.per_window = g.B;
		if (forced > 0) plan.per_window = (uint32_t) std::min <int64_t>(forced, g.B);
		else if ((size_t) nl * per_col + other > avail)
		{
			size_t const room = avail > other ? avail - other : 0;
			uint64_t blocks = room / per_col / g.C;
			// whole resident waves when more than one fits
			uint64_t const wave = std::max <uint64_t>(1, (uint64_t) ctx->prop.multiProcessorCount);
			if (blocks > wave) blocks = blocks / wave * wave;
			if (0 == blocks)
				return fail(ctx, FSQ_ERR_NOMEM, "the top lists of one block of %u columns (%zu bytes per column at bound %u) do not fit the %zu bytes available", g.C, per_col, R->cap, room);
			plan.per_window = (uint32_t) std::min <uint64_t>(blocks, g.B);
		}
		plan.whole = plan.per_window >= g.B;
		// Pass B keeps the state of every column when 4 m n bytes fit comfortably (option store_states: -1 auto, 0 never,
		// 1 always); the states at the segment boundaries are then a gather instead of a third scan.
		plan.keep = g.chunked && 0 != want_states && plan.whole && ((1 == want_states) || states + (size_t) nl * per_col + other < avail);
		// the ranks' collectives must line up: they all take the cheaper path only if every one of them can
		uint32_t bits = (plan.whole ? 1u : 0u) | (plan.keep ? 2u : 0u);
		int const rc = run_agree_and(R, &bits);
		if (rc) return rc;
		plan.whole = 0 != (bits & 1u);
		plan.keep = plan.whole && 0 != (bits & 2u);
		ctx->plans[key] = plan;
	}
	R->whole = plan.whole;
	R->windows.clear();
	for (uint32_t b = 0; b < g.B; b += plan.per_window)
	{
		fsq_run::window w;
		w.b0 = b; w.b1 = std::min(g.B, b + plan.per_window);
		w.col0 = (uint64_t) w.b0 * g.C; w.col1 = std::min <uint64_t>(nl, (uint64_t) w.b1 * g.C);
		R->windows.push_back(w);
	}
	bool keep = plan.keep;
	if (keep && !R->states.p && cudaSuccess != R->states.alloc(ctx, (size_t) (R->state_stride * 4 * nl)))
	{
		cudaGetLastError();
		if (R->size > 1) return fail(ctx, FSQ_ERR_NOMEM, "the states of %llu columns no longer fit (the plan the ranks agreed on keeps them)", (unsigned long long) nl);
		keep = false;
	}
	if (!keep) R->states.release();
	ctx->stats["store_states"] = keep ? 1 : 0;
	ctx->stats["windows"] = (double) R->windows.size();
	uint64_t widest = 0;
	for (auto const &w : R->windows) widest = std::max(widest, w.col1 - w.col0);
	if (cudaSuccess != R->toplist.alloc(ctx, (size_t) widest * per_col))
	{
		cudaGetLastError();
		return fail(ctx, FSQ_ERR_NOMEM, "%zu bytes of top lists (%llu columns at bound %u) do not fit", (size_t) widest * per_col, (unsigned long long) widest, R->cap);
	}
	return FSQ_OK;
}

// Tables and control words as they stand before the first round.
int run_reset_tables(fsq_run *R)
{
	fsq_ctx *ctx = R->ctx;
	uint64_t const nl = R->msa->n;
	if (!R->tbx_max.p)
	{
		CK(R->tbx_max.alloc(ctx, R->tb_len * 4)); CK(R->tbx_lb.alloc(ctx, R->tb_len * 4)); CK(R->tbx_size.alloc(ctx, R->tb_len * 4));
		CK(R->mk.alloc(ctx, R->dp_size * 8)); CK(R->lev.alloc(ctx, (size_t) R->n_levels * R->n_blk * 4)); CK(R->ctl.alloc(ctx, N_CONTROL * 4));
		CK(R->cnt.alloc(ctx, nl * 4)); CK(R->h0.alloc(ctx, nl * 4));
		CK(R->chunk_level.alloc(ctx, (size_t) R->n_chunks * 8));
		if (!R->xflags.p) CK(R->xflags.alloc(ctx, ((size_t) R->size + 1) * 16));
		CK(R->h_ctl.alloc(ctx, N_CONTROL * 4));
	}
	R->tails = 0;
	if (R->size > 1 && 0xffffffffu != R->back_bound && 0 != ctx->opt("exchange_tails", 1))
	{
		R->tails = (uint32_t) std::min <uint64_t>(R->n_per, (uint64_t) R->back_bound + 64);
		CK(R->xtails.alloc(ctx, (size_t) R->size * R->tails * 4));
		CK(R->xlevels.alloc(ctx, (size_t) R->size * R->n_chunks * 4));
	}
	// entries the per-column DP writes start from 0 (the first round's stand-in for "nothing known"); the ones the
	// reference never writes keep its defaults; masks and rmq cells start from values that index inside the table.
	// With tails only this rank touches its own slice and the window below it: the rest of its copy is never read.
	uint64_t x0 = 0, x1 = R->tb_len;                    // in table positions (entry t at t + L - 1)
	if (R->tails)
	{
		uint64_t const below = (uint64_t) R->tails + 128;
		x0 = (R->c0 > below ? R->c0 - below : 0) & ~63ull;
		x1 = std::min <uint64_t>(R->tb_len, R->c0 + R->n_per + 64);
	}
	uint64_t const shift = R->L - 1;
	uint64_t const e0 = x0 > shift ? x0 - shift : 0;    // in entries
	uint64_t const e1 = x1 > shift ? x1 - shift : 0;
	CK(cudaMemsetAsync(R->tbx_max.as <uint32_t>() + x0, 0xff, (x1 - x0) * 4, ctx->stream));
	if (std::min <uint64_t>(e1, R->n_entries) > e0) CK(cudaMemsetAsync(R->tb_max() + e0, 0, (std::min <uint64_t>(e1, R->n_entries) - e0) * 4, ctx->stream));
	CK(cudaMemsetAsync(R->tbx_size.as <uint32_t>() + x0, 0xff, (x1 - x0) * 4, ctx->stream));
	CK(cudaMemsetAsync(R->tbx_lb.as <uint32_t>() + x0, 0, (x1 - x0) * 4, ctx->stream));
	{
		uint64_t const k0 = e0 & ~63ull, k1 = std::min <uint64_t>(e1, R->dp_size);
		if (k1 > k0) CK(cudaMemsetAsync(R->mk.as <uint64_t>() + k0, 0xff, (k1 - k0) * 8, ctx->stream));
		// (a stale mask of all zeros in block 0 would point one entry below the table)
		if (k0 > 0) CK(cudaMemsetAsync(R->mk.p, 0xff, (size_t) std::min <uint64_t>(64, R->dp_size) * 8, ctx->stream));
	}
	CK(cudaMemsetAsync(R->lev.p, 0, (size_t) R->n_levels * R->n_blk * 4, ctx->stream));
	CK(cudaMemsetAsync(R->ctl.p, 0, N_CONTROL * 4, ctx->stream));
	R->single_chain = false;
	return FSQ_OK;
}

// Pass B over the blocks of one window with the current top-list bound.
int run_scan_window(fsq_run *R, fsq_run::window const &w)
{
	fsq_ctx *ctx = R->ctx;
	geometry const &g = R->S.g;
	phase pb(ctx, "pass_b");
	block_args A;
	fill_common(A, R->msa, g);
	A.mode = MODE_EXACT;
	A.first_block = w.b0;
	A.st_a = R->S.st_a.as <uint32_t>();
	A.code = R->S.code.as <uint16_t>();
	A.oldvals = R->S.oldvals.as <uint32_t>();
	A.R = R->S.R.as <uint32_t>();
	A.toplist = R->toplist.as <uint2>();
	A.list_col0 = w.col0;
	A.cap = R->cap;
	A.cnt = R->cnt.as <uint32_t>();
	A.h0 = R->h0.as <uint32_t>();
	A.states = R->states.as <uint32_t>();
	A.state_stride = R->state_stride;
	CK(launch_scan(A, w.b1 - w.b0, g, true, ctx->stream));
	pb.stop();
	ctx->stats["kernel_launches"] += 1;
	return FSQ_OK;
}

// The DP arguments for the entries computed at the local columns [col0, col1) of this rank (a window), or all of them.
dp_args run_dp_args(fsq_run const *R, fsq_run::window const *w)
{
	dp_args D;
	memset(&D, 0, sizeof D);
	uint64_t const L = R->L;
	D.m = R->m; D.n = R->n; D.L = L; D.col_base = R->c0;
	D.t_first = R->t_first; D.t_last = R->t_last;
	D.list_base = R->c0;
	if (w)
	{
		uint64_t const g0 = R->c0 + w->col0, g1 = R->c0 + w->col1;
		uint64_t const lo = g0 + 1 > L ? g0 + 1 - L : 0, hi = g1 + 1 > L ? g1 + 1 - L : 0;
		D.t_first = (uint32_t) std::min <uint64_t>(lo, R->n_entries);
		D.t_last = (uint32_t) std::min <uint64_t>(hi, R->n_entries);
		D.list_base = g0;
	}
	D.toplist = R->toplist.as <uint2>(); D.cap = R->cap; D.cnt = R->cnt.as <uint32_t>(); D.h0 = R->h0.as <uint32_t>();
	D.tb_max = R->tb_max(); D.tb_lb = R->tb_lb(); D.tb_size = R->tb_size();
	D.mk = R->mk.as <uint64_t>();
	D.lev = R->lev.as <uint32_t>(); D.n_blk = R->n_blk; D.n_levels = R->n_levels;
	D.n_entries = R->n_entries;
	D.chunk = R->single_chain ? (R->n_entries + 63) / 64 * 64 : R->chunk;
	D.n_chunks = R->single_chain ? 1 : R->n_chunks;
	D.chunk_level = R->chunk_level.as <uint32_t>();
	D.flags = R->control();
	D.bound_on_device = 1;
	D.skip_when_done = 1;
	// what this launch's chains can read: their own entries and the look-back before them
	uint64_t const back = R->back_bound;
	D.w_first = (uint32_t) ((D.t_first > back ? D.t_first - back : 0) & ~63ull);
	D.w_last = (R->c1 == R->n && (!w || R->c0 + w->col1 == R->n)) ? R->n_entries : D.t_last;
	// the guess levels what no round has made final: everything while several ranks iterate together, the window's
	// entries when a single rank walks its windows in order
	D.lvl_first = 0; D.lvl_last = R->n_entries;
	if (1 == R->size && w) { D.lvl_first = D.t_first; D.lvl_last = D.t_last; }
	D.own_first = 0; D.own_last = R->n_entries;
	D.lvl_gathered = nullptr; D.lvl_ranks = 0;
	if (R->tails)
	{
		// what this rank holds of the table: its own entries and the window below them
		D.lvl_first = D.w_first; D.lvl_last = R->t_last;
		D.own_first = R->t_first; D.own_last = R->t_last;
		D.lvl_gathered = R->xlevels.as <uint32_t>(); D.lvl_ranks = (uint32_t) R->size;
	}
	return D;
}

// One DP round over the given windows: between rounds the guess (or only the rebuild), then every chain once against
// the table as it stands, then the ranks' slices and flags travel and the decision is merged on the device.
int run_enqueue_round(fsq_run *R, uint32_t round_no, size_t w_begin, size_t w_end, bool guess, bool rescan)
{
	fsq_ctx *ctx = R->ctx;
	uint32_t *ctl = R->control();
	int rc;
	uint32_t chains = 0;
	for (size_t wi = w_begin; wi < w_end; ++wi)
	{
		fsq_run::window const &w = R->windows[wi];
		bool const whole = 1 == R->windows.size();      // (of this rank: the lists of all its columns are in the buffer)
		dp_args D = run_dp_args(R, whole ? nullptr : &w);
		if (round_no > 1 && (whole || 1 == R->size || wi == w_begin))
		{
			// several ranks with windows: one guess over the whole table per round, before the first window
			dp_args G = (whole || 1 == R->size) ? D : run_dp_args(R, nullptr);
			phase pd(ctx, "dp");
			if (guess) { CK(launch_dp_guess(G, ctx->stream)); ctx->stats["kernel_launches"] += 5; }
			else { CK(launch_dp_rebuild(G, ctx->stream)); ctx->stats["kernel_launches"] += 2; }
			pd.stop();
		}
		if (rescan && (rc = run_scan_window(R, w))) return rc;
		phase pd(ctx, "dp");
		if (1 == round_no) { D.bound_on_device = 0; D.reach_bound = 0; }
		D.round_no = round_no;
		D.chain_base = 1 == R->size ? D.t_first / D.chunk : 0;
		if (D.t_first < D.t_last) chains += (D.t_last - 1) / D.chunk - D.t_first / D.chunk + 1;
		CK(launch_dp_chains(D, ctx->stream));
		pd.stop();
		ctx->stats["kernel_launches"] += 1;
	}
	uint32_t *xs = R->xflags.as <uint32_t>();
	uint32_t *mine = xs + 4 * (size_t) R->rank;
	{
		phase pd(ctx, "dp");
		CK(launch_dp_round_end(ctl, 1 == round_no ? 1u : 0u, mine, ctx->stream));
		pd.stop();
	}
	if (R->size > 1)
	{
		dp_args G = run_dp_args(R, nullptr);
		if (R->tails)
		{
			// the window minima at the ends of the own chains, for the guess that may follow
			phase pd(ctx, "dp");
			CK(launch_dp_chunk_minima(G, ctx->stream));
			pd.stop();
			ctx->stats["kernel_launches"] += 1;
		}
		phase px(ctx, "exchange");
		R->comm->group_begin();
		cudaError_t e1, e4 = cudaSuccess;
		if (R->tails)
		{
			e1 = R->comm->all_gather(R->tbx_max.as <uint32_t>() + ((size_t) R->rank + 1) * R->n_per - R->tails, R->xtails.p, (size_t) R->tails * 4, ctx->stream);
			e4 = R->comm->all_gather(G.chunk_level + R->n_chunks, R->xlevels.p, (size_t) R->n_chunks * 4, ctx->stream);
		}
		else
			e1 = R->comm->all_gather(R->tbx_max.as <uint32_t>() + (size_t) R->rank * R->n_per, R->tbx_max.p, (size_t) R->n_per * 4, ctx->stream);
		cudaError_t e2 = R->comm->all_gather(mine, xs, 16, ctx->stream);
		cudaError_t e3 = R->comm->group_end();
		if (cudaSuccess != e1 || cudaSuccess != e2 || cudaSuccess != e3 || cudaSuccess != e4)
			return fail(ctx, FSQ_ERR_CUDA, "exchange of the DP table slices failed: %s", R->comm->error.empty() ? "CUDA error" : R->comm->error.c_str());
		if (R->tails)
		{
			CK(launch_dp_scatter_tails(R->xtails.as <uint32_t>(), (uint32_t) R->size, (uint32_t) R->rank, R->tails, R->n_per, R->tbx_max.as <uint32_t>(), ctx->stream));
			ctx->stats["kernel_launches"] += 1;
		}
		px.stop();
	}
	phase pd(ctx, "dp");
	CK(launch_dp_round_merge(ctl, xs, (uint32_t) R->size, round_no, R->back_bound, (1 == R->size && chains <= 1) ? 1u : 0u, ctx->stream));
	pd.stop();
	ctx->stats["kernel_launches"] += 2;
	return FSQ_OK;
}

// Every rank receives the whole table (three in-place all-gathers): only the traceback over the whole table and the
// keep_dp test tap need it; the rounds exchange tb_max alone and the sharded traceback reads own entries only.
int run_gather_tables(fsq_run *R)
{
	fsq_ctx *ctx = R->ctx;
	if (R->size <= 1 || R->tables_gathered) return FSQ_OK;
	phase px(ctx, "exchange");
	R->comm->group_begin();
	size_t const off = (size_t) R->rank * R->n_per, bytes = (size_t) R->n_per * 4;
	cudaError_t e = R->comm->all_gather(R->tbx_max.as <uint32_t>() + off, R->tbx_max.p, bytes, ctx->stream);
	if (cudaSuccess == e) e = R->comm->all_gather(R->tbx_lb.as <uint32_t>() + off, R->tbx_lb.p, bytes, ctx->stream);
	if (cudaSuccess == e) e = R->comm->all_gather(R->tbx_size.as <uint32_t>() + off, R->tbx_size.p, bytes, ctx->stream);
	cudaError_t const e2 = R->comm->group_end();
	if (cudaSuccess != e || cudaSuccess != e2)
		return fail(ctx, FSQ_ERR_CUDA, "exchange of the DP table failed: %s", R->comm->error.empty() ? "CUDA error" : R->comm->error.c_str());
	px.stop();
	R->tables_gathered = true;
	return FSQ_OK;
}

// The last entry (segmentation_lp_context.cc:165-185), on the rank that holds the last column; then every rank learns
// whether any list was too short, and (several ranks) receives the rest of the table.
int run_enqueue_final(fsq_run *R, bool rescan_last_window, bool traceback_follows = false)
{
	fsq_ctx *ctx = R->ctx;
	uint32_t *ctl = R->control();
	if (R->c1 == R->n)
	{
		if (rescan_last_window)
		{
			int const rc = run_scan_window(R, R->windows.back());
			if (rc) return rc;
		}
		phase pd(ctx, "dp");
		dp_args D = run_dp_args(R, 1 == R->windows.size() ? nullptr : &R->windows.back());
		CK(launch_dp_final(D, ctx->stream));
		pd.stop();
		ctx->stats["kernel_launches"] += 1;
	}
	if (R->size > 1)
	{
		R->tables_gathered = false;
		// (the ranks' verdicts on their lists travel with the traceback's exports when that is enqueued right behind)
		if (traceback_follows && !R->trace_replicated && !ctx->opt("keep_dp", 0)) return FSQ_OK;
		uint32_t *xs = R->xflags.as <uint32_t>();
		uint32_t *mine = xs + 4 * (size_t) R->rank;
		CK(launch_dp_round_end(ctl, 0, mine, ctx->stream));
		phase px(ctx, "exchange");
		if (cudaSuccess != R->comm->all_gather(mine, xs, 16, ctx->stream))
			return fail(ctx, FSQ_ERR_CUDA, "exchange of the DP flags failed: %s", R->comm->error.empty() ? "CUDA error" : R->comm->error.c_str());
		px.stop();
		R->tables_gathered = false;
		if (R->trace_replicated || ctx->opt("keep_dp", 0))
		{
			int const rc = run_gather_tables(R);
			if (rc) return rc;
		}
		CK(launch_dp_overflow_merge(ctl, xs, (uint32_t) R->size, ctx->stream));
		ctx->stats["kernel_launches"] += 2;
	}
	return FSQ_OK;
}

// The control words to the host (blocking).
int run_read_control(fsq_run *R)
{
	fsq_ctx *ctx = R->ctx;
	CK(cudaMemcpyAsync(R->h_ctl.p, R->ctl.p, N_CONTROL * 4, cudaMemcpyDeviceToHost, ctx->stream));
	{
		// how long the host spent enqueueing since the call (or the last read) began, and how long it then waited for the
		// device: a host that waits little is the one holding the device up
		auto const t0 = std::chrono::steady_clock::now();
		ctx->stats["hms_enqueue"] += std::chrono::duration <double, std::milli>(t0 - R->t_mark).count();
		CK(cudaStreamSynchronize(ctx->stream));
		R->t_mark = std::chrono::steady_clock::now();
		ctx->stats["hms_wait"] += std::chrono::duration <double, std::milli>(R->t_mark - t0).count();
	}
	ctx->stats["host_syncs"] += 1;
	memcpy(R->h, R->h_ctl.p, N_CONTROL * 4);
	if (R->comm && !R->comm->error.empty()) return fail(ctx, FSQ_ERR_CUDA, "%s", R->comm->error.c_str());
	return FSQ_OK;
}

// ---- traceback, reduction, reduced segments, their states: everything behind the DP that needs no host decision ----------
int run_alloc_tail(fsq_run *R)
{
	fsq_ctx *ctx = R->ctx;
	if (R->tr_lb.p) return FSQ_OK;
	uint32_t const tb = R->t_bound, rb = R->rows_bound;
	CK(R->tr_lb.alloc(ctx, (size_t) tb * 4)); CK(R->tr_rb.alloc(ctx, (size_t) tb * 8)); CK(R->tr_max.alloc(ctx, (size_t) tb * 4)); CK(R->tr_size.alloc(ctx, (size_t) tb * 4));
	CK(R->next_own.alloc(ctx, ((size_t) rb + 1) * 4)); CK(R->nsize_own.alloc(ctx, ((size_t) rb + 1) * 4));
	if (R->size > 1)
	{
		CK(R->pos_all.alloc(ctx, ((size_t) R->size + 1) * 4));
		CK(R->next_all.alloc(ctx, (size_t) R->size * (rb + 1) * 4)); CK(R->nsize_all.alloc(ctx, (size_t) R->size * (rb + 1) * 4));
		CK(R->next.alloc(ctx, (size_t) tb * 4)); CK(R->nsize.alloc(ctx, (size_t) tb * 4));
	}
	uint32_t wl = 1;
	while ((1ull << wl) <= tb) ++wl;
	CK(R->walk_jump.alloc(ctx, (size_t) wl * tb * 4));
	CK(R->red_lb.alloc(ctx, (size_t) tb * 8)); CK(R->red_rb.alloc(ctx, (size_t) tb * 8)); CK(R->red_size.alloc(ctx, (size_t) tb * 4)); CK(R->red_pos.alloc(ctx, (size_t) tb * 4));
	return FSQ_OK;
}

int run_enqueue_traceback(fsq_run *R)
{
	fsq_ctx *ctx = R->ctx;
	int rc = run_alloc_tail(R);
	if (rc) return rc;
	phase pt(ctx, "traceback");
	if (R->size > 1 && !R->trace_replicated)
	{
		// every rank over its own entries; the ranks exchange where a path can enter them and their parts of the path
		ltrace_args A;
		memset(&A, 0, sizeof A);
		A.tb_lb = R->tb_lb(); A.tb_max = R->tb_max(); A.tb_size = R->tb_size();
		A.L = (uint32_t) R->L; A.dp_size = (uint32_t) R->dp_size;
		A.own_first = R->t_first; A.own_last = R->t_last;
		A.has_final = R->c1 == R->n ? 1u : 0u;
		A.own_n = R->t_last - R->t_first + A.has_final;
		// the top table must span any path inside the rank (its jumps are the exits): 2^(levels - 1) hops
		A.levels = 2;
		while ((1ull << (A.levels - 1)) <= (uint64_t) A.own_n / R->L + 2) ++A.levels;
		A.window = (uint32_t) std::min <uint64_t>(std::max <uint64_t>(R->n_per, 1), (uint64_t) std::max <int64_t>(1, ctx->opt("trace_window", 1 << 15)));
		A.rank = (uint32_t) R->rank; A.size = (uint32_t) R->size; A.n_per = R->n_per;
		A.rows_bound = R->rows_bound; A.t_bound = R->t_bound;
		size_t const xs = (size_t) A.window + 3;
		if (!R->lt_jump.p)
		{
			CK(R->lt_jump.alloc(ctx, (size_t) A.levels * std::max <uint32_t>(A.own_n, 1) * 4)); CK(R->lt_hops.alloc(ctx, (size_t) A.levels * std::max <uint32_t>(A.own_n, 1) * 4));
			CK(R->lt_xport.alloc(ctx, xs * 8)); CK(R->lt_xport_all.alloc(ctx, xs * 8 * R->size));
			CK(R->lt_entry.alloc(ctx, (size_t) R->size * 4));
			CK(R->lt_own.alloc(ctx, (size_t) A.rows_bound * 20)); CK(R->lt_own_all.alloc(ctx, (size_t) A.rows_bound * 20 * R->size));
		}
		A.jump = R->lt_jump.as <uint32_t>(); A.hops = R->lt_hops.as <uint32_t>();
		A.xport = R->lt_xport.as <uint2>(); A.xport_all = R->lt_xport_all.as <uint2>();
		A.ctl = R->control();
		A.pos_first_all = R->pos_all.as <uint32_t>(); A.entry_all = R->lt_entry.as <uint32_t>();
		A.own = R->lt_own.as <uint32_t>(); A.own_all = R->lt_own_all.as <uint32_t>();
		A.out_lb = R->tr_lb.as <uint32_t>(); A.out_rb = R->tr_rb.as <uint64_t>(); A.out_max = R->tr_max.as <uint32_t>(); A.out_size = R->tr_size.as <uint32_t>();
		cudaError_t e = launch_ltrace_tables(A, ctx->stream);
		if (cudaErrorNotSupported == e) { cudaGetLastError(); R->trace_replicated = true; }
		else
		{
			CK(e);
			if (cudaSuccess != R->comm->all_gather(R->lt_xport.p, R->lt_xport_all.p, xs * 8, ctx->stream))
				return fail(ctx, FSQ_ERR_CUDA, "exchange of the traceback exports failed: %s", R->comm->error.c_str());
			CK(launch_ltrace_resolve(A, ctx->stream));
			if (cudaSuccess != R->comm->all_gather(R->lt_own.p, R->lt_own_all.p, (size_t) A.rows_bound * 20, ctx->stream))
				return fail(ctx, FSQ_ERR_CUDA, "exchange of the traceback parts failed: %s", R->comm->error.c_str());
			CK(launch_ltrace_scatter(A, ctx->stream));
			ctx->stats["kernel_launches"] += 3;
			pt.stop();
			return FSQ_OK;
		}
	}
	if ((rc = run_gather_tables(R))) return rc;
	trace_args TA;
	memset(&TA, 0, sizeof TA);
	TA.tb_lb = R->tb_lb(); TA.tb_max = R->tb_max(); TA.tb_size = R->tb_size();
	TA.L = (uint32_t) R->L; TA.dp_size = (uint32_t) R->dp_size;
	TA.levels = 1;
	while ((1ull << TA.levels) <= R->dp_size / R->L + 1) ++TA.levels;
	if (!R->jump.p) CK(R->jump.alloc(ctx, (size_t) TA.levels * R->dp_size * 4));
	TA.jump = R->jump.as <uint32_t>(); TA.ctl = R->control();
	TA.t_bound = R->t_bound;
	TA.out_lb = R->tr_lb.as <uint32_t>(); TA.out_rb = R->tr_rb.as <uint64_t>(); TA.out_max = R->tr_max.as <uint32_t>(); TA.out_size = R->tr_size.as <uint32_t>();
	TA.c0 = R->c0; TA.c1 = R->c1;
	CK(launch_trace_tables(TA, ctx->stream));
	CK(launch_trace_gather(TA, ctx->stream));
	ctx->stats["kernel_launches"] += 3;
	pt.stop();
	return FSQ_OK;
}

// Greedy reduction (segmentation_lp_context.cc:335-390) for the own traceback positions, from rows that are either list
// rows (filled here from the top lists) or divergence rows (filled by the caller from a re-scan); then the walk.
int run_enqueue_reduce(fsq_run *R, uint32_t row_kind)
{
	fsq_ctx *ctx = R->ctx;
	uint32_t const m = R->m;
	phase pr(ctx, "reduce");
	reduce_args A;
	memset(&A, 0, sizeof A);
	A.ctl = R->control();
	A.m = m; A.width = R->cap + 1; A.row_kind = row_kind;
	A.rows_bound = R->rows_bound;
	A.tb_rb = R->tr_rb.as <uint64_t>(); A.tb_size = R->tr_size.as <uint32_t>();
	size_t const row_bytes = row_kind ? (size_t) m * 4 : (size_t) A.width * sizeof(uint2);
	if (0 == row_kind)
	{
		CK(R->rows.alloc(ctx, (size_t) R->rows_bound * row_bytes));
		A.own_rows = R->rows.p;
		A.toplist = R->toplist.as <uint2>(); A.cap = R->cap; A.cnt = R->cnt.as <uint32_t>(); A.col_base = R->c0;
		CK(launch_reduce_lists(A, ctx->stream));
		ctx->stats["kernel_launches"] += 1;
	}
	A.own_rows = R->rows.p;
	A.rank = (uint32_t) R->rank; A.size = (uint32_t) R->size;
	if (R->size > 1)
	{
		// the first rows of every rank: where a run of this rank's last positions continues
		A.halo = std::min(R->halo, R->rows_bound);
		CK(R->halo_rows.alloc(ctx, (size_t) R->size * A.halo * row_bytes));
		phase px(ctx, "exchange");
		if (cudaSuccess != R->comm->all_gather(R->rows.p, R->halo_rows.p, (size_t) A.halo * row_bytes, ctx->stream))
			return fail(ctx, FSQ_ERR_CUDA, "exchange of the reduction rows failed: %s", R->comm->error.c_str());
		px.stop();
		CK(launch_rank_positions(A.tb_rb, A.ctl, R->n_per, (uint32_t) R->size, R->pos_all.as <uint32_t>(), ctx->stream));
		A.halo_rows = R->halo_rows.p;
		A.pos_first_all = R->pos_all.as <uint32_t>();
		ctx->stats["kernel_launches"] += 1;
	}
	A.next = R->next_own.as <uint32_t>(); A.next_size = R->nsize_own.as <uint32_t>();
	CK(launch_reduce(A, ctx->stream));
	ctx->stats["kernel_launches"] += 1;
	uint32_t const *next = A.next, *nsize = A.next_size;
	if (R->size > 1)
	{
		phase px(ctx, "exchange");
		R->comm->group_begin();
		cudaError_t e = R->comm->all_gather(R->next_own.p, R->next_all.p, ((size_t) R->rows_bound + 1) * 4, ctx->stream);
		if (cudaSuccess == e) e = R->comm->all_gather(R->nsize_own.p, R->nsize_all.p, ((size_t) R->rows_bound + 1) * 4, ctx->stream);
		cudaError_t const e2 = R->comm->group_end();
		if (cudaSuccess != e || cudaSuccess != e2)
			return fail(ctx, FSQ_ERR_CUDA, "exchange of the reduction's runs failed: %s", R->comm->error.c_str());
		px.stop();
		CK(launch_scatter_next(R->next_all.as <uint32_t>(), R->nsize_all.as <uint32_t>(), R->pos_all.as <uint32_t>(), (uint32_t) R->size, R->rows_bound, R->t_bound,
			R->next.as <uint32_t>(), R->nsize.as <uint32_t>(), R->control(), ctx->stream));
		ctx->stats["kernel_launches"] += 1;
		next = R->next.as <uint32_t>(); nsize = R->nsize.as <uint32_t>();
	}
	// the walk u -> next[u] + 1 from 0, on the device
	walk_args W;
	memset(&W, 0, sizeof W);
	W.t_bound = R->t_bound;
	W.levels = 1;
	while ((1ull << W.levels) <= R->t_bound) ++W.levels;
	W.ctl = R->control();
	W.next = next; W.next_size = nsize; W.tb_rb = R->tr_rb.as <uint64_t>();
	W.jump = R->walk_jump.as <uint32_t>();
	W.out_lb = R->red_lb.as <uint64_t>(); W.out_rb = R->red_rb.as <uint64_t>(); W.out_size = R->red_size.as <uint32_t>(); W.out_pos = R->red_pos.as <uint32_t>();
	W.c0 = R->c0; W.c1 = R->c1;
	CK(launch_walk(W, ctx->stream));
	ctx->stats["kernel_launches"] += 1;
	pr.stop();
	return FSQ_OK;
}

// (a, d) at the right bounds of this rank's reduced segments out of the states pass B kept; `bound` segments at most.
int run_enqueue_state_gather(fsq_run *R, uint32_t bound)
{
	fsq_ctx *ctx = R->ctx;
	uint32_t const m = R->m;
	CK(R->red_a.alloc(ctx, (size_t) std::max <uint32_t>(bound, 1) * m * 4)); CK(R->red_d.alloc(ctx, (size_t) std::max <uint32_t>(bound, 1) * m * 4));
	if (0 == bound) return FSQ_OK;
	phase pc(ctx, "pass_c");
	gather_stored_states_kernel <<<bound, 256, 0, ctx->stream>>>(R->states.as <uint32_t>(), R->state_stride, R->red_rb.as <uint64_t>(), R->control(), m, R->S.g.C, R->c0,
		R->S.oldvals.as <uint32_t>(), R->S.R.as <uint32_t>(), R->red_a.as <uint32_t>(), R->red_d.as <uint32_t>());
	CK(cudaGetLastError());
	pc.stop();
	ctx->stats["kernel_launches"] += 1;
	return FSQ_OK;
}

// The reduced bounds to the host (what segmentation_container holds); `count` entries.
int run_enqueue_bounds_copy(fsq_run *R, uint32_t count)
{
	fsq_ctx *ctx = R->ctx;
	CK(R->h_red.alloc(ctx, (size_t) std::max <uint32_t>(count, 1) * 20));
	uint64_t *h_lb = R->h_red.as <uint64_t>(), *h_rb = h_lb + count;
	uint32_t *h_sz = (uint32_t *) (h_rb + count);
	CK(cudaMemcpyAsync(h_lb, R->red_lb.p, (size_t) count * 8, cudaMemcpyDeviceToHost, ctx->stream));
	CK(cudaMemcpyAsync(h_rb, R->red_rb.p, (size_t) count * 8, cudaMemcpyDeviceToHost, ctx->stream));
	CK(cudaMemcpyAsync(h_sz, R->red_size.p, (size_t) count * 4, cudaMemcpyDeviceToHost, ctx->stream));
	return FSQ_OK;
}

// Single-chain fallback when the rounds do not settle: the reference's own order, exact in one pass.  Several ranks take
// turns, each behind the complete prefix.
int run_sequential_dp(fsq_run *R, bool rescan)
{
	fsq_ctx *ctx = R->ctx;
	int rc;
	R->single_chain = true;
	R->tails = 0;               // whole slices from here on: every turn ends with one
	ctx->stats["dp_fallbacks"] += 1;
	CK(cudaMemsetAsync(R->control() + F_DONE, 0, 4, ctx->stream));
	for (int r = 0; r < R->size; ++r)
	{
		dp_args G = run_dp_args(R, nullptr);
		G.skip_when_done = 0;
		CK(launch_dp_rebuild(G, ctx->stream));
		if (r == R->rank)
			for (auto const &w : R->windows)
			{
				if (rescan && (rc = run_scan_window(R, w))) return rc;
				dp_args D = run_dp_args(R, 1 == R->windows.size() ? nullptr : &w);
				D.skip_when_done = 0;
				D.bound_on_device = 0; D.reach_bound = 0xffffffffu;
				// one chain over the window's entries
				D.chunk = (R->n_entries + 63) / 64 * 64; D.n_chunks = 1;
				CK(launch_dp_chains(D, ctx->stream));
			}
		if (R->size > 1 && cudaSuccess != R->comm->all_gather(R->tbx_max.as <uint32_t>() + (size_t) R->rank * R->n_per, R->tbx_max.p, (size_t) R->n_per * 4, ctx->stream))
			return fail(ctx, FSQ_ERR_CUDA, "exchange of the DP table slices failed: %s", R->comm->error.c_str());
	}
	// the lists were judged by the chains just run: merge the ranks' verdicts
	CK(cudaMemsetAsync(R->control() + F_WINDOW_ERR, 0, 4, ctx->stream));
	return FSQ_OK;
}

} // namespace

namespace {

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


namespace {

// The tail behind a settled DP table when pass B kept every column's state and the top lists are resident: traceback,
// reduction, walk, state gather and the copies of the results, all enqueued; nothing here waits for the host.
int run_enqueue_fast_tail(fsq_run *R, bool with_states, bool with_bounds)
{
	int rc;
	if ((rc = run_enqueue_traceback(R))) return rc;
	if ((rc = run_enqueue_reduce(R, 0))) return rc;
	if (with_states && (rc = run_enqueue_state_gather(R, R->rows_bound))) return rc;
	if (with_bounds && (rc = run_enqueue_bounds_copy(R, R->t_bound))) return rc;
	return FSQ_OK;
}

__global__ void gather_rows_kernel(uint32_t const *rows_a, uint32_t const *rows_d, uint32_t const *red_pos, uint32_t const *ctl, uint32_t m, uint32_t *dst_a, uint32_t *dst_d)
{
	uint32_t const s = ctl[I_SEG_FIRST] + blockIdx.x;
	if (s >= ctl[I_SEG_LAST]) return;
	uint32_t const i = blockIdx.y * blockDim.x + threadIdx.x;
	if (i >= m) return;
	size_t const from = (size_t) (red_pos[s] - ctl[I_POS_FIRST]) * m + i, to = (size_t) blockIdx.x * m + i;
	dst_a[to] = rows_a[from];
	dst_d[to] = rows_d[from];
}

// The whole long path for this rank's columns.  On FSQ_ERR_NOT_REDUCIBLE *out still receives the traceback.
int run_segment(fsq_run *R, fsq_segmentation **out)
{
	fsq_ctx *ctx = R->ctx;
	uint32_t const m = R->m;
	int rc;
	if ((rc = run_prepare(R))) return rc;
	geometry const &g = R->S.g;

	// A round costs one chain's length of sequential work, the single-chain fallback as much as n_chunks rounds (and round r
	// leaves the first r chains exact, so n_chunks rounds always suffice): give up on the rounds only after half of that.
	// Slowly rising levels (thousands of distinct rows per segment) take 15-20 rounds where config 2 takes two.
	int64_t max_rounds = ctx->opt("dp_max_rounds", 0);
	if (max_rounds <= 0) max_rounds = std::max <int64_t>(16, R->n_chunks / 2);
	max_rounds = std::max <int64_t>(2, max_rounds);
	int64_t guess_rounds = ctx->opt("dp_guess_rounds", -1);           // after these the table is left to settle by itself
	if (guess_rounds < 0) guess_rounds = max_rounds;

	bool tail_done = false;
	uint32_t resume_block = 0;            // one rank, windows: everything before this block's first entry is final
	uint32_t total_rounds = 0;
	auto grow_cap = [&]() -> int {
		if (R->cap >= m) return fail(ctx, FSQ_ERR_INTERNAL, "DP reported an incomplete top list with cap = m");
		// The lists double.  Every value computed under lists that are too short is an upper bound of the true one, so lists
		// that reach as far as the largest of them (F_OVERFLOW) are proven to suffice -- but that bound is loose (one entry
		// without a good candidate in its list inflates everything computed from it; measured on config 3: 4,095 where 1,023
		// does), so it only serves when it is the smaller of the two.  A retry is cheap: the chains that are final already
		// (F_SURE) report a short list after the first rounds, before the rest of the table has settled.
		uint64_t const entries = (uint64_t) R->cap + 1;
		uint64_t next = 2 * entries;
		if (R->h[F_OVERFLOW] && !R->h[F_BLIND]) next = std::min <uint64_t>(next, (uint64_t) R->h[F_OVERFLOW] + R->h[F_OVERFLOW] / 8 + 8);
		next = std::max <uint64_t>(next, entries + 1);
		next = (next + 7) & ~7ull;
		R->cap = (uint32_t) std::min <uint64_t>(m, next - 1);
		ctx->stats["cap_retries"] += 1;
		R->toplist.release();
		R->rows.release();
		return FSQ_OK;
	};
	auto lists_too_short = [&]() { return 0 != R->h[F_OVERFLOW] || 0 != R->h[F_BLIND]; };

	while (true)
	{
		if ((rc = run_plan(R))) return rc;
		bool const whole = R->whole;
		tail_done = false;
		if (whole || R->size > 1)
		{
			// rounds over everything this rank holds: the one window, or (several ranks) every window in every round
			if ((rc = run_reset_tables(R))) return rc;
			bool const rescan = R->windows.size() > 1;
			if (!rescan && (rc = run_scan_window(R, R->windows[0]))) return rc;
			bool const can_fast_tail = whole && nullptr != R->states.p && (size_t) R->rows_bound * m * 8 <= ((size_t) 1 << 30) && 0 == ctx->opt("keep_dp", 0);
			uint32_t rounds = 0, batch = (1 == R->size && 1 == R->n_chunks) ? 1 : 2;
			bool restart = false;
			while (true)
			{
				for (uint32_t i = 0; i < batch && rounds < (uint32_t) max_rounds; ++i)
				{
					++rounds;
					if ((rc = run_enqueue_round(R, rounds, 0, R->windows.size(), (int64_t) rounds - 1 <= guess_rounds, rescan))) return rc;
				}
				bool const with_tail = can_fast_tail && rounds <= 2;
				if ((rc = run_enqueue_final(R, rescan, with_tail))) return rc;
				if (with_tail && (rc = run_enqueue_fast_tail(R, true, (size_t) R->t_bound * 20 <= ((size_t) 4 << 20)))) return rc;
				if ((rc = run_read_control(R))) return rc;
				if (R->h[F_WINDOW_ERR] && R->size > 1 && 0xffffffffu != R->back_bound) { R->back_bound = 0xffffffffu; restart = true; ctx->stats["dp_window_restarts"] += 1; break; }
				if (R->h[F_DONE]) { tail_done = with_tail; break; }
				if (R->h[F_SURE] && lists_too_short()) break;          // chains that are final already ran out of list: no point in settling the rest
				if (rounds >= (uint32_t) max_rounds)
				{
					// the guesses keep failing: the chains one after the other, exact in a single pass
					R->back_bound = 0xffffffffu;
					if ((rc = run_sequential_dp(R, rescan))) return rc;
					if ((rc = run_enqueue_final(R, rescan))) return rc;
					if ((rc = run_read_control(R))) return rc;
					++rounds;
					break;
				}
				batch = std::min <uint32_t>(8, rounds);
			}
			total_rounds += R->h[F_DONE] ? R->h[F_ROUNDS] : rounds;
			if (restart) continue;
		}
		else
		{
			// one rank, the columns in windows over one list buffer: a window's entries settle (behind final ones) before the
			// next window is scanned
			if (0 == resume_block && (rc = run_reset_tables(R))) return rc;
			bool overflow = false;
			for (size_t wi = 0; wi < R->windows.size() && !overflow; ++wi)
			{
				fsq_run::window const &w = R->windows[wi];
				if (w.b1 <= resume_block) continue;
				if ((rc = run_scan_window(R, w))) return rc;
				CK(cudaMemsetAsync(R->control(), 0, 8 * 4, ctx->stream));          // the DP words; the traceback's stay
				R->single_chain = false;
				uint32_t rounds = 0, batch = 2;
				while (true)
				{
					for (uint32_t i = 0; i < batch && rounds < (uint32_t) max_rounds; ++i)
					{
						++rounds;
						if ((rc = run_enqueue_round(R, rounds, wi, wi + 1, (int64_t) rounds - 1 <= guess_rounds, false))) return rc;
					}
					bool const last = wi + 1 == R->windows.size();
					if (last && (rc = run_enqueue_final(R, false))) return rc;
					if ((rc = run_read_control(R))) return rc;
					if (R->h[F_WINDOW_ERR] && 0xffffffffu != R->back_bound) { R->back_bound = 0xffffffffu; rounds = 0; CK(cudaMemsetAsync(R->control(), 0, 8 * 4, ctx->stream)); continue; }
					if (R->h[F_DONE]) break;
					if (R->h[F_SURE] && lists_too_short()) break;
					if (rounds >= (uint32_t) max_rounds)
					{
						// the guesses keep failing: this window's entries as one chain behind the final ones
						R->back_bound = 0xffffffffu;
						CK(cudaMemsetAsync(R->control() + F_DONE, 0, 4, ctx->stream));
						{
							dp_args G = run_dp_args(R, &w);
							G.skip_when_done = 0;
							CK(launch_dp_rebuild(G, ctx->stream));
							dp_args D = G;
							D.bound_on_device = 0; D.reach_bound = 0xffffffffu;
							D.chunk = (R->n_entries + 63) / 64 * 64; D.n_chunks = 1;
							CK(launch_dp_chains(D, ctx->stream));
							ctx->stats["dp_fallbacks"] += 1;
						}
						if (last && (rc = run_enqueue_final(R, false))) return rc;
						if ((rc = run_read_control(R))) return rc;
						++rounds;
						break;
					}
					batch = std::min <uint32_t>(8, rounds);
				}
				total_rounds += rounds;
				if (lists_too_short()) { overflow = true; resume_block = w.b0; }
			}
			if (overflow)
			{
				if ((rc = grow_cap())) return rc;
				continue;
			}
		}
		if (!lists_too_short()) break;
		if ((rc = grow_cap())) return rc;
		resume_block = 0;
	}
	ctx->learned_cap[std::make_tuple(R->m, R->n, R->L)] = R->cap;      // later calls on this shape start from the bound that worked
	ctx->stats["cap"] = R->cap;
	ctx->stats["block_columns"] = g.C;
	ctx->stats["blocks"] = g.B;
	ctx->stats["warps"] = g.chunked ? g.T / 32 : g.n_warps;
	ctx->stats["chunk"] = g.chunked ? g.E : 0;
	ctx->stats["kernel"] = g.chunked ? 3 : 1;
	ctx->stats["dp_rounds"] += total_rounds;
	ctx->stats["dp_chunks"] = R->n_chunks;
	ctx->stats["dp_reach"] = R->h[F_REACH];
	ctx->stats["exchange_tail_entries"] = R->tails;

	fsq_segmentation *seg = new (std::nothrow) fsq_segmentation;
	if (!seg) return fail(ctx, FSQ_ERR_NOMEM, "out of host memory");
	struct guard { fsq_segmentation *&s; bool ok{false}; ~guard() { if (!ok) { fsq_segmentation_free(s); s = nullptr; } } } gd{seg};
	seg->ctx = ctx; seg->m = m; seg->n = R->n; seg->L = R->L;

	// ---- traceback (reference: segmentation_lp_context.cc:191-224) ----
	bool const whole = R->whole;
	if (!tail_done)
	{
		if ((rc = run_enqueue_traceback(R))) return rc;
		if (whole && (rc = run_enqueue_reduce(R, 0))) return rc;       // harmless when K == m turns out below
		if ((rc = run_read_control(R))) return rc;
	}
	if (R->h[I_TRACE_ERR])
	{
		// a jump longer than the window of entries the ranks export to one another: the traceback over the whole table instead
		R->trace_replicated = true;
		ctx->stats["trace_fallbacks"] += 1;
		CK(cudaMemsetAsync(R->control() + I_TRACE_ERR, 0, 4, ctx->stream));
		if ((rc = run_enqueue_traceback(R))) return rc;
		if (whole && (rc = run_enqueue_reduce(R, 0))) return rc;
		if ((rc = run_read_control(R))) return rc;
		tail_done = false;
	}
	seg->T = R->h[I_T];
	seg->K = R->h[I_K];
	if (ctx->opt("keep_dp", 0))
	{
		uint64_t const dp_size = R->dp_size;
		seg->dp_lb.resize(dp_size); seg->dp_max.resize(dp_size); seg->dp_size.resize(dp_size);
		CK(cudaMemcpyAsync(seg->dp_lb.data(), R->tb_lb(), dp_size * 4, cudaMemcpyDeviceToHost, ctx->stream));
		CK(cudaMemcpyAsync(seg->dp_max.data(), R->tb_max(), dp_size * 4, cudaMemcpyDeviceToHost, ctx->stream));
		CK(cudaMemcpyAsync(seg->dp_size.data(), R->tb_size(), dp_size * 4, cudaMemcpyDeviceToHost, ctx->stream));
		CK(cudaStreamSynchronize(ctx->stream));
	}
	// the traceback stays on the device with the handle (handed over when the run is through with it); it is copied when asked for
	auto hand_over_traceback = [&] { seg->tr_lb.adopt(R->tr_lb); seg->tr_rb.adopt(R->tr_rb); seg->tr_max.adopt(R->tr_max); seg->tr_size.adopt(R->tr_size); };
	if (!(seg->K < m))
	{
		hand_over_traceback();
		// generate_context.cc:192-200; the traceback stays available through the handle for inspection
		gd.ok = true;
		*out = seg;
		return fail(ctx, FSQ_ERR_NOT_REDUCIBLE, "the maximum segment size %u is equal to the number of input sequences", seg->K);
	}

	// ---- greedy reduction (segmentation_lp_context.cc:335-390), then the states at the reduced right bounds
	//      (segmentation_lp_context.cc:229-332, update_pbwt_task.cc:13-35) ----
	uint32_t const pos_first = R->h[I_POS_FIRST], pos_last = R->h[I_POS_LAST];
	if (!tail_done && !whole)
	{
		// the lists are gone (windows): the divergence arrays at the own traceback positions, from a re-scan, serve instead
		R->toplist.release();         // (back to the context's cache, which gives memory back to the device when an allocation needs it)
		host_watch hw(ctx, "reduced_states");
		uint32_t const cnt = pos_last - pos_first;
		std::vector <uint64_t> cols(cnt);
		if (cnt)
		{
			CK(cudaMemcpyAsync(cols.data(), R->tr_rb.as <uint64_t>() + pos_first, (size_t) cnt * 8, cudaMemcpyDeviceToHost, ctx->stream));
			CK(cudaStreamSynchronize(ctx->stream));
		}
		for (auto &c : cols) c -= R->c0;
		CK(R->rows.alloc(ctx, (size_t) R->rows_bound * m * 4)); CK(R->rows_a.alloc(ctx, (size_t) std::max <uint32_t>(cnt, 1) * m * 4));
		if (cnt && (rc = dump_states(ctx, R->msa, R->S, cols, R->rows_a.as <uint32_t>(), R->rows.as <uint32_t>()))) return rc;
		if ((rc = run_enqueue_reduce(R, 1))) return rc;
		if ((rc = run_read_control(R))) return rc;
	}
	while (R->h[I_HALO_ERR])
	{
		// a run of merged segments went past the rows at hand of the ranks behind this one: more rows (every rank decides alike:
		// the flag is merged below)
		if (R->halo >= R->rows_bound) return fail(ctx, FSQ_ERR_INTERNAL, "the reduction ran out of rows with the complete halo");
		R->halo = std::min <uint32_t>(R->rows_bound, R->halo * 8);
		CK(cudaMemsetAsync(R->control() + I_HALO_ERR, 0, 4, ctx->stream));
		if ((rc = run_enqueue_reduce(R, whole ? 0 : 1))) return rc;
		if ((rc = run_read_control(R))) return rc;
		tail_done = false;
	}
	uint32_t const n_red = R->h[I_NRED], seg_first = R->h[I_SEG_FIRST], seg_last = R->h[I_SEG_LAST];
	uint32_t const own = seg_last - seg_first;
	if (!tail_done)
	{
		if (R->states.p)
		{
			if ((rc = run_enqueue_state_gather(R, own))) return rc;
		}
		else if (!whole)
		{
			CK(R->red_a.alloc(ctx, (size_t) std::max <uint32_t>(own, 1) * m * 4)); CK(R->red_d.alloc(ctx, (size_t) std::max <uint32_t>(own, 1) * m * 4));
			if (own)
			{
				gather_rows_kernel <<<dim3(own, (m + 255) / 256), 256, 0, ctx->stream>>>(R->rows_a.as <uint32_t>(), R->rows.as <uint32_t>(), R->red_pos.as <uint32_t>(), R->control(), m,
					R->red_a.as <uint32_t>(), R->red_d.as <uint32_t>());
				CK(cudaGetLastError());
				ctx->stats["kernel_launches"] += 1;
			}
		}
		else
		{
			// the states were not kept: a re-scan of the blocks that hold the reduced right bounds (pass C)
			host_watch hw(ctx, "reduced_states");
			std::vector <uint64_t> cols(own);
			if (own)
			{
				CK(cudaMemcpyAsync(cols.data(), R->red_rb.as <uint64_t>() + seg_first, (size_t) own * 8, cudaMemcpyDeviceToHost, ctx->stream));
				CK(cudaStreamSynchronize(ctx->stream));
			}
			for (auto &c : cols) c -= R->c0;
			CK(R->red_a.alloc(ctx, (size_t) std::max <uint32_t>(own, 1) * m * 4)); CK(R->red_d.alloc(ctx, (size_t) std::max <uint32_t>(own, 1) * m * 4));
			if (own && (rc = dump_states(ctx, R->msa, R->S, cols, R->red_a.as <uint32_t>(), R->red_d.as <uint32_t>()))) return rc;
		}
	}
	// The reduced bounds stay on the device with the handle and reach the host when asked for -- unless they were small
	// enough to ride along with the control words already.
	seg->n_reduced = n_red;
	if (tail_done && (size_t) R->t_bound * 20 <= ((size_t) 4 << 20))
	{
		// (the pinned block goes with the handle as it is: unpacking it here would cost every call 0.1 - 0.7 ms of host time,
		// for arrays that only some callers ask for)
		seg->h_bounds.adopt(R->h_red);
		seg->h_bounds_stride = R->t_bound;
	}
	seg->d_lb.adopt(R->red_lb); seg->d_rb.adopt(R->red_rb); seg->d_size.adopt(R->red_size);
	seg->own_first = seg_first; seg->own_last = seg_last;
	seg->red_a.adopt(R->red_a); seg->red_d.adopt(R->red_d);
	hand_over_traceback();
	gd.ok = true;
	*out = seg;
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
	for (auto e : ctx->event_pool) cudaEventDestroy(e);
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
	static char const *known[] = {"block_columns", "cap", "keep_dp", "warps", "kernel", "chunk", "stitch_sort", "dp_chunks", "dp_max_rounds", "dp_guess_rounds", "store_states", "whole_waves", "mem_limit", "window_blocks", "halo", "trace_window", "timeline", "exchange_tails", "dp_back_bound", nullptr};
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

// ---- identity columns (remove-identity-columns/main.cc:107-160) --------------------------------------------------------

int fsq_identity_columns(fsq_ctx *ctx, const fsq_msa *msa, uint8_t *identity, uint64_t *n_identity, fsq_msa **without)
{
	int rc = check_shape(ctx, msa);
	if (rc) return rc;
	if (!identity && !n_identity && !without) return fail(ctx, FSQ_ERR_ARG, "nothing asked for");
	CK(cudaSetDevice(ctx->device));
	uint64_t const n = msa->n;
	dbuf d_id, d_off, d_total;
	uint32_t const n_blocks = (uint32_t) ((n + 1023) / 1024);
	CK(d_id.alloc(ctx, n)); CK(d_off.alloc(ctx, (size_t) n_blocks * 4)); CK(d_total.alloc(ctx, 8));
	phase pi(ctx, "identity_columns");
	CK(launch_identity_columns(msa->cols, msa->col_stride, msa->m, n, d_id.as <uint8_t>(), ctx->stream));
	CK(launch_kept_offsets(d_id.as <uint8_t>(), n, d_off.as <uint32_t>(), d_total.as <uint64_t>(), ctx->stream));
	ctx->stats["kernel_launches"] += 3;
	uint64_t kept = 0;
	CK(cudaMemcpyAsync(&kept, d_total.p, 8, cudaMemcpyDeviceToHost, ctx->stream));
	if (identity) CK(cudaMemcpyAsync(identity, d_id.p, n, cudaMemcpyDeviceToHost, ctx->stream));
	CK(cudaStreamSynchronize(ctx->stream));
	if (n_identity) *n_identity = n - kept;
	if (without)
	{
		*without = nullptr;
		if (0 == kept) { pi.stop(); return fail(ctx, FSQ_ERR_ARG, "every column is an identity column: nothing is left"); }
		fsq_msa *out = nullptr;
		if ((rc = new_msa(ctx, msa->m, kept, 0, msa->nsym, &out))) return rc;
		out->sym_mask = msa->sym_mask;
		cudaError_t e = launch_compact_columns(msa->cols, msa->col_stride, n, d_id.as <uint8_t>(), d_off.as <uint32_t>(), out->cols, ctx->stream);
		if (cudaSuccess == e) e = cudaStreamSynchronize(ctx->stream);
		if (cudaSuccess != e) { fsq_msa_free(out); return fail(ctx, FSQ_ERR_CUDA, "column compaction: %s", cudaGetErrorString(e)); }
		ctx->stats["kernel_launches"] += 1;
		*without = out;
	}
	pi.stop();
	return FSQ_OK;
}

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

int fsq_segment_sharded(fsq_ctx *ctx, const fsq_msa *msa, fsq_comm *comm, uint64_t n_total, uint64_t L, fsq_segmentation **out)
{
	int rc = check_shape(ctx, msa);
	if (rc) return rc;
	if (!out) return fail(ctx, FSQ_ERR_ARG, "null out pointer");
	*out = nullptr;
	if (0 == L) return fail(ctx, FSQ_ERR_ARG, "segment length bound must be positive");
	if (n_total >= (1ull << 31)) return fail(ctx, FSQ_ERR_UNSUPPORTED, "n = %llu columns: positions are kept in 32 bits (n < 2^31)", (unsigned long long) n_total);
	if (n_total < 2 * L) return fail(ctx, FSQ_ERR_SHORT, "n = %llu < 2L = %llu: short path", (unsigned long long) n_total, (unsigned long long) (2 * L));
	CK(cudaSetDevice(ctx->device));
	// (the phase timers of earlier calls are read when somebody asks for a statistic, not here: twenty event pairs take
	// 0.2 ms of host time during which the device would sit idle; the offsets from the call's start, "at_<phase>", are kept
	// for the call before a read only)
	if (ctx->pending.size() > 4096) flush_phases(ctx);
	else ctx->call_first_pending = ctx->pending.size();
	if (!ctx->call_start) CK(cudaEventCreate(&ctx->call_start));
	CK(cudaEventRecord(ctx->call_start, ctx->stream));
	auto const t_entry = std::chrono::steady_clock::now();
	{
		fsq_run run;
		if ((rc = run_init(&run, ctx, msa, comm && comm->size > 1 ? comm : nullptr, n_total, L))) return rc;
		rc = run_segment(&run, out);
		// the run's buffers go back to the cache when `run` dies: nothing may still be using them
		cudaStreamSynchronize(ctx->stream);
		ctx->stats["hms_after"] += std::chrono::duration <double, std::milli>(std::chrono::steady_clock::now() - run.t_mark).count();   // since the last read of the control words
	}
	ctx->stats["hms_call"] += std::chrono::duration <double, std::milli>(std::chrono::steady_clock::now() - t_entry).count();
	return rc;
}

int fsq_segment(fsq_ctx *ctx, const fsq_msa *msa, uint64_t L, fsq_segmentation **out)
{
	if (msa && 0 != msa->col0) return fail(ctx, FSQ_ERR_ARG, "fsq_segment needs an alignment that starts at column 0 (column shards go through fsq_segment_sharded)");
	return fsq_segment_sharded(ctx, msa, nullptr, msa ? msa->n : 0, L, out);
}

void fsq_segmentation_free(fsq_segmentation *seg)
{
	// the buffers go back to the context's workspace cache (the context must outlive its segmentations)
	delete seg;
}

uint32_t fsq_segmentation_count(const fsq_segmentation *seg) { return seg ? seg->n_reduced : 0; }
uint32_t fsq_segmentation_max_segment_size(const fsq_segmentation *seg) { return seg ? seg->K : 0; }

int fsq_segmentation_own_range(const fsq_segmentation *seg, uint32_t *first, uint32_t *last)
{
	if (!seg || !first || !last) return FSQ_ERR_ARG;
	*first = seg->own_first;
	*last = seg->own_last;
	return FSQ_OK;
}

static int fetch_bounds(fsq_segmentation *seg)
{
	fsq_ctx *ctx = seg->ctx;
	if (seg->bounds_on_host) return FSQ_OK;
	uint32_t const n = seg->n_reduced;
	if (seg->h_bounds.p)
	{
		uint64_t const *h_lb = seg->h_bounds.as <uint64_t>(), *h_rb = h_lb + seg->h_bounds_stride;
		uint32_t const *h_sz = (uint32_t const *) (h_rb + seg->h_bounds_stride);
		seg->lb.assign(h_lb, h_lb + n);
		seg->rb.assign(h_rb, h_rb + n);
		seg->size.assign(h_sz, h_sz + n);
		seg->h_bounds.release();
		seg->bounds_on_host = true;
		return FSQ_OK;
	}
	seg->lb.resize(n); seg->rb.resize(n); seg->size.resize(n);
	if (n)
	{
		CK(cudaSetDevice(ctx->device));
		CK(cudaMemcpyAsync(seg->lb.data(), seg->d_lb.p, (size_t) n * 8, cudaMemcpyDeviceToHost, ctx->stream));
		CK(cudaMemcpyAsync(seg->rb.data(), seg->d_rb.p, (size_t) n * 8, cudaMemcpyDeviceToHost, ctx->stream));
		CK(cudaMemcpyAsync(seg->size.data(), seg->d_size.p, (size_t) n * 4, cudaMemcpyDeviceToHost, ctx->stream));
		CK(cudaStreamSynchronize(ctx->stream));
	}
	seg->bounds_on_host = true;
	return FSQ_OK;
}

int fsq_segmentation_bounds(const fsq_segmentation *cseg, uint64_t *lb, uint64_t *rb, uint32_t *size)
{
	fsq_segmentation *seg = const_cast <fsq_segmentation *>(cseg);
	if (!seg || !lb || !rb || !size) return FSQ_ERR_ARG;
	{
		int const rc = fetch_bounds(seg);
		if (rc) return rc;
	}
	std::copy(seg->lb.begin(), seg->lb.end(), lb);
	std::copy(seg->rb.begin(), seg->rb.end(), rb);
	std::copy(seg->size.begin(), seg->size.end(), size);
	return FSQ_OK;
}

static int fetch_states(fsq_segmentation *seg)
{
	fsq_ctx *ctx = seg->ctx;
	if (seg->h_a.p) return FSQ_OK;
	size_t const bytes = (size_t) (seg->own_last - seg->own_first) * seg->m * 4;
	CK(cudaSetDevice(ctx->device));
	phase pd(ctx, "states_d2h");
	CK(seg->h_a.alloc(ctx, bytes)); CK(seg->h_d.alloc(ctx, bytes));
	if (bytes)
	{
		CK(cudaMemcpyAsync(seg->h_a.p, seg->red_a.p, bytes, cudaMemcpyDeviceToHost, ctx->stream));
		CK(cudaMemcpyAsync(seg->h_d.p, seg->red_d.p, bytes, cudaMemcpyDeviceToHost, ctx->stream));
	}
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
	size_t const bytes = (size_t) (seg->own_last - seg->own_first) * seg->m * 4;
	memcpy(a, seg->h_a.p, bytes);
	memcpy(d, seg->h_d.p, bytes);
	return FSQ_OK;
}

int fsq_segmentation_states_range(const fsq_segmentation *cseg, uint32_t first, uint32_t count, uint32_t *a, uint32_t *d)
{
	fsq_segmentation *seg = const_cast <fsq_segmentation *>(cseg);
	if (!seg || !a || !d) return FSQ_ERR_ARG;
	fsq_ctx *ctx = seg->ctx;
	if (first < seg->own_first || (uint64_t) first + count > seg->own_last)
		return fail(ctx, FSQ_ERR_ARG, "segments [%u, %u) are not among the ones whose state this handle holds: [%u, %u)", first, first + count, seg->own_first, seg->own_last);
	if (0 == count) return FSQ_OK;
	CK(cudaSetDevice(ctx->device));
	size_t const at = (size_t) (first - seg->own_first) * seg->m, bytes = (size_t) count * seg->m * 4;
	CK(cudaMemcpyAsync(a, seg->red_a.as <uint32_t>() + at, bytes, cudaMemcpyDeviceToHost, ctx->stream));
	CK(cudaMemcpyAsync(d, seg->red_d.as <uint32_t>() + at, bytes, cudaMemcpyDeviceToHost, ctx->stream));
	CK(cudaStreamSynchronize(ctx->stream));
	return FSQ_OK;
}

int fsq_segmentation_states_view(const fsq_segmentation *cseg, const uint32_t **a, const uint32_t **d)
{
	fsq_segmentation *seg = const_cast <fsq_segmentation *>(cseg);
	if (!seg || !a || !d) return FSQ_ERR_ARG;
	int const rc = fetch_states(seg);
	if (rc) return rc;
	*a = seg->h_a.as <uint32_t>();
	*d = seg->h_d.as <uint32_t>();
	return FSQ_OK;
}

uint32_t fsq_segmentation_traceback_count(const fsq_segmentation *seg) { return seg ? seg->T : 0; }

// rb u64[T], then lb, max, size u32[T] each
int fsq_segmentation_traceback(const fsq_segmentation *cseg, uint64_t *lb, uint64_t *rb, uint32_t *max_size, uint32_t *size)
{
	fsq_segmentation *seg = const_cast <fsq_segmentation *>(cseg);
	if (!seg || !lb || !rb || !max_size || !size) return FSQ_ERR_ARG;
	fsq_ctx *ctx = seg->ctx;
	uint32_t const T = seg->T;
	if (0 == T || !seg->tr_rb.p) return FSQ_OK;
	if (!seg->h_tb.p)
	{
		CK(cudaSetDevice(ctx->device));
		CK(seg->h_tb.alloc(ctx, (size_t) T * 20));
		uint64_t *h_rb = seg->h_tb.as <uint64_t>();
		uint32_t *h_lb = (uint32_t *) (h_rb + T), *h_max = h_lb + T, *h_size = h_max + T;
		CK(cudaMemcpyAsync(h_rb, seg->tr_rb.p, (size_t) T * 8, cudaMemcpyDeviceToHost, ctx->stream));
		CK(cudaMemcpyAsync(h_lb, seg->tr_lb.p, (size_t) T * 4, cudaMemcpyDeviceToHost, ctx->stream));
		CK(cudaMemcpyAsync(h_max, seg->tr_max.p, (size_t) T * 4, cudaMemcpyDeviceToHost, ctx->stream));
		CK(cudaMemcpyAsync(h_size, seg->tr_size.p, (size_t) T * 4, cudaMemcpyDeviceToHost, ctx->stream));
		CK(cudaStreamSynchronize(ctx->stream));
	}
	uint64_t const *h_rb = seg->h_tb.as <uint64_t>();
	uint32_t const *h_lb = reinterpret_cast <uint32_t const *>(h_rb + T), *h_max = h_lb + T, *h_size = h_max + T;
	std::copy(h_lb, h_lb + T, lb);
	std::copy(h_rb, h_rb + T, rb);
	std::copy(h_max, h_max + T, max_size);
	std::copy(h_size, h_size + T, size);
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
	if (s < seg->own_first || s >= seg->own_last) return fail(ctx, FSQ_ERR_ARG, "segment %u is not among the ones whose state this handle holds: [%u, %u)", s, seg->own_first, seg->own_last);
	CK(cudaSetDevice(ctx->device));
	{
		int const rcb = fetch_bounds(const_cast <fsq_segmentation *>(seg));
		if (rcb) return rcb;
	}
	uint32_t const m = seg->m;
	dbuf off, ids, cls, nruns;
	size_t const at = (size_t) (s - seg->own_first) * m;
	int rc = sets_on_device(ctx, seg->red_a.as <uint32_t>() + at, seg->red_d.as <uint32_t>() + at, m, seg->lb[s], off, ids, cls, nruns);
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
	if (s < seg->own_first || (size_t) s + 1 >= seg->own_last) return fail(ctx, FSQ_ERR_ARG, "segment pair (%u, %u) is not among the segments whose state this handle holds: [%u, %u)", s, s + 1, seg->own_first, seg->own_last);
	CK(cudaSetDevice(ctx->device));
	uint32_t const m = seg->m;
	dbuf off_l, ids_l, cls_l, n_l, off_r, ids_r, cls_r, n_r, dW;
	int rc;
	if ((rc = fetch_bounds(const_cast <fsq_segmentation *>(seg)))) return rc;
	size_t const at = (size_t) (s - seg->own_first) * m;
	uint32_t const *sa = seg->red_a.as <uint32_t>() + at, *sd = seg->red_d.as <uint32_t>() + at;
	if ((rc = sets_on_device(ctx, sa, sd, m, seg->lb[s], off_l, ids_l, cls_l, n_l))) return rc;
	if ((rc = sets_on_device(ctx, sa + m, sd + m, m, seg->lb[s + 1], off_r, ids_r, cls_r, n_r))) return rc;
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

int fsq_pair_counts(fsq_ctx *ctx, const fsq_segmentation *seg, uint32_t s, uint32_t *lhs, uint32_t *rhs, uint32_t *count, uint32_t *n_pairs, uint32_t *runs_lhs, uint32_t *runs_rhs)
{
	if (!ctx || !seg || !lhs || !rhs || !count || !n_pairs) return fail(ctx, FSQ_ERR_ARG, "null argument");
	if (s < seg->own_first || (size_t) s + 1 >= seg->own_last) return fail(ctx, FSQ_ERR_ARG, "segment pair (%u, %u) is not among the segments whose state this handle holds: [%u, %u)", s, s + 1, seg->own_first, seg->own_last);
	CK(cudaSetDevice(ctx->device));
	uint32_t const m = seg->m;
	dbuf off_l, ids_l, cls_l, n_l, off_r, ids_r, cls_r, n_r, dW, d_lhs, d_rhs, d_cnt, d_n;
	int rc;
	if ((rc = fetch_bounds(const_cast <fsq_segmentation *>(seg)))) return rc;
	size_t const at = (size_t) (s - seg->own_first) * m;
	uint32_t const *sa = seg->red_a.as <uint32_t>() + at, *sd = seg->red_d.as <uint32_t>() + at;
	if ((rc = sets_on_device(ctx, sa, sd, m, seg->lb[s], off_l, ids_l, cls_l, n_l))) return rc;
	if ((rc = sets_on_device(ctx, sa + m, sd + m, m, seg->lb[s + 1], off_r, ids_r, cls_r, n_r))) return rc;
	uint32_t rl = 0, rr = 0;
	CK(cudaMemcpyAsync(&rl, n_l.p, 4, cudaMemcpyDeviceToHost, ctx->stream));
	CK(cudaMemcpyAsync(&rr, n_r.p, 4, cudaMemcpyDeviceToHost, ctx->stream));
	CK(cudaStreamSynchronize(ctx->stream));
	size_t const bytes = (size_t) rl * rr * 4;
	CK(dW.alloc(ctx, bytes)); CK(d_lhs.alloc(ctx, (size_t) m * 4)); CK(d_rhs.alloc(ctx, (size_t) m * 4)); CK(d_cnt.alloc(ctx, (size_t) m * 4)); CK(d_n.alloc(ctx, 4));
	CK(cudaMemsetAsync(dW.p, 0, bytes, ctx->stream));
	CK(launch_pair_weights(cls_l.as <uint32_t>(), cls_r.as <uint32_t>(), m, rr, dW.as <int32_t>(), ctx->stream));
	CK(launch_pair_counts(dW.as <int32_t>(), rl, rr, d_lhs.as <uint32_t>(), d_rhs.as <uint32_t>(), d_cnt.as <uint32_t>(), d_n.as <uint32_t>(), ctx->stream));
	ctx->stats["kernel_launches"] += 2;
	uint32_t np = 0;
	CK(cudaMemcpyAsync(&np, d_n.p, 4, cudaMemcpyDeviceToHost, ctx->stream));
	CK(cudaStreamSynchronize(ctx->stream));
	CK(cudaMemcpyAsync(lhs, d_lhs.p, (size_t) np * 4, cudaMemcpyDeviceToHost, ctx->stream));
	CK(cudaMemcpyAsync(rhs, d_rhs.p, (size_t) np * 4, cudaMemcpyDeviceToHost, ctx->stream));
	CK(cudaMemcpyAsync(count, d_cnt.p, (size_t) np * 4, cudaMemcpyDeviceToHost, ctx->stream));
	CK(cudaStreamSynchronize(ctx->stream));
	*n_pairs = np;
	if (runs_lhs) *runs_lhs = rl;
	if (runs_rhs) *runs_rhs = rr;
	return FSQ_OK;
}

// ---- match_founder_sequences (match-sequences-to-founders/match_founder_sequences.cc:127-214) -----------------------------

int fsq_match_founders(fsq_ctx *ctx, const fsq_msa *sequences, const fsq_msa *founders, uint64_t min_segment_length,
	uint32_t *seq_idx, uint64_t *lb, uint64_t *rb, uint32_t *founder_bits, uint64_t capacity, uint64_t *count)
{
	int rc = check_shape(ctx, sequences);
	if (rc) return rc;
	if ((rc = check_shape(ctx, founders))) return rc;
	if (!count) return fail(ctx, FSQ_ERR_ARG, "null argument");
	if (sequences->n != founders->n) return fail(ctx, FSQ_ERR_ARG, "sequences have %llu columns, founders %llu", (unsigned long long) sequences->n, (unsigned long long) founders->n);
	uint32_t const K = founders->m, m = sequences->m, W = (K + 31) / 32;
	if (K > 8192) return fail(ctx, FSQ_ERR_UNSUPPORTED, "%u founders: the matching kernel holds 8,192", K);
	uint64_t const n = sequences->n;
	uint32_t const nsym = std::max(sequences->nsym, founders->nsym);
	CK(cudaSetDevice(ctx->device));
	dbuf masks, row_count, row_off, d_seq, d_lb, d_rb, d_bits;
	if (cudaSuccess != masks.alloc(ctx, (size_t) n * nsym * W * 4)) { cudaGetLastError(); return fail(ctx, FSQ_ERR_NOMEM, "%llu bytes of founder masks do not fit", (unsigned long long) n * nsym * W * 4); }
	CK(row_count.alloc(ctx, (size_t) m * 8)); CK(row_off.alloc(ctx, ((size_t) m + 1) * 8));
	phase pm(ctx, "match_founders");
	CK(launch_founder_masks(founders->cols, founders->col_stride, K, n, nsym, W, masks.as <uint32_t>(), ctx->stream));
	CK(launch_match_founders(sequences->cols, sequences->col_stride, m, n, masks.as <uint32_t>(), nsym, K, W, min_segment_length, 0,
		row_count.as <uint64_t>(), row_off.as <uint64_t>(), nullptr, nullptr, nullptr, nullptr, ctx->stream));
	uint64_t total = 0;
	CK(cudaMemcpyAsync(&total, row_off.as <uint64_t>() + m, 8, cudaMemcpyDeviceToHost, ctx->stream));
	CK(cudaStreamSynchronize(ctx->stream));
	ctx->stats["kernel_launches"] += 3;
	*count = total;
	if (!seq_idx || !lb || !rb || !founder_bits || capacity < total) { pm.stop(); return FSQ_OK; }       // the caller learns the size and calls again
	CK(d_seq.alloc(ctx, (size_t) std::max <uint64_t>(total, 1) * 4)); CK(d_lb.alloc(ctx, (size_t) std::max <uint64_t>(total, 1) * 8));
	CK(d_rb.alloc(ctx, (size_t) std::max <uint64_t>(total, 1) * 8)); CK(d_bits.alloc(ctx, (size_t) std::max <uint64_t>(total, 1) * W * 4));
	CK(launch_match_founders(sequences->cols, sequences->col_stride, m, n, masks.as <uint32_t>(), nsym, K, W, min_segment_length, 1,
		row_count.as <uint64_t>(), row_off.as <uint64_t>(), d_seq.as <uint32_t>(), d_lb.as <uint64_t>(), d_rb.as <uint64_t>(), d_bits.as <uint32_t>(), ctx->stream));
	ctx->stats["kernel_launches"] += 1;
	pm.stop();
	CK(cudaMemcpyAsync(seq_idx, d_seq.p, (size_t) total * 4, cudaMemcpyDeviceToHost, ctx->stream));
	CK(cudaMemcpyAsync(lb, d_lb.p, (size_t) total * 8, cudaMemcpyDeviceToHost, ctx->stream));
	CK(cudaMemcpyAsync(rb, d_rb.p, (size_t) total * 8, cudaMemcpyDeviceToHost, ctx->stream));
	CK(cudaMemcpyAsync(founder_bits, d_bits.p, (size_t) total * W * 4, cudaMemcpyDeviceToHost, ctx->stream));
	CK(cudaStreamSynchronize(ctx->stream));
	return FSQ_OK;
}

// ---- column-sharded operation -------------------------------------------------------------------------------------------
// One process (or thread) per GPU; rank r holds the columns fsq_shard_columns() names of one alignment and calls
// fsq_segment_sharded() with its communicator.  The exchanges are all-gathers on the context's stream (fsq_comm.cu).

int fsq_shard_columns(uint64_t n_total, int size, int rank, uint64_t *c0, uint64_t *c1)
{
	if (!c0 || !c1 || size < 1 || rank < 0 || rank >= size) return FSQ_ERR_ARG;
	uint64_t const per = columns_per_rank(n_total, size);
	*c0 = std::min <uint64_t>(n_total, (uint64_t) rank * per);
	*c1 = std::min <uint64_t>(n_total, *c0 + per);
	return *c0 < *c1 ? FSQ_OK : FSQ_ERR_ARG;
}

int fsq_comm_unique_id(uint8_t id[FSQ_COMM_ID_BYTES])
{
	if (!id) return FSQ_ERR_ARG;
	std::string err;
	if (!nccl_unique_id(id, &err)) return fail(nullptr, FSQ_ERR_CUDA, "%s", err.c_str());
	return FSQ_OK;
}

int fsq_comm_create(fsq_ctx *ctx, const uint8_t id[FSQ_COMM_ID_BYTES], int rank, int size, fsq_comm **out)
{
	if (!ctx || !id || !out || size < 1 || rank < 0 || rank >= size) return fail(ctx, FSQ_ERR_ARG, "bad argument");
	CK(cudaSetDevice(ctx->device));
	std::string err;
	fsq_comm *c = nccl_comm_create(id, rank, size, &err);
	if (!c) return fail(ctx, FSQ_ERR_CUDA, "%s", err.c_str());
	*out = c;
	return FSQ_OK;
}

int fsq_comm_adopt_nccl(fsq_ctx *ctx, void *nccl_comm, int rank, int size, fsq_comm **out)
{
	if (!ctx || !nccl_comm || !out || size < 1 || rank < 0 || rank >= size) return fail(ctx, FSQ_ERR_ARG, "bad argument");
	std::string err;
	fsq_comm *c = nccl_comm_adopt(nccl_comm, rank, size, &err);
	if (!c) return fail(ctx, FSQ_ERR_CUDA, "%s", err.c_str());
	*out = c;
	return FSQ_OK;
}

int fsq_local_group_create(int size, fsq_local_group **out)
{
	if (!out || size < 1) return FSQ_ERR_ARG;
	*out = local_group_create(size);
	return *out ? FSQ_OK : FSQ_ERR_NOMEM;
}

void fsq_local_group_free(fsq_local_group *g) { local_group_free(g); }

int fsq_comm_create_local(fsq_ctx *ctx, fsq_local_group *g, int rank, fsq_comm **out)
{
	if (!ctx || !g || !out) return fail(ctx, FSQ_ERR_ARG, "null argument");
	fsq_comm *c = local_comm_create(g, rank);
	if (!c) return fail(ctx, FSQ_ERR_ARG, "rank %d is outside the group", rank);
	*out = c;
	return FSQ_OK;
}

void fsq_comm_free(fsq_comm *c) { delete c; }

int fsq_comm_rank(const fsq_comm *c) { return c ? c->rank : 0; }
int fsq_comm_size(const fsq_comm *c) { return c ? c->size : 1; }

} // extern "C"
