/*
 * TEST INFRASTRUCTURE — NOT PRODUCT CODE.
 *
 * oracle/libfsq_oracle.so: a plain-C, single-threaded, streaming restatement of the reference's
 * PBWT + segmentation path (tsnorri/founder-sequences), used only as the checker by tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.  The product
 * (founder-sequences_b200/) never loads it.
 *
 * Parity pinning: the reference ships no tests or golden vectors for this path, so
 * this restatement is pinned against the reference's own code compiled in place
 * (oracle/_ref/libfsq_ref.so, see ref_harness.cc) by tests/test_oracle_vs_ref.py, and against the
 * fixtures that harness generated (tests/golden/, made by tests/golden/make_golden.py).
 *
 * Every function cites the reference lines it follows (paths relative to the reference root).
 */

#define _POSIX_C_SOURCE 200809L
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include <math.h>

typedef struct
{
	uint64_t lb, rb;
	uint32_t max, size;
} dp_arg; /* include/founder_sequences/segmentation_dp_arg.hh:18-56 */

typedef struct
{
	uint32_t m;
	uint64_t n;
	uint8_t *rows; /* row-major copy */
	uint8_t comp[256];
	uint32_t sigma; /* number of comps incl. NUL = 0 */

	/* results */
	uint64_t L;
	dp_arg *dp; uint64_t dp_size;
	dp_arg *tb; uint64_t tb_size;
	uint32_t K;
	uint32_t *tb_a, *tb_d;
	dp_arg *red; uint64_t red_size;
	uint32_t *red_a, *red_d;
	double first_pass_seconds;
} oracle;


/* ---- alphabet: lib/libbio/include/libbio/consecutive_alphabet.hh:259-264, 309-326.
 * NUL is always comp 0; the bytes present get comps 1.. in byte order. */
static void build_alphabet(oracle *o)
{
	uint8_t seen[256];
	memset(seen, 0, sizeof seen);
	size_t const total = (size_t) o->m * o->n;
	for (size_t i = 0; i < total; ++i)
		seen[o->rows[i]] = 1;
	seen[0] = 1;
	uint32_t next = 0;
	for (unsigned c = 0; c < 256; ++c)
		o->comp[c] = seen[c] ? (uint8_t) next++ : 0;
	o->sigma = next;
}


/* ---- PBWT state ---- */
typedef struct
{
	uint32_t m;
	uint64_t n;
	uint64_t k; /* sequence_idx: next column to process */
	uint32_t *a, *d, *a2, *d2;
	/* histogram as a value-indexed doubly linked list, lib/libbio/include/libbio/array_list.hh */
	int with_hist;
	uint32_t *cnt; /* n + 2 */
	uint64_t *nxt, *prv;
	uint64_t head, tail; /* UINT64_MAX = none */
} pbwt;

#define NONE UINT64_MAX

/* lib/libbio/include/libbio/pbwt_context.hh:259-268 */
static void pbwt_init(pbwt *p, uint32_t m, uint64_t n, int with_hist)
{
	p->m = m; p->n = n; p->k = 0; p->with_hist = with_hist;
	p->a = malloc(sizeof(uint32_t) * m); p->d = malloc(sizeof(uint32_t) * m);
	p->a2 = malloc(sizeof(uint32_t) * m); p->d2 = malloc(sizeof(uint32_t) * m);
	for (uint32_t i = 0; i < m; ++i) { p->a[i] = i; p->d[i] = 0; }
	p->cnt = NULL; p->nxt = p->prv = NULL; p->head = p->tail = NONE;
	if (with_hist)
	{
		p->cnt = calloc(n + 2, sizeof(uint32_t));
		p->nxt = malloc(sizeof(uint64_t) * (n + 2));
		p->prv = malloc(sizeof(uint64_t) * (n + 2));
		p->cnt[0] = m; p->nxt[0] = NONE; p->prv[0] = NONE; p->head = p->tail = 0;
	}
}

static void pbwt_free(pbwt *p)
{
	free(p->a); free(p->d); free(p->a2); free(p->d2); free(p->cnt); free(p->nxt); free(p->prv);
}

/* One column: lib/libbio/include/libbio/pbwt.hh:66-142 (stable counting sort + divergence) and
 * :157-199 (histogram).  The SDSL range-max over d_k[prev+1 .. i] is restated as per-symbol running
 * maxima; only the returned value matters (pbwt_context.hh:38-43). */
static void pbwt_step(pbwt *p, oracle const *o)
{
	uint32_t const m = p->m;
	uint64_t const k = p->k;
	uint32_t counts[257];
	uint32_t runmax[256];
	uint8_t seen[256];
	uint32_t const sigma = o->sigma;
	memset(counts, 0, sizeof(uint32_t) * (sigma + 1));
	memset(seen, 0, sigma);
	uint8_t const *rows = o->rows;
	uint64_t const n = o->n;

	for (uint32_t i = 0; i < m; ++i)
		++counts[o->comp[rows[(size_t) p->a[i] * n + k]]];
	{
		uint32_t sum = 0;
		for (uint32_t c = 0; c < sigma; ++c) { uint32_t t = counts[c]; counts[c] = sum; sum += t; }
	}
	for (uint32_t i = 0; i < m; ++i)
	{
		uint32_t const x = p->a[i];
		uint32_t const c = o->comp[rows[(size_t) x * n + k]];
		uint32_t const di = p->d[i];
		/* every symbol's running max since its last occurrence absorbs d_k[i] */
		for (uint32_t q = 0; q < sigma; ++q)
			if (seen[q] && runmax[q] < di) runmax[q] = di;
		uint32_t const dst = counts[c]++;
		p->a2[dst] = x;
		if (!seen[c])
			p->d2[dst] = (uint32_t) (k + 1);
		else
			p->d2[dst] = runmax[c];
		seen[c] = 1;
		runmax[c] = 0;
	}

	if (p->with_hist)
	{
		/* pbwt.hh:163-181: add output values; a value not yet listed is the new one, k+1, linked at the tail */
		for (uint32_t i = 0; i < m; ++i)
		{
			uint64_t const v = p->d2[i];
			if (0 == p->cnt[v])
			{
				p->prv[v] = p->tail; p->nxt[v] = NONE;
				if (NONE != p->tail) p->nxt[p->tail] = v; else p->head = v;
				p->tail = v;
			}
			++p->cnt[v];
		}
		/* pbwt.hh:184-198: remove input values */
		for (uint32_t i = 0; i < m; ++i)
		{
			uint64_t const v = p->d[i];
			if (0 == --p->cnt[v])
			{
				uint64_t const pr = p->prv[v], nx = p->nxt[v];
				if (NONE != pr) p->nxt[pr] = nx; else p->head = nx;
				if (NONE != nx) p->prv[nx] = pr; else p->tail = pr;
			}
		}
	}

	{ uint32_t *t = p->a; p->a = p->a2; p->a2 = t; }
	{ uint32_t *t = p->d; p->d = p->d2; p->d2 = t; }
	++p->k;
}


/* ---- include/founder_sequences/rmq.hh:61-118, block size 64, values compared by .max with std::less ---- */
typedef struct
{
	dp_arg const *vals;
	uint64_t **lev; uint64_t *lev_size; uint64_t *lev_cap; uint32_t n_lev;
} rmq64;

static uint64_t naive_min(rmq64 const *r, uint64_t first, uint64_t last)
{
	/* std::min_element: leftmost minimal element (rmq.hh:107-118) */
	uint64_t best = first;
	for (uint64_t i = first + 1; i < last; ++i)
		if (r->vals[i].max < r->vals[best].max) best = i;
	return best;
}

static void rmq_push(rmq64 *r, uint32_t level, uint64_t v)
{
	if (r->n_lev <= level)
	{
		r->lev = realloc(r->lev, sizeof(uint64_t *) * (level + 1));
		r->lev_size = realloc(r->lev_size, sizeof(uint64_t) * (level + 1));
		r->lev_cap = realloc(r->lev_cap, sizeof(uint64_t) * (level + 1));
		for (uint32_t l = r->n_lev; l <= level; ++l) { r->lev[l] = NULL; r->lev_size[l] = 0; r->lev_cap[l] = 0; }
		r->n_lev = level + 1;
	}
	if (r->lev_size[level] == r->lev_cap[level])
	{
		r->lev_cap[level] = r->lev_cap[level] ? 2 * r->lev_cap[level] : 64;
		r->lev[level] = realloc(r->lev[level], sizeof(uint64_t) * r->lev_cap[level]);
	}
	r->lev[level][r->lev_size[level]++] = v;
}

/* rmq.hh:61-84, including the same-cell double read at :76-77 (smp1 and smp2 use the same index). */
static void rmq_update(rmq64 *r, uint64_t last_idx)
{
	if (((1 + last_idx) & 63) != 0) return;
	uint64_t const bnum = 1 + last_idx / 64;
	uint64_t const new_smp = naive_min(r, (bnum - 1) * 64, bnum * 64);
	uint32_t const new_val = r->vals[new_smp].max;
	rmq_push(r, 0, new_smp);
	for (uint32_t pow2 = 1; (1ull << pow2) <= bnum; ++pow2)
	{
		uint64_t const smp1 = r->lev[pow2 - 1][bnum - (1ull << pow2)];
		uint64_t const smp2 = r->lev[pow2 - 1][bnum - ((1ull << pow2) - 1) - 1];
		uint64_t const smp = (r->vals[smp2].max < r->vals[smp1].max) ? smp2 : smp1;
		rmq_push(r, pow2, (new_val < r->vals[smp].max) ? new_smp : smp);
	}
}

/* Test switch: answer every query with the exact leftmost minimum instead.  The reference does NOT do this (SURVEY.md
 * F3); the switch exists so that tests can show an input on which the two differ (tests/golden/rmq_dp_quirk.npz). */
static int g_exact_rmq = 0;
void fsqo_set_exact_rmq(int on) { g_exact_rmq = on; }

/* rmq.hh:87-105 */
static uint64_t rmq_query(rmq64 const *r, uint64_t beg, uint64_t end)
{
	if (g_exact_rmq) return naive_min(r, beg, end);
	uint64_t const beg_block = beg / 64 + 1;
	uint64_t const end_block = end / 64;
	if (beg_block >= end_block)
		return naive_min(r, beg, end);
	uint32_t const pow2 = 63 - (uint32_t) __builtin_clzll(end_block - beg_block);
	uint64_t const smp1 = r->lev[pow2][beg_block];
	uint64_t const smp2 = r->lev[pow2][end_block - (1ull << pow2)];
	uint64_t smp = (r->vals[smp2].max < r->vals[smp1].max) ? smp2 : smp1;
	uint64_t const left = naive_min(r, beg, beg_block * 64);
	smp = (r->vals[left].max < r->vals[smp].max) ? left : smp;
	if (end == end_block * 64)
		return smp;
	uint64_t const right = naive_min(r, end_block * 64, end);
	return (r->vals[right].max < r->vals[smp].max) ? right : smp;
}

static void rmq_free(rmq64 *r)
{
	for (uint32_t l = 0; l < r->n_lev; ++l) free(r->lev[l]);
	free(r->lev); free(r->lev_size); free(r->lev_cap);
	memset(r, 0, sizeof *r);
}


/* ---- founder-sequences/segmentation_lp_context.cc:393-481 ---- */
static void dp_eval(pbwt const *p, dp_arg const *tb, rmq64 const *r, uint32_t m, uint64_t L, uint64_t text_pos, dp_arg *min_arg)
{
	uint64_t it = p->head;
	uint64_t diff = p->cnt[it];
	uint64_t dp_lb = 0, dp_rb = it;
	if (0 == dp_rb)
	{
		uint32_t const sz = (uint32_t) (m - diff);
		if (sz < min_arg->max) { min_arg->lb = 0; min_arg->rb = 1 + text_pos; min_arg->max = sz; min_arg->size = sz; }
		it = p->nxt[it];
		dp_rb = it;
		diff += p->cnt[it];
	}
	it = p->nxt[it];
	while (NONE != it)
	{
		dp_lb = dp_rb;
		dp_rb = it;
		uint64_t dp_rb_c = dp_rb;
		int ok = 1;
		if (text_pos + 2 - L < dp_rb_c) dp_rb_c = text_pos + 2 - L;
		if (dp_lb < L)
		{
			if (L < dp_rb_c) dp_lb = L; else ok = 0;
		}
		if (ok && dp_lb < dp_rb_c)
		{
			uint64_t const idx = rmq_query(r, dp_lb - L, dp_rb_c - L);
			uint32_t const lhs = tb[idx].max;
			uint32_t const rhs = (uint32_t) (m - diff);
			uint32_t const mx = lhs < rhs ? rhs : lhs;
			if (mx < min_arg->max) { min_arg->lb = idx + L; min_arg->rb = 1 + text_pos; min_arg->max = mx; min_arg->size = rhs; }
		}
		diff += p->cnt[it];
		it = p->nxt[it];
	}
}


static double now_s(void)
{
	struct timespec ts;
	clock_gettime(CLOCK_MONOTONIC, &ts);
	return ts.tv_sec + 1e-9 * ts.tv_nsec;
}


/* ---- public API ---- */

void *fsqo_create(uint8_t const *rows, uint32_t m, uint64_t n)
{
	oracle *o = calloc(1, sizeof *o);
	o->m = m; o->n = n;
	o->rows = malloc((size_t) m * n);
	memcpy(o->rows, rows, (size_t) m * n);
	build_alphabet(o);
	return o;
}

static void free_results(oracle *o)
{
	free(o->dp); free(o->tb); free(o->tb_a); free(o->tb_d); free(o->red); free(o->red_a); free(o->red_d);
	o->dp = o->tb = o->red = NULL; o->tb_a = o->tb_d = o->red_a = o->red_d = NULL;
	o->dp_size = o->tb_size = o->red_size = 0;
}

void fsqo_destroy(void *hp)
{
	oracle *o = hp;
	free_results(o);
	free(o->rows);
	free(o);
}

void fsqo_alphabet(void *hp, uint8_t *comp256) { memcpy(comp256, ((oracle *) hp)->comp, 256); }

long long fsqo_scan_dump(void *hp, uint64_t ncols, uint32_t *a_out, uint32_t *d_out,
	uint64_t *hist_off, uint32_t *hist_v, uint32_t *hist_c, uint64_t hist_cap)
{
	oracle *o = hp;
	pbwt p;
	pbwt_init(&p, o->m, o->n, 1);
	uint64_t pairs = 0; int overflow = 0;
	if (hist_off) hist_off[0] = 0;
	if (ncols > o->n) ncols = o->n;
	for (uint64_t k = 0; k < ncols; ++k)
	{
		pbwt_step(&p, o);
		if (a_out) memcpy(a_out + k * o->m, p.a, sizeof(uint32_t) * o->m);
		if (d_out) memcpy(d_out + k * o->m, p.d, sizeof(uint32_t) * o->m);
		if (hist_off)
		{
			for (uint64_t v = p.head; NONE != v; v = p.nxt[v])
			{
				if (pairs < hist_cap) { hist_v[pairs] = (uint32_t) v; hist_c[pairs] = p.cnt[v]; } else overflow = 1;
				++pairs;
			}
			hist_off[k + 1] = pairs;
		}
	}
	pbwt_free(&p);
	return overflow ? -1 : (long long) pairs;
}

/* First pass over the first ncols columns only (PBWT + histogram + DP), timed.  Returns seconds. */
double fsqo_time_first_pass(void *hp, uint64_t L, uint64_t ncols)
{
	oracle *o = hp;
	uint64_t const n = ncols < o->n ? ncols : o->n;
	if (n < 2 * L) return -1.0;
	pbwt p;
	pbwt_init(&p, o->m, o->n, 1);
	dp_arg *tb = malloc(sizeof(dp_arg) * (n - L + 1));
	rmq64 r; memset(&r, 0, sizeof r); r.vals = tb;
	double const t0 = now_s();
	while (p.k < n - L)
	{
		uint64_t const k = p.k;
		pbwt_step(&p, o);
		if (k + 1 < L) continue;
		dp_arg arg = {0, k + 1, o->m, o->m};
		dp_eval(&p, tb, &r, o->m, L, k, &arg);
		tb[k + 1 - L] = arg;
		rmq_update(&r, k + 1 - L);
	}
	double const t1 = now_s();
	rmq_free(&r); free(tb); pbwt_free(&p);
	return t1 - t0;
}

/* The long path.  founder-sequences/segmentation_lp_context.cc:26-390.
 * Returns 0 on success, 1 if n < 2L, 2 if K == m (generate_context.cc:192-200). */
int fsqo_segment(void *hp, uint64_t L)
{
	oracle *o = hp;
	uint32_t const m = o->m;
	uint64_t const n = o->n;
	if (n < 2 * L) return 1;
	free_results(o);
	o->L = L;

	/* (a, d) checkpoints every `rate` columns instead of the reference's full samples
	 * (pbwt_context.hh:461-468); the re-scan result does not depend on the start state's position. */
	uint64_t rate = (uint64_t) ceil(4.0 * sqrt((double) n));
	if (rate < 1) rate = 1;
	uint64_t const n_ckpt = (n - 1) / rate + 1;	/* checkpoints exist for columns k < n only */
	uint32_t *ck_a = malloc(sizeof(uint32_t) * m * n_ckpt);
	uint32_t *ck_d = malloc(sizeof(uint32_t) * m * n_ckpt);

	double const t0 = now_s();
	pbwt p;
	pbwt_init(&p, m, n, 1);
	o->dp_size = n - L + 1;
	o->dp = malloc(sizeof(dp_arg) * o->dp_size);
	for (uint64_t i = 0; i < o->dp_size; ++i) { o->dp[i].lb = 0; o->dp[i].rb = 0; o->dp[i].max = UINT32_MAX; o->dp[i].size = UINT32_MAX; }
	dp_arg *tb = o->dp;
	rmq64 r; memset(&r, 0, sizeof r); r.vals = tb;

	while (p.k < n)
	{
		uint64_t const k = p.k;
		if (0 == k % rate)
		{
			memcpy(ck_a + (k / rate) * m, p.a, sizeof(uint32_t) * m);
			memcpy(ck_d + (k / rate) * m, p.d, sizeof(uint32_t) * m);
		}
		pbwt_step(&p, o);
		if (k + 1 < L || k >= n - L) continue;			/* :52-58 and :156-163 */
		uint64_t const limit2 = (2 * L < n - L ? 2 * L : n - L) - 1;
		dp_arg arg;
		if (k < limit2)
		{
			/* :72-98 */
			uint32_t diff = 0;
			if (0 == p.head) diff = p.cnt[0];
			arg.lb = 0; arg.rb = k + 1; arg.max = m - diff; arg.size = m - diff;
		}
		else
		{
			/* :113-142 */
			arg.lb = 0; arg.rb = k + 1; arg.max = m; arg.size = m;
			dp_eval(&p, tb, &r, m, L, k, &arg);
		}
		tb[k + 1 - L] = arg;
		rmq_update(&r, k + 1 - L);
	}
	{
		/* :165-185 */
		dp_arg arg = {0, n, m, m};
		dp_eval(&p, tb, &r, m, L, n - 1, &arg);
		tb[o->dp_size - 1] = arg;
	}
	o->first_pass_seconds = now_s() - t0;
	rmq_free(&r);

	/* :191-224 */
	{
		uint64_t cnt = 0, idx = o->dp_size - 1;
		while (1) { ++cnt; if (0 == tb[idx].lb) break; idx = tb[idx].lb - L; }
		o->tb_size = cnt;
		o->tb = malloc(sizeof(dp_arg) * cnt);
		idx = o->dp_size - 1;
		for (uint64_t i = cnt; i-- > 0; ) { o->tb[i] = tb[idx]; if (i) idx = tb[idx].lb - L; }
		o->K = o->tb[cnt - 1].max;
	}
	if (!(o->K < m)) { pbwt_free(&p); free(ck_a); free(ck_d); return 2; }

	/* :229-332, update_pbwt_task.cc:13-35: (a_rb, d_rb) for every traceback entry */
	o->tb_a = malloc(sizeof(uint32_t) * m * o->tb_size);
	o->tb_d = malloc(sizeof(uint32_t) * m * o->tb_size);
	{
		pbwt q;
		pbwt_init(&q, m, n, 0);
		uint64_t cur_ck = NONE;
		for (uint64_t t = 0; t < o->tb_size; ++t)
		{
			uint64_t const rb = o->tb[t].rb;
			uint64_t ck = rb / rate;
			if (ck >= n_ckpt) ck = n_ckpt - 1;
			if (ck != cur_ck || q.k > rb)
			{
				memcpy(q.a, ck_a + ck * m, sizeof(uint32_t) * m);
				memcpy(q.d, ck_d + ck * m, sizeof(uint32_t) * m);
				q.k = ck * rate;
				cur_ck = ck;
			}
			while (q.k < rb) pbwt_step(&q, o);
			memcpy(o->tb_a + t * m, q.a, sizeof(uint32_t) * m);
			memcpy(o->tb_d + t * m, q.d, sizeof(uint32_t) * m);
		}
		pbwt_free(&q);
	}

	/* :335-390; pbwt.hh:206-216 */
	{
		o->red = malloc(sizeof(dp_arg) * o->tb_size);
		o->red_a = malloc(sizeof(uint32_t) * m * o->tb_size);
		o->red_d = malloc(sizeof(uint32_t) * m * o->tb_size);
		uint64_t cur_lb = 0, prev = 0, cnt = 0;
		uint32_t prev_size = o->tb[0].size;
		for (uint64_t t = 1; t < o->tb_size; ++t)
		{
			uint32_t s = 0;
			uint32_t const *d = o->tb_d + t * m;
			for (uint32_t i = 0; i < m; ++i) if (cur_lb < d[i]) ++s;
			if (s <= o->K)
				prev_size = s;
			else
			{
				dp_arg e = {cur_lb, o->tb[prev].rb, UINT32_MAX, prev_size};
				o->red[cnt] = e;
				memcpy(o->red_a + cnt * m, o->tb_a + prev * m, sizeof(uint32_t) * m);
				memcpy(o->red_d + cnt * m, o->tb_d + prev * m, sizeof(uint32_t) * m);
				++cnt;
				prev_size = o->tb[t].size;
				cur_lb = o->tb[prev].rb;
			}
			prev = t;
		}
		dp_arg e = {cur_lb, o->tb[prev].rb, UINT32_MAX, prev_size};
		o->red[cnt] = e;
		memcpy(o->red_a + cnt * m, o->tb_a + prev * m, sizeof(uint32_t) * m);
		memcpy(o->red_d + cnt * m, o->tb_d + prev * m, sizeof(uint32_t) * m);
		o->red_size = cnt + 1;
	}

	pbwt_free(&p); free(ck_a); free(ck_d);
	return 0;
}

double fsqo_first_pass_seconds(void *hp) { return ((oracle *) hp)->first_pass_seconds; }
uint32_t fsqo_K(void *hp) { return ((oracle *) hp)->K; }
uint64_t fsqo_dp_size(void *hp) { return ((oracle *) hp)->dp_size; }
uint64_t fsqo_traceback_size(void *hp) { return ((oracle *) hp)->tb_size; }
uint64_t fsqo_reduced_size(void *hp) { return ((oracle *) hp)->red_size; }

static void copy_args(dp_arg const *v, uint64_t cnt, uint64_t *lb, uint64_t *rb, uint32_t *mx, uint32_t *sz)
{
	for (uint64_t i = 0; i < cnt; ++i) { lb[i] = v[i].lb; rb[i] = v[i].rb; mx[i] = v[i].max; sz[i] = v[i].size; }
}

void fsqo_get_dp(void *hp, uint64_t *lb, uint64_t *rb, uint32_t *mx, uint32_t *sz)
{ oracle *o = hp; copy_args(o->dp, o->dp_size, lb, rb, mx, sz); }
void fsqo_get_traceback(void *hp, uint64_t *lb, uint64_t *rb, uint32_t *mx, uint32_t *sz)
{ oracle *o = hp; copy_args(o->tb, o->tb_size, lb, rb, mx, sz); }
void fsqo_get_reduced(void *hp, uint64_t *lb, uint64_t *rb, uint32_t *mx, uint32_t *sz)
{ oracle *o = hp; copy_args(o->red, o->red_size, lb, rb, mx, sz); }
void fsqo_get_traceback_states(void *hp, uint32_t *a, uint32_t *d)
{ oracle *o = hp; memcpy(a, o->tb_a, sizeof(uint32_t) * o->m * o->tb_size); memcpy(d, o->tb_d, sizeof(uint32_t) * o->m * o->tb_size); }
void fsqo_get_reduced_states(void *hp, uint32_t *a, uint32_t *d)
{ oracle *o = hp; memcpy(a, o->red_a, sizeof(uint32_t) * o->m * o->red_size); memcpy(d, o->red_d, sizeof(uint32_t) * o->m * o->red_size); }

static int cmp_u32(void const *x, void const *y)
{
	uint32_t const a = *(uint32_t const *) x, b = *(uint32_t const *) y;
	return a < b ? -1 : (a > b);
}

/* Runs of reduced segment s in PBWT order (lib/libbio/include/libbio/pbwt.hh:253-280; join_context.cc:63-126)
 * and the ascending row-id set of every run (create_segment_texts_task.cc:22-35), CSR.
 * run_off has m+1 capacity, ids has m entries, class_of_row has m entries.  Returns the number of runs. */
uint32_t fsqo_segment_sets(void *hp, uint64_t s, uint32_t *run_off, uint32_t *ids, uint32_t *class_of_row)
{
	oracle *o = hp;
	uint32_t const m = o->m;
	uint32_t const *a = o->red_a + s * m, *d = o->red_d + s * m;
	uint64_t const lb = o->red[s].lb;
	uint32_t runs = 0;
	for (uint32_t i = 0; i < m; ++i)
	{
		if (0 == i || lb < d[i]) run_off[runs++] = i;
		ids[i] = a[i];
		class_of_row[a[i]] = runs - 1;
	}
	run_off[runs] = m;
	for (uint32_t r = 0; r < runs; ++r)
		qsort(ids + run_off[r], run_off[r + 1] - run_off[r], sizeof(uint32_t), cmp_u32);
	return runs;
}

/* merge_segments_task.cc:133-176 restricted to the non-copied slots: W[i][j] = |S_i ∩ T_j| for the runs (PBWT order)
 * of reduced segments s and s+1; lib/libbio/include/libbio/algorithm/set_count.hh:61-89.  W is runs_l × runs_r. */
void fsqo_pair_weights(void *hp, uint64_t s, int32_t *W)
{
	oracle *o = hp;
	uint32_t const m = o->m;
	uint32_t *off_l = malloc(sizeof(uint32_t) * (m + 1)), *off_r = malloc(sizeof(uint32_t) * (m + 1));
	uint32_t *ids_l = malloc(sizeof(uint32_t) * m), *ids_r = malloc(sizeof(uint32_t) * m);
	uint32_t *cl = malloc(sizeof(uint32_t) * m);
	uint32_t const rl = fsqo_segment_sets(hp, s, off_l, ids_l, cl);
	uint32_t const rr = fsqo_segment_sets(hp, s + 1, off_r, ids_r, cl);
	for (uint32_t i = 0; i < rl; ++i)
		for (uint32_t j = 0; j < rr; ++j)
		{
			/* sorted-merge intersection count */
			uint32_t x = off_l[i], y = off_r[j], c = 0;
			while (x < off_l[i + 1] && y < off_r[j + 1])
			{
				if (ids_l[x] < ids_r[y]) ++x;
				else if (ids_r[y] < ids_l[x]) ++y;
				else { ++c; ++x; ++y; }
			}
			W[(size_t) i * rr + j] = (int32_t) c;
		}
	free(off_l); free(off_r); free(ids_l); free(ids_r); free(cl);
}

/* Short path: founder-sequences/segmentation_sp_context.cc:21-28; pbwt.hh:253-280 with lb = 0. */
uint64_t fsqo_short_path(void *hp, uint32_t *row_of_run, uint32_t *run_len)
{
	oracle *o = hp;
	pbwt p;
	pbwt_init(&p, o->m, o->n, 0);
	while (p.k < o->n) pbwt_step(&p, o);
	uint64_t runs = 0;
	for (uint32_t i = 0; i < o->m; ++i)
	{
		if (0 == i || 0 < p.d[i]) { row_of_run[runs] = p.a[i]; run_len[runs] = 0; ++runs; }
		++run_len[runs - 1];
	}
	pbwt_free(&p);
	return runs;
}

void fsqo_rmq_queries(uint32_t const *vals, uint64_t count, uint64_t const *beg, uint64_t const *end, uint64_t nq, uint64_t *out)
{
	dp_arg *tb = malloc(sizeof(dp_arg) * count);
	rmq64 r; memset(&r, 0, sizeof r); r.vals = tb;
	for (uint64_t i = 0; i < count; ++i)
	{
		tb[i].lb = tb[i].rb = 0; tb[i].max = tb[i].size = vals[i];
		rmq_update(&r, i);
	}
	for (uint64_t q = 0; q < nq; ++q) out[q] = rmq_query(&r, beg[q], end[q]);
	rmq_free(&r); free(tb);
}


/* ---- match-sequences-to-founders/match_founder_sequences.cc:127-214 (match_sequence_and_report) ----
 * Ranges (sequence, lb, rb, founder set) in the order the reference prints them for one sequence after the other
 * (--single-threaded); a founder set is K bits in `words` = ceil(K / 32) uint32.  Returns the number of ranges, or
 * (uint64_t) -1 when `cap` is too small.  The reference's messages to stderr are not part of the result. */
uint64_t fsqo_match_founders(
	uint8_t const *seqs, uint32_t m, uint8_t const *founders, uint32_t K, uint64_t n, uint64_t min_len,
	uint32_t *out_seq, uint64_t *out_lb, uint64_t *out_rb, uint32_t *out_bits, uint64_t cap
)
{
	uint32_t const words = (K + 31) / 32;
	uint8_t *cand = malloc(K ? K : 1), *dst = malloc(K ? K : 1);
	uint64_t count = 0;
	int overflow = 0;
#define FSQO_EMIT(rb_) do { \
		if (count < cap) { \
			out_seq[count] = r; out_lb[count] = lb; out_rb[count] = (rb_); \
			for (uint32_t w = 0; w < words; ++w) out_bits[count * words + w] = 0; \
			for (uint32_t f = 0; f < K; ++f) if (cand[f]) out_bits[count * words + f / 32] |= 1u << (f % 32); \
		} else overflow = 1; \
		++count; } while (0)
	for (uint32_t r = 0; r < m; ++r)
	{
		uint8_t const *seq = seqs + (size_t) r * n;
		uint64_t lb = 0, n_cand = K, n_dst = 0;
		memset(cand, 1, K);
		for (uint64_t chr = 0; chr < n; ++chr)
		{
			uint8_t const c = seq[chr];
			int recheck = (0 != min_len && min_len <= chr - lb);                 /* :163-165 */
			if (!recheck)
			{
				n_dst = 0;                                                          /* compare_to_sequences, :89-108 */
				for (uint32_t f = 0; f < K; ++f) { dst[f] = cand[f] && founders[(size_t) f * n + chr] == c; n_dst += dst[f]; }
				if (0 == n_dst) recheck = 1;                                        /* :169-175 */
			}
			if (recheck)
			{
				FSQO_EMIT(chr);                                                     /* :181 */
				lb = chr;                                                           /* :184-191 */
				n_dst = 0;
				for (uint32_t f = 0; f < K; ++f) { dst[f] = founders[(size_t) f * n + chr] == c; n_dst += dst[f]; }
			}
			n_cand = n_dst;                                                         /* :197-201 */
			{ uint8_t *t = cand; cand = dst; dst = t; }
		}
		if (0 != n_cand) FSQO_EMIT(n);                                             /* :206-215 */
	}
#undef FSQO_EMIT
	free(cand); free(dst);
	return overflow ? (uint64_t) -1 : count;
}
