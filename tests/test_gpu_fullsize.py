"""Parity at BASELINE.json's full sizes (through the C ABI, on the GPU).

Config 1 (64 x 100k, L = 10) is small enough for the oracle to run whole: everything is compared bit for bit.
Config 2 (1000 x 1M, sigma = 5, L = 20) is too long for the CPU oracle, so it is pinned three ways:
  * prefix parity -- PBWT states and DP entries depend only on the columns before them, so the states at columns of a
    prefix and the DP entries t < P - 2L + 1 must equal what the oracle computes on the first P columns alone;
  * size-independent properties of the result -- the reduced segments tile [0, n) with lengths >= L, every segment's
    size is the number of distinct rows over it, recomputed from the returned (a, d) state (pbwt.hh:206-216), no size
    exceeds K, K is the largest size on the traceback, every a is a permutation;
  * invariance -- another block length, another DP chaining and the kept-states / re-scan paths give identical bytes.
"""
import zlib

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_config1_full_size_against_oracle(ctx, oracle, synth):
    m, n, L = 64, 100000, 10
    rows = synth.synth_rows(m, n, seed=1, founders=8, gaps=False)
    with ctx.msa_from_rows(rows) as msa, ctx.segment(msa, L) as seg:
        lb, rb, sz = seg.reduced_traceback()
        a, d = seg.reduced_pbwt_samples()
        tlb, trb, tmx, tsz = seg.traceback()
        K = seg.max_segment_size
    with oracle.open(rows) as o:
        assert o.segment(L) == 0
        assert o.K() == K
        olb, orb, _, osz = o.reduced()
        assert np.array_equal(lb, olb) and np.array_equal(rb, orb) and np.array_equal(sz, osz)
        oa, od = o.reduced_states()
        assert np.array_equal(a, oa) and np.array_equal(d, od)
        xlb, xrb, xmx, xsz = o.traceback()
        assert np.array_equal(tlb, xlb) and np.array_equal(trb, xrb) and np.array_equal(tmx, xmx) and np.array_equal(tsz, xsz)


def _digest(seg):
    lb, rb, sz = seg.reduced_traceback()
    a, d = seg.reduced_pbwt_samples_view()
    return (seg.max_segment_size, lb.tobytes(), rb.tobytes(), sz.tobytes(), zlib.crc32(a.tobytes()), zlib.crc32(d.tobytes()))


def _full_size_case(ctx, oracle, synth, m, n, L, F, seed, P, variants):
    """P = length of the prefix the oracle runs; variants = option sets whose result must have the same bytes."""
    msa = ctx.msa_synth(m, n, seed, F, True)
    try:
        ctx.set_option("keep_dp", 1)
        seg = ctx.segment(msa, L)
        ctx.set_option("keep_dp", 0)
        try:
            K = seg.max_segment_size
            lb, rb, sz = seg.reduced_traceback()
            a, d = seg.reduced_pbwt_samples_view()
            tlb, trb, tmx, tsz = seg.traceback()
            dlb, drb, dmx, dsz = seg.dp_table(n - L + 1)
            base = _digest(seg)

            # ---- properties ----
            assert K < m and len(lb) > 0
            assert lb[0] == 0 and rb[-1] == n and np.array_equal(lb[1:], rb[:-1])
            assert np.all(rb - lb >= L)
            assert sz.max() <= K and tmx[-1] == K and tmx.max() == K
            assert tlb[0] == 0 and trb[-1] == n and np.array_equal(tlb[1:], trb[:-1]) and np.all(trb - tlb >= L)
            distinct = (d > lb[:, None].astype(np.uint32)).sum(axis=1)      # pbwt.hh:206-216
            assert np.array_equal(distinct.astype(np.uint32), sz)
            assert np.all(d[:, 0] == rb.astype(np.uint32))                  # d[0] = k + 1 after column k
            ident = np.arange(m, dtype=np.uint32)
            for i in range(0, len(lb), 97):
                assert np.array_equal(np.sort(a[i]), ident)

            # ---- prefix parity with the oracle ----
            rows = synth.synth_rows(m, P, seed=seed, founders=F, gaps=True)
            with oracle.open(rows) as o:
                assert o.segment(L) == 0
                olb, orb, omx, osz = o.dp()
                cnt = P - 2 * L + 1                                      # entries whose column lies inside the prefix
                assert np.array_equal(dlb[:cnt], olb[:cnt]) and np.array_equal(drb[:cnt], orb[:cnt])
                assert np.array_equal(dmx[:cnt], omx[:cnt]) and np.array_equal(dsz[:cnt], osz[:cnt])
                n_dump = min(2000, P)
                oa, od = o.scan_dump(n_dump, hist=False)
            cols = np.array([c for c in (1, 2, 451, 452, 999, 2000) if c <= n_dump], np.uint64)
            ga, gd = ctx.pbwt_states(msa, cols)
            for i, c in enumerate(cols):
                assert np.array_equal(ga[i], oa[int(c) - 1]) and np.array_equal(gd[i], od[int(c) - 1])
        finally:
            seg.free()

        # ---- invariance under the decomposition ----
        for opts in variants:
            for k, v in opts.items():
                ctx.set_option(k, v)
            try:
                with ctx.segment(msa, L) as other:
                    assert _digest(other) == base, opts
            finally:
                for k in opts:
                    ctx.set_option(k, {"block_columns": 0, "dp_chunks": 0, "store_states": -1, "stitch_sort": 1}[k])
    finally:
        msa.free()


def test_config2_full_size(ctx, oracle, synth):
    _full_size_case(ctx, oracle, synth, 1000, 1000000, 20, 16, 2, 30000,
                    ({"block_columns": 1000}, {"dp_chunks": 37}, {"store_states": 0}, {"stitch_sort": 0, "block_columns": 333}))


# Config 4 (BASELINE.json: segment-length-bound sweep on 2,500 x 5M; SURVEY.md section 8d: F = 24, seed 4): L = 50 and
# 200 at the full 5M columns; L = 5 and 10 at 1M columns, where the reduced states the test reads back (segments x m x
# 8 B on the host) stay below 4 GB.  Three warps per scan CTA (chunk length 27), DP wavefronts of 5 .. 16 columns.
@pytest.mark.parametrize("n,L", [(5_000_000, 50), (5_000_000, 200), (1_000_000, 5), (1_000_000, 10)])
def test_config4_length_bound_sweep(ctx, oracle, synth, n, L):
    _full_size_case(ctx, oracle, synth, 2500, n, L, 24, 4, 12000,
                    ({"block_columns": 700}, {"dp_chunks": 37}, {"store_states": 0}))


# Config 3's share of one GPU at 8 GPUs (5,000 x 1.25M, L = 50; the column-sharded whole is tests/sharded_torchrun.py)
def test_config3_one_gpu_share(ctx, oracle, synth):
    _full_size_case(ctx, oracle, synth, 5000, 1_250_000, 50, 32, 3, 6000,
                    ({"block_columns": 900}, {"dp_chunks": 37}, {"store_states": 0}))


# Config 5's rows (20,000, L = 100, F = 64, seed 5) on a 250k-column slice: chunk length 53, one CTA per SM, K in the
# thousands (cap 8,128) and about 15 DP rounds.  The full 50M columns are open.
def test_config5_rows_slice(ctx, oracle, synth):
    _full_size_case(ctx, oracle, synth, 20000, 250_000, 100, 64, 5, 3000,
                    ({"dp_chunks": 37}, {"store_states": 0}))


# ---- shapes whose reduced states do not fit the host at once: read back in slices --------------------------------------
def _sliced_digest_and_checks(seg, m, n, L, check=True, step=40000):
    """K, the bounds' bytes and a running digest of the states, slice by slice; with `check`, the size-independent
    properties of _full_size_case on every slice."""
    lb, rb, sz = seg.reduced_traceback()
    K = seg.max_segment_size
    if check:
        assert K < m and len(lb) > 0
        assert lb[0] == 0 and rb[-1] == n and np.array_equal(lb[1:], rb[:-1])
        assert np.all(rb - lb >= L) and sz.max() <= K
    acc = np.uint64(0)
    ident = np.arange(m, dtype=np.uint32)
    for f in range(0, len(lb), step):
        c = min(step, len(lb) - f)
        a, d = seg.reduced_pbwt_samples_range(f, c)
        if check:
            distinct = (d > lb[f:f + c, None].astype(np.uint32)).sum(axis=1)      # pbwt.hh:206-216
            assert np.array_equal(distinct.astype(np.uint32), sz[f:f + c])
            assert np.all(d[:, 0] == rb[f:f + c].astype(np.uint32))
            assert np.array_equal(np.sort(a[0]), ident) and np.array_equal(np.sort(a[c - 1]), ident)
        # (a cheap running digest: wrapping sums and xors of the states read as 64-bit words, chained over the slices)
        va, vd = a.reshape(-1)[: c * m // 2 * 2].view(np.uint64), d.reshape(-1)[: c * m // 2 * 2].view(np.uint64)
        with np.errstate(over="ignore"):
            acc = acc * np.uint64(1000003) + va.sum() + np.bitwise_xor.reduce(va) * np.uint64(3) + vd.sum() * np.uint64(5) + np.bitwise_xor.reduce(vd) * np.uint64(7)
    return K, lb.tobytes(), rb.tobytes(), sz.tobytes(), int(acc)


def _large_case(ctx, oracle, synth, m, n, L, F, seed, P, variants):
    msa = ctx.msa_synth(m, n, seed, F, True)
    try:
        ctx.set_option("keep_dp", 1)
        with ctx.segment(msa, L) as seg:
            ctx.set_option("keep_dp", 0)
            base = _sliced_digest_and_checks(seg, m, n, L)
            dlb, drb, dmx, dsz = seg.dp_table(n - L + 1)
        rows = synth.synth_rows(m, P, seed=seed, founders=F, gaps=True)
        with oracle.open(rows) as o:
            assert o.segment(L) == 0
            olb, orb, omx, osz = o.dp()
            cnt = P - 2 * L + 1
            assert np.array_equal(dlb[:cnt], olb[:cnt]) and np.array_equal(drb[:cnt], orb[:cnt])
            assert np.array_equal(dmx[:cnt], omx[:cnt]) and np.array_equal(dsz[:cnt], osz[:cnt])
        del dlb, drb, dmx, dsz
        for opts in variants:
            for k, v in opts.items():
                ctx.set_option(k, v)
            try:
                with ctx.segment(msa, L) as other:
                    assert _sliced_digest_and_checks(other, m, n, L, check=False) == base, opts
            finally:
                for k in opts:
                    ctx.set_option(k, {"block_columns": 0, "dp_chunks": 0, "store_states": -1, "window_blocks": 0}[k])
    finally:
        ctx.set_option("keep_dp", 0)
        msa.free()


# Config 4 at L = 5 and 10 over the full 5M columns: 660,505 and 364,395 reduced segments (13 and 7 GB of states).
@pytest.mark.parametrize("L", [5, 10])
def test_config4_short_bounds_full_length(ctx, oracle, synth, L):
    _large_case(ctx, oracle, synth, 2500, 5_000_000, L, 24, 4, 8000, ({"dp_chunks": 37},))


# Config 3 whole on one GPU with default options (25 GB packed, 82 GB of top lists at the bound the DP settles on, states
# by re-scan); the same bytes with the lists in windows over one buffer.
def test_config3_whole_on_one_gpu(ctx, oracle, synth):
    _large_case(ctx, oracle, synth, 5000, 10_000_000, 50, 32, 3, 5000, ({"window_blocks": 700},))
