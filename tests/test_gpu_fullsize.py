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


def test_config2_full_size(ctx, oracle, synth):
    m, n, L, F, seed = 1000, 1000000, 20, 16, 2
    P = 30000                                   # prefix the oracle runs
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
                oa, od = o.scan_dump(2000, hist=False)
            cols = np.array([1, 2, 451, 452, 999, 2000], np.uint64)
            ga, gd = ctx.pbwt_states(msa, cols)
            for i, c in enumerate(cols):
                assert np.array_equal(ga[i], oa[int(c) - 1]) and np.array_equal(gd[i], od[int(c) - 1])
        finally:
            seg.free()

        # ---- invariance under the decomposition ----
        for opts in ({"block_columns": 1000}, {"dp_chunks": 37}, {"store_states": 0}, {"stitch_sort": 0, "block_columns": 333}):
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
