"""Parity tests proper: the CUDA path, called through the C ABI (ctypes over libfsq_b200.so), against the oracle on
the same seeded inputs.  Everything here is integer work, so the bar is bit-exact equality."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

KAT_ROWS = [b"ACGTACGTAC", b"ACGTACGAAC", b"ACCTACGTAC", b"ACGTACGTAC", b"TCGTACGAAC"]


def as_rows(lst):
    return np.frombuffer(b"".join(lst), np.uint8).reshape(len(lst), -1).copy()


def check_states(ctx, oracle, rows, block_columns=None, warps=None, columns=None, kernel=0, chunk=0):
    m, n = rows.shape
    if block_columns:
        ctx.set_option("block_columns", block_columns)
    ctx.set_option("warps", warps or 0)
    ctx.set_option("kernel", kernel)      # 0 auto, 1 striped step, 2 blocked step
    ctx.set_option("chunk", chunk)        # elements per thread of the blocked step (9, 17, 33; 0 auto)
    if columns is None:
        columns = np.arange(1, n + 1)
    with oracle.open(rows) as o:
        oa, od = o.scan_dump(n, hist=False)
    with ctx.msa_from_rows(rows) as msa:
        a, d = ctx.pbwt_states(msa, columns)
    for i, c in enumerate(columns):
        assert np.array_equal(a[i], oa[c - 1]), "a differs at column %d (m=%d n=%d C=%s W=%s)\n%s\n%s" % (c, m, n, block_columns, warps, a[i], oa[c - 1])
        assert np.array_equal(d[i], od[c - 1]), "d differs at column %d (m=%d n=%d C=%s W=%s)\n%s\n%s" % (c, m, n, block_columns, warps, d[i], od[c - 1])


def test_kat_states(ctx, oracle):
    """The 5 x 10 known-answer example of SURVEY.md Appendix B (reference pbwt_context output)."""
    rows = as_rows(KAT_ROWS)
    with ctx.msa_from_rows(rows) as msa:
        a, d = ctx.pbwt_states(msa, [3, 8, 10])
    assert a[0].tolist() == [2, 0, 1, 3, 4] and d[0].tolist() == [3, 3, 0, 0, 1]
    assert a[1].tolist() == [1, 4, 2, 0, 3] and d[1].tolist() == [8, 1, 8, 3, 0]
    assert a[2].tolist() == [1, 4, 2, 0, 3] and d[2].tolist() == [10, 1, 8, 3, 0]


@pytest.mark.parametrize("m,n,C,W", [
    (5, 40, 7, 0), (33, 100, 16, 0), (64, 300, 64, 0), (64, 300, 13, 2), (100, 257, 50, 4),
    (257, 200, 33, 0), (1000, 150, 40, 1), (1000, 150, 40, 2), (1000, 120, 64, 8), (2500, 90, 32, 0),
])
def test_states_every_column(ctx, oracle, synth, m, n, C, W):
    rows = synth.synth_rows(m, n, seed=m + n, founders=max(2, m // 12), gaps=(m % 2 == 0), recomb_block=37, mu=0.02, gap=0.02)
    check_states(ctx, oracle, rows, C, W, kernel=1)


@pytest.mark.parametrize("m,n,C,E", [
    (5, 40, 7, 0), (33, 100, 16, 9), (64, 300, 64, 0), (64, 300, 13, 17), (100, 257, 50, 33), (288, 90, 30, 9), (289, 90, 30, 0),
    (257, 200, 33, 0), (545, 100, 25, 0), (1000, 150, 40, 0), (1000, 150, 40, 9), (1000, 120, 64, 17), (1056, 80, 20, 33),
    (1057, 80, 20, 33), (2500, 90, 32, 0), (5000, 70, 35, 0),
])
def test_states_every_column_blocked(ctx, oracle, synth, m, n, C, E):
    """The blocked column step (E consecutive positions per thread) in every chunk length and CTA shape."""
    rows = synth.synth_rows(m, n, seed=m + n, founders=max(2, m // 12), gaps=(m % 2 == 0), recomb_block=37, mu=0.02, gap=0.02)
    check_states(ctx, oracle, rows, C, kernel=2, chunk=E)


@pytest.mark.parametrize("m,n,C,E", [
    (5, 40, 7, 0), (33, 100, 16, 9), (64, 300, 64, 0), (64, 300, 13, 17), (100, 257, 50, 33), (288, 90, 30, 9), (289, 90, 30, 0),
    (257, 200, 33, 0), (545, 100, 25, 0), (1000, 150, 40, 0), (1000, 150, 40, 9), (1000, 120, 64, 17), (1056, 80, 20, 33),
    (1057, 80, 20, 33), (2500, 90, 32, 0), (5000, 70, 35, 0),
])
def test_states_every_column_chunked(ctx, oracle, synth, m, n, C, E):
    """The chunked column step (fsq_cstep.cuh: float one-hot arithmetic, padded rows) in every chunk length and CTA shape."""
    rows = synth.synth_rows(m, n, seed=m + n, founders=max(2, m // 12), gaps=(m % 2 == 0), recomb_block=37, mu=0.02, gap=0.02)
    check_states(ctx, oracle, rows, C, kernel=3, chunk=E)


def test_states_blocked_nul_and_wide_alphabets(ctx, oracle):
    """NUL bytes in the rows (comp 0 is then a real symbol), 8 symbols, and more than 8 (falls back to the striped step)."""
    rng = np.random.default_rng(11)
    for alphabet in (b"\0AC", b"ACGTNRYK", b"ACGTNRYKMSW"):
        rows = rng.choice(np.frombuffer(alphabet, np.uint8), size=(150, 60))
        rows[:, 20:40] = rows[0, 20:40]
        check_states(ctx, oracle, rows, 11)


def test_states_identical_rows_and_columns(ctx, oracle):
    """Degenerate classes: all rows equal; then one differing row; identity columns."""
    rows = np.full((40, 90), ord("A"), np.uint8)
    check_states(ctx, oracle, rows, 16)
    rows[7, 50] = ord("C")
    rows[21, 10:14] = ord("T")
    check_states(ctx, oracle, rows, 16)


def test_states_random_binary(ctx, oracle):
    rng = np.random.default_rng(5)
    for it in range(6):
        m = int(rng.integers(2, 200))
        n = int(rng.integers(2, 120))
        rows = rng.choice(np.frombuffer(b"AC", np.uint8), size=(m, n), p=[0.85, 0.15])
        check_states(ctx, oracle, rows, int(rng.integers(1, 40)))


def test_synth_device_matches_host(ctx, fsq, synth, oracle):
    """The device generator (packed) is the same matrix as synth.py: same PBWT states."""
    m, n = 70, 400
    rows = synth.synth_rows(m, n, seed=3, founders=5, gaps=True)
    cols = [1, 7, 199, 400]
    with ctx.msa_synth(m, n, 3, 5, gaps=True) as msa_dev, ctx.msa_from_rows(rows) as msa_host:
        a1, d1 = ctx.pbwt_states(msa_dev, cols)
        a2, d2 = ctx.pbwt_states(msa_host, cols)
    assert np.array_equal(a1, a2) and np.array_equal(d1, d2)


def compare_segmentation(ctx, oracle, rows, L, block_columns=None, cap=None, sets=True, all_sets=False):
    m, n = rows.shape
    if block_columns:
        ctx.set_option("block_columns", block_columns)
    if cap:
        ctx.set_option("cap", cap)
    ctx.set_option("keep_dp", 1)
    with oracle.open(rows) as o, ctx.msa_from_rows(rows) as msa:
        status = o.segment(L)
        assert status in (0, 2)
        try:
            seg = ctx.segment(msa, L)
            assert status == 0
        except Exception as e:      # K == m: the reference exits; we report NOT_REDUCIBLE and keep the traceback
            assert status == 2 and getattr(e, "code", None) == 4, e
            seg = e.segmentation
        with seg:
            olb, orb, omx, osz = o.dp()
            lb, rb, mx, sz = seg.dp_table(len(olb))
            bad = np.nonzero((lb != olb) | (mx != omx) | (sz != osz) | (rb != orb))[0]
            assert len(bad) == 0, "DP table differs first at entry %d: got (%d,%d,%d,%d) want (%d,%d,%d,%d)" % (
                bad[0], lb[bad[0]], rb[bad[0]], mx[bad[0]], sz[bad[0]], olb[bad[0]], orb[bad[0]], omx[bad[0]], osz[bad[0]])
            for x, y in zip(seg.traceback(), o.traceback()):
                assert np.array_equal(x, y)
            assert seg.max_segment_size == o.K()
            if status == 2:
                return None
            rlb, rrb, rsz = seg.reduced_traceback()
            qlb, qrb, _, qsz = o.reduced()
            assert np.array_equal(rlb, qlb) and np.array_equal(rrb, qrb) and np.array_equal(rsz, qsz)
            a, d = seg.reduced_pbwt_samples()
            oa, od = o.reduced_states()
            assert np.array_equal(a, oa) and np.array_equal(d, od)
            if sets:
                for s in range(seg.count if all_sets else min(seg.count, 6)):
                    off, ids, cls = seg.segment_sets(s)
                    ooff, oids, ocls = o.segment_sets(s)
                    assert np.array_equal(off, ooff) and np.array_equal(ids, oids) and np.array_equal(cls, ocls)
                    if s + 1 < seg.count:
                        W = seg.pair_weights(s)
                        noff, _, ncls = o.segment_sets(s + 1)
                        assert np.array_equal(W, o.pair_weights(s, len(off) - 1, len(noff) - 1))
                        # greedy_matcher.cc:293-343: every row's (lhs class, rhs class) encoded, sorted, runs counted
                        code, cnt = np.unique(ocls.astype(np.uint64) * (len(noff) - 1) + ncls, return_counts=True)
                        pl, pr, pc = seg.pair_counts(s)
                        assert np.array_equal(pl, code // (len(noff) - 1)) and np.array_equal(pr, code % (len(noff) - 1)) and np.array_equal(pc, cnt)
            return seg.max_segment_size, len(rlb)


def test_rmq_quirk_decides_the_table(ctx, oracle):
    """The fixture on which the reference's rmq (include/founder_sequences/rmq.hh:61-106: sparse levels over blocks
    {i, i+1, i+3, ...}, block-table candidate ahead of the left partial block on ties) decides the DP table, against the
    values the reference's own code produced (tests/golden/make_golden.py).  tests/test_oracle_golden.py shows that an
    exact leftmost range-min gives a different table, traceback and segmentation on it."""
    import os
    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "rmq_dp_quirk_4x1599_L4.npz"))
    rows, L = z["rows"], int(z["L"])
    for opts in ({}, {"block_columns": 100}, {"dp_chunks": 1}, {"dp_chunks": 5}):
        for k, v in opts.items():
            ctx.set_option(k, v)
        ctx.set_option("keep_dp", 1)
        with ctx.msa_from_rows(rows) as msa, ctx.segment(msa, L) as seg:
            lb, rb, mx, sz = seg.dp_table(len(z["dp_lb"]))
            assert np.array_equal(lb, z["dp_lb"]) and np.array_equal(rb, z["dp_rb"]) and np.array_equal(mx, z["dp_max"]) and np.array_equal(sz, z["dp_size"])
            for got, key in zip(seg.traceback(), ("tb_lb", "tb_rb", "tb_max", "tb_size")):
                assert np.array_equal(got, z[key]), key
            rlb, rrb, rsz = seg.reduced_traceback()
            assert np.array_equal(rlb, z["red_lb"]) and np.array_equal(rrb, z["red_rb"]) and np.array_equal(rsz, z["red_size"])
            a, d = seg.reduced_pbwt_samples()
            assert np.array_equal(a, z["red_a"]) and np.array_equal(d, z["red_d"])
    compare_segmentation(ctx, oracle, rows, L)


def test_kat_segmentation(ctx, oracle):
    rows = as_rows(KAT_ROWS)
    compare_segmentation(ctx, oracle, rows, 2, block_columns=4)


@pytest.mark.parametrize("m,n,L,C,founders", [
    (16, 400, 5, 32, 3), (48, 6000, 10, 256, 6), (64, 20000, 10, 0, 8), (200, 3000, 20, 100, 8),
    (1000, 2000, 20, 128, 16), (96, 2000, 1, 64, 6), (96, 2100, 70, 300, 6), (2500, 600, 5, 64, 24),
])
def test_segmentation_matches_oracle(ctx, oracle, synth, m, n, L, C, founders):
    rows = synth.synth_rows(m, n, seed=m + L, founders=founders, gaps=(founders % 2 == 0))
    res = compare_segmentation(ctx, oracle, rows, L, C or None)
    assert res is not None


@pytest.mark.parametrize("m,n,L,founders,seed", [(1500, 8000, 20, 16, 7), (2500, 6000, 50, 24, 4), (5000, 4000, 50, 32, 3), (8000, 1500, 30, 40, 9),
                                                    (20000, 1200, 100, 64, 11)])
def test_dp_table_with_several_warps_per_block(ctx, oracle, synth, m, n, L, founders, seed):
    """More rows than one warp's chunks hold: the scan CTA has several warps, and only the last one emits the top
    lists.  (A missing barrier between that emission and the next step's histogram moves once put single DP entries off
    by one here while every smaller case passed.)  The whole DP table, K > cap included, against the oracle, twice."""
    rows = synth.synth_rows(m, n, seed=seed, founders=founders, gaps=True)
    with oracle.open(rows) as o:
        assert o.segment(L) == 0
        odp = o.dp()
        olb, orb, _, osz = o.reduced()
        oa, od = o.reduced_states()
        K = o.K()
    ctx.set_option("keep_dp", 1)
    try:
        for _ in range(2):
            with ctx.msa_from_rows(rows) as msa, ctx.segment(msa, L) as seg:
                assert seg.max_segment_size == K
                for name, x, y in zip(("lb", "rb", "max", "size"), seg.dp_table(n - L + 1), odp):
                    assert np.array_equal(x, y), "dp %s differs first at entry %d" % (name, int(np.nonzero(x != y)[0][0]))
                lb, rb, sz = seg.reduced_traceback()
                assert np.array_equal(lb, olb) and np.array_equal(rb, orb) and np.array_equal(sz, osz)
                a, d = seg.reduced_pbwt_samples()
                assert np.array_equal(a, oa) and np.array_equal(d, od)
        assert ctx.stat("warps") > 1
    finally:
        ctx.set_option("keep_dp", 0)


@pytest.mark.parametrize("store,kernel", [(0, 0), (1, 0), (-1, 2), (-1, 1)])
@pytest.mark.parametrize("m,n,L,C", [(200, 3000, 20, 100), (1000, 1500, 20, 128), (1101, 900, 8, 77)])
def test_segmentation_stored_states_and_rescan(ctx, oracle, synth, store, kernel, m, n, L, C):
    """States at the traceback positions: out of the states pass B keeps (TMA bulk store) or by the re-scan (pass C); and
    the older step kernels (striped, blocked) through the same pipeline."""
    rows = synth.synth_rows(m, n, seed=m + L + 1, founders=9, gaps=True)
    ctx.set_option("store_states", store)
    ctx.set_option("kernel", kernel)
    ctx.reset_stats()
    res = compare_segmentation(ctx, oracle, rows, L, C, sets=False)
    assert res is not None
    if kernel == 0:
        assert ctx.stat("store_states") == (1 if store else 0)


def test_segmentation_edge_lengths(ctx, oracle, synth):
    """n == 2L, n == 2L + 1, n == 3L - 1: the staging boundaries of the first pass."""
    for n, L in [(20, 10), (21, 10), (29, 10), (30, 10), (31, 10), (130, 64), (200, 65)]:
        rows = synth.synth_rows(24, n, seed=n, founders=3, recomb_block=11, mu=0.01)
        compare_segmentation(ctx, oracle, rows, L, block_columns=9, sets=False)


def test_segmentation_small_cap_retries(ctx, oracle, synth):
    """A top-list bound far below K must be detected by the DP and retried, not silently accepted."""
    rows = synth.synth_rows(128, 1500, seed=9, founders=40, recomb_block=50, mu=0.01)
    ctx.reset_stats()
    res = compare_segmentation(ctx, oracle, rows, 6, block_columns=100, cap=2, sets=False)
    assert res is None or res[0] > 2
    assert ctx.stat("cap_retries") >= 1


def test_segmentation_invariant_under_block_size(ctx, oracle, synth):
    rows = synth.synth_rows(80, 1800, seed=21, founders=7, gaps=True)
    results = []
    for C in (1, 17, 64, 500, 4000):
        ctx.set_option("block_columns", C)
        with ctx.msa_from_rows(rows) as msa, ctx.segment(msa, 12) as seg:
            results.append((seg.max_segment_size,) + tuple(x.tobytes() for x in seg.reduced_traceback()) + tuple(x.tobytes() for x in seg.reduced_pbwt_samples()))
    assert all(r == results[0] for r in results)


@pytest.mark.parametrize("B", [1, 2, 3, 4, 5, 7, 8, 9, 16, 17, 31, 32, 33, 63])
def test_states_for_every_shape_of_the_sweep(ctx, oracle, synth, B):
    """The block summaries are composed by an up-sweep and a down-sweep over a binary tree of B blocks: powers of two, their
    neighbours, and numbers with many one bits (partial nodes at every level)."""
    n = 504
    C = (n + B - 1) // B
    rows = synth.synth_rows(130, n, seed=77 + B, founders=9, gaps=True, recomb_block=41, mu=0.01)
    check_states(ctx, oracle, rows, C, columns=np.arange(1, n + 1, 7))


@pytest.mark.parametrize("sort,m", [(0, 130), (1, 130), (1, 33), (1, 1000), (0, 1000), (1, 1024), (1, 1025), (2, 130), (2, 33), (2, 1000), (2, 1024), (2, 1025), (2, 31)])
def test_states_invariant_under_stitch_kernel(ctx, oracle, synth, sort, m):
    """The boundary stitch has three formulations (for m <= 1024 a sorting network or a count per (class, warp), both with a
    sparse table for the divergences; radix passes otherwise)."""
    rows = synth.synth_rows(m, 420, seed=31 + m, founders=11, gaps=True, recomb_block=37, mu=0.01)
    ctx.set_option("stitch_sort", sort)
    ctx.set_option("block_columns", 30)
    try:
        check_states(ctx, oracle, rows, 12, columns=np.arange(1, 421, 11))
    finally:
        ctx.set_option("stitch_sort", 2)


@pytest.mark.parametrize("chunks,max_rounds", [(1, 12), (3, 12), (0, 12), (1000, 12), (1000, 2)])
def test_segmentation_invariant_under_dp_chaining(ctx, oracle, synth, chunks, max_rounds):
    """The DP table is cut into chains that run at once and repeat until a round changes nothing; one chain, many chains
    and the give-up-and-run-one-chain path must all produce the reference's table."""
    rows = synth.synth_rows(120, 9000, seed=5, founders=10, gaps=True, recomb_block=300, mu=0.004)
    ctx.set_option("dp_chunks", chunks)
    ctx.set_option("dp_max_rounds", max_rounds)
    ctx.reset_stats()
    compare_segmentation(ctx, oracle, rows, 6, block_columns=700, sets=False)
    if chunks == 1:
        assert ctx.stat("dp_rounds") == 1
    if chunks == 1000:
        assert ctx.stat("dp_chunks") > 8


def test_not_reducible_and_short(ctx, fsq, oracle):
    rng = np.random.default_rng(1)
    rows = rng.choice(np.frombuffer(b"ACGT", np.uint8), size=(6, 40))
    with ctx.msa_from_rows(rows) as msa:
        with pytest.raises(fsq.FsqError) as ei:
            ctx.segment(msa, 30)
        assert ei.value.code == fsq.ERR_SHORT
        row, ln = ctx.distinct_rows(msa)
        with oracle.open(rows) as o:
            orow, oln = o.short_path()
        assert np.array_equal(row, orow) and np.array_equal(ln, oln)
    compare_segmentation(ctx, oracle, rows, 3, block_columns=8, sets=False)


def test_errors_are_loud(ctx, fsq):
    with pytest.raises(fsq.FsqError):
        ctx.set_option("no_such_option", 1)
    rows = np.arange(20 * 64, dtype=np.uint8).reshape(64, 20) % 23 + 65     # 23 distinct characters
    with pytest.raises(fsq.FsqError) as ei:
        ctx.msa_from_rows(rows)
    assert ei.value.code == fsq.ERR_ARG


@pytest.mark.parametrize("m,n,L,C,wb", [(96, 3000, 8, 64, 3), (300, 2500, 12, 100, 1), (1500, 1200, 10, 0, 2)])
def test_segmentation_with_windowed_top_lists(ctx, oracle, synth, m, n, L, C, wb):
    """The exact scan and the DP alternate over windows of `wb` blocks that share one top-list buffer (what a shape whose
    lists do not fit at once runs as); the reduction then works from the divergence arrays of a re-scan."""
    rows = synth.synth_rows(m, n, seed=m + n, founders=7, gaps=True)
    ctx.set_option("window_blocks", wb)
    ctx.reset_stats()
    compare_segmentation(ctx, oracle, rows, L, block_columns=C)
    assert ctx.stat("windows") > 1


def test_windowed_top_lists_grow_the_bound_where_it_fails(ctx, oracle, synth):
    """A bound that is too small is found out in some window; the scan resumes there with the bound the DP proved."""
    rows = synth.synth_rows(128, 2000, seed=9, founders=40, recomb_block=50, mu=0.01)
    ctx.set_option("window_blocks", 2)
    ctx.reset_stats()
    compare_segmentation(ctx, oracle, rows, 6, block_columns=100, cap=2, sets=False)
    assert ctx.stat("cap_retries") >= 1 and ctx.stat("windows") > 1


def test_memory_limit_turns_into_windows(ctx, oracle, synth):
    """mem_limit stands in for the device's answer: with room for the lists of a few blocks only, the plan is windows."""
    rows = synth.synth_rows(200, 4000, seed=3, founders=9, gaps=True)
    ctx.set_option("mem_limit", (1 << 30) + 4000 * 64 + 4100 * 32 + 402 * 200 * 16 + 300 * 128 * 8)       # what run_plan sets aside + the lists of 300 columns
    ctx.reset_stats()
    compare_segmentation(ctx, oracle, rows, 10, block_columns=100, sets=False)
    assert ctx.stat("windows") > 1


@pytest.mark.parametrize("m,n", [(5, 37), (64, 5000), (1000, 4100), (2500, 1025)])
def test_identity_columns(ctx, oracle, synth, m, n):
    """remove-identity-columns/main.cc:107-160: a column is dropped when the set of its characters has one element; the
    alignment that is left must scan exactly like the rows with those columns cut out on the host."""
    rows = synth.synth_rows(m, n, seed=m * 7 + n, founders=3, gaps=True, mu=2e-4)
    rows[:, n // 3: n // 3 + 40] = rows[0, n // 3: n // 3 + 40]          # a run of identity columns for sure
    want = (rows == rows[0]).all(axis=0).astype(np.uint8)
    comp = np.zeros(256, np.uint8)
    for i, ch in enumerate(sorted(set(rows.tobytes()) | {0})):
        comp[ch] = i
    with ctx.msa_from_rows(rows, comp) as msa:
        mask, reduced = ctx.identity_columns(msa, want_msa=True)
        assert np.array_equal(mask, want) and 0 < want.sum() < n
        kept = np.ascontiguousarray(rows[:, want == 0])
        assert reduced.n == kept.shape[1]
        cols = np.unique(np.linspace(1, kept.shape[1], 7).astype(np.uint64))
        a, d = ctx.pbwt_states(reduced, cols)
        with ctx.msa_from_rows(kept, comp) as direct:
            a2, d2 = ctx.pbwt_states(direct, cols)
        reduced.free()
    assert np.array_equal(a, a2) and np.array_equal(d, d2)
    with oracle.open(kept) as o:
        oa, od = o.scan_dump(int(cols[-1]), hist=False)[:2]
    for i, c in enumerate(cols):
        assert np.array_equal(a[i], oa[int(c) - 1]) and np.array_equal(d[i], od[int(c) - 1])


def mosaic_rows(m, n, letters, seed, founders=8, block=150, mu=2e-3):
    """A founder mosaic over an arbitrary alphabet (the stock generator knows ACGT and '-')."""
    rng = np.random.default_rng(seed)
    al = np.frombuffer(letters, np.uint8)
    f = al[rng.integers(0, len(al), size=(founders, n))]
    rows = np.empty((m, n), np.uint8)
    for r in range(m):
        starts = np.arange(0, n, block)
        pick = rng.integers(0, founders, size=len(starts))
        for s, p in zip(starts, pick):
            rows[r, s:s + block] = f[p, s:s + block]
    mut = rng.random((m, n)) < mu
    rows[mut] = al[rng.integers(0, len(al), size=int(mut.sum()))]
    return rows


@pytest.mark.parametrize("m,n,L,letters", [
    (40, 1500, 6, b"ACGTNRYKM"),                       # 9 comps + NUL
    (1000, 3000, 15, b"ACGTNRYKMSWB"),                # 12
    (1000, 2500, 12, b"-ACGTNRYKMSWBDH"),             # 15: the most the 4-bit packing holds
    (2500, 1200, 20, b"ACGTacgtNn-"),                 # lower case and N: FASTA as it comes
])
def test_wide_alphabets(ctx, oracle, m, n, L, letters):
    """Alphabets of 9 .. 15 characters run on the chunked scan as two steps per column over the base-4 digits of the
    comps (consecutive_alphabet.hh:259-264 gives any byte a comp; DNA with IUPAC codes or lower case needs more than
    eight)."""
    rows = mosaic_rows(m, n, letters, seed=m + len(letters))
    ctx.reset_stats()
    with ctx.msa_from_rows(rows) as msa:
        cols = np.unique(np.linspace(1, n, 9).astype(np.uint64))
        a, d = ctx.pbwt_states(msa, cols)
    with oracle.open(rows) as o:
        oa, od = o.scan_dump(n, hist=False)[:2]
    for i, c in enumerate(cols):
        assert np.array_equal(a[i], oa[int(c) - 1]) and np.array_equal(d[i], od[int(c) - 1]), int(c)
    compare_segmentation(ctx, oracle, rows, L)
    assert ctx.stat("kernel") == 3        # the chunked step, not the striped fallback


def test_alphabet_of_more_than_15_characters_is_refused(ctx):
    rows = np.frombuffer(bytes(range(65, 65 + 16)) * 4, np.uint8).reshape(4, 16).copy()
    with pytest.raises(Exception) as e:
        ctx.msa_from_rows(rows)
    assert e.value.code == 1 and "15 distinct characters" in str(e.value)


@pytest.mark.parametrize("m,n,K,min_len", [(7, 300, 3, 0), (40, 2500, 9, 0), (64, 4000, 70, 0), (25, 1500, 33, 40), (12, 900, 5, 7)])
def test_match_founders(ctx, oracle, synth, m, n, K, min_len):
    """match_founder_sequences (match-sequences-to-founders/match_founder_sequences.cc:127-214): greedy longest match of
    every input row against the founders; rows that carry characters no founder has produce ranges with empty sets."""
    rng = np.random.default_rng(m * 31 + K)
    founders = synth.synth_rows(K, n, seed=K + 3, founders=max(2, K // 2), gaps=True, recomb_block=120)
    pick = rng.integers(0, K, size=(m, (n + 79) // 80))
    seqs = np.empty((m, n), np.uint8)
    for r in range(m):
        for b in range(pick.shape[1]):
            seqs[r, 80 * b: 80 * b + 80] = founders[pick[r, b], 80 * b: 80 * b + 80]
    noise = rng.random((m, n)) < 0.004
    seqs[noise] = np.frombuffer(b"ACGTN", np.uint8)[rng.integers(0, 5, size=int(noise.sum()))]      # 'N': in no founder
    comp = np.zeros(256, np.uint8)
    for i, ch in enumerate(sorted(set(seqs.tobytes()) | set(founders.tobytes()) | {0})):
        comp[ch] = i
    with ctx.msa_from_rows(seqs, comp) as ms, ctx.msa_from_rows(founders, comp) as mf:
        got = ctx.match_founders(ms, mf, min_len)
    want = oracle.match_founders(seqs, founders, min_len)
    assert len(got[0]) == len(want[0]) and len(want[0]) >= m
    for g, w in zip(got, want):
        assert np.array_equal(g, w)
    assert (want[3].sum(axis=1) == 0).any() or min_len        # some 'N' produced an empty set



@pytest.mark.parametrize("m,n,L,founders,cap", [(1500, 4000, 20, 16, None), (128, 1500, 6, 40, 2), (2500, 1500, 30, 24, 8)])
def test_sets_weights_and_pair_counts_over_all_segments(ctx, oracle, synth, m, n, L, founders, cap):
    """Segment sets (create_segment_texts_task.cc:15-35), edge weights (merge_segments_task.cc:133-176) and the greedy
    matcher's pair counts (greedy_matcher.cc:293-343) for EVERY reduced segment and adjacent pair: a scan CTA of several
    warps (1,500 and 2,500 rows) and bounds far below K that the DP has to grow."""
    kw = dict(recomb_block=50, mu=0.01) if founders == 40 else {}
    rows = synth.synth_rows(m, n, seed=m + L, founders=founders, gaps=True, **kw)
    ctx.reset_stats()
    res = compare_segmentation(ctx, oracle, rows, L, cap=cap, all_sets=True)
    assert res is not None and res[1] > 6
    if cap:
        assert ctx.stat("cap_retries") >= 1


@pytest.mark.parametrize("m,n,gaps", [(1, 9, False), (33, 64, True), (300, 199, True), (1057, 120, False), (2500, 398, True), (6000, 60, True)])
def test_short_path_shapes(ctx, fsq, oracle, synth, m, n, gaps):
    """n < 2L (generate_context.cc:379-390 -> segmentation_sp_context.cc:21-28): the distinct rows in PBWT order with their
    multiplicities; one row, more rows than one warp's chunks hold, gaps, duplicates."""
    rows = synth.synth_rows(m, n, seed=m + n, founders=max(1, min(5, m)), gaps=gaps, recomb_block=40, mu=0.002)
    with ctx.msa_from_rows(rows) as msa:
        with pytest.raises(fsq.FsqError) as ei:
            ctx.segment(msa, n // 2 + 1)
        assert ei.value.code == fsq.ERR_SHORT
        row, ln = ctx.distinct_rows(msa)
    with oracle.open(rows) as o:
        orow, oln = o.short_path()
    assert np.array_equal(row, orow) and np.array_equal(ln, oln)
    assert ln.sum() == m and len(np.unique(rows, axis=0)) == len(row)
