"""The C restatement (oracle/libfsq_oracle.so) against the reference's own code compiled in place
(oracle/_ref/libfsq_ref.so, `make -C oracle ref`).  Only runs where /root/reference was available at build time;
elsewhere the committed fixtures (test_oracle_golden.py) carry the pin."""
import numpy as np
import pytest

import oraclelib

pytestmark = pytest.mark.skipif(not oraclelib.have_ref(), reason="oracle/_ref/libfsq_ref.so not built (needs /root/reference)")


@pytest.fixture(scope="module")
def ref():
    return oraclelib.ref()


def compare(oracle, ref, rows, L, ncols_dump=60):
    with oracle.open(rows) as o, ref.open(rows) as r:
        present = np.union1d(np.unique(rows), [0])
        assert np.array_equal(o.alphabet()[present], r.alphabet()[present])
        nc = min(ncols_dump, rows.shape[1])
        for x, y in zip(o.scan_dump(nc), r.scan_dump(nc)):
            assert np.array_equal(x, y)
        so, sr = o.segment(L), r.segment(L)
        assert so == sr
        if so == 1:
            for x, y in zip(o.short_path(), r.short_path()):
                assert np.array_equal(x, y)
            return so
        for x, y in zip(o.dp(), r.dp()):
            assert np.array_equal(x, y)
        for x, y in zip(o.traceback(), r.traceback()):
            assert np.array_equal(x, y)
        assert o.K() == r.K()
        if so == 0:
            for x, y in zip(o.reduced(), r.reduced()):
                assert np.array_equal(x, y)
            for x, y in zip(o.reduced_states(), r.reduced_states()):
                assert np.array_equal(x, y)
        return so


@pytest.mark.parametrize("m,n,L,founders,gaps", [
    (16, 300, 4, 3, False), (48, 3000, 10, 6, True), (64, 8000, 10, 8, False), (200, 1200, 20, 8, True),
    (96, 900, 1, 6, False), (96, 1500, 70, 6, True), (500, 800, 16, 12, True),
])
def test_founder_mosaic(oracle, ref, synth, m, n, L, founders, gaps):
    rows = synth.synth_rows(m, n, seed=m + L, founders=founders, gaps=gaps)
    assert compare(oracle, ref, rows, L) == 0


def test_staging_boundaries(oracle, ref, synth):
    """n == 2L, 2L + 1, 3L - 1, 3L, 3L + 1 and a short-path case."""
    for n, L in [(20, 10), (21, 10), (29, 10), (30, 10), (31, 10), (130, 64), (19, 10)]:
        rows = synth.synth_rows(24, n, seed=n, founders=3, recomb_block=11, mu=0.01)
        compare(oracle, ref, rows, L)


def test_random_inputs(oracle, ref):
    rng = np.random.default_rng(3)
    for it in range(25):
        m = int(rng.integers(2, 40))
        n = int(rng.integers(4, 150))
        L = int(rng.integers(1, 12))
        sigma = int(rng.integers(2, 5))
        rows = rng.choice(np.frombuffer(b"ACGT", np.uint8)[:sigma], size=(m, n), p=None)
        compare(oracle, ref, rows, L)


def test_rmq_against_compiled_rmq_hh(oracle, ref):
    rng = np.random.default_rng(11)
    vals = rng.integers(0, 50, size=5000).astype(np.uint32)
    beg = rng.integers(0, 4999, size=4000).astype(np.uint64)
    end = np.minimum(5000, beg + rng.integers(1, 4000, size=4000).astype(np.uint64)).astype(np.uint64)
    assert np.array_equal(oracle.rmq_queries(vals, beg, end), ref.rmq_queries(vals, beg, end))
