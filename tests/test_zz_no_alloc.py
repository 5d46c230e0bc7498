"""fsq_pair_counts_noalloc (include/fsq.h): the counting step of fsq_pair_counts on the caller's device memory, its
contingency table in a static device array instead of cudaMalloc'd scratch.  Checked through the C ABI against
fsq_pair_counts itself, bit for bit (integers), on calls in a row that share the static table.

This file is collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_CELLS = 1 << 20          # FSQ_PAIR_COUNTS_NOALLOC_MAX_CELLS


def _call(lib, torch, cls_l, cls_r, runs_l, runs_r, out=None):
    """One call on device copies of the class maps; returns (rc, message, lhs, rhs, count, n_pairs)."""
    m = len(cls_l)
    dev = "cuda:0"
    d_l = torch.from_numpy(cls_l.astype(np.int32)).to(dev)
    d_r = torch.from_numpy(cls_r.astype(np.int32)).to(dev)
    if out is None:
        # outputs start from garbage: whatever the call leaves untouched must not be mistaken for a result
        out = [torch.full((m,), -7, dtype=torch.int32, device=dev) for _ in range(3)] + [torch.full((1,), -7, dtype=torch.int32, device=dev)]
    err = C.create_string_buffer(256)
    torch.cuda.synchronize()
    rc = lib.fsq_pair_counts_noalloc(d_l.data_ptr(), d_r.data_ptr(), m, runs_l, runs_r,
                                     out[0].data_ptr(), out[1].data_ptr(), out[2].data_ptr(), out[3].data_ptr(),
                                     torch.cuda.current_stream().cuda_stream, err, len(err))
    torch.cuda.synchronize()
    n = int(out[3].cpu().numpy().view(np.uint32)[0])
    k = min(n, m)
    res = [o.cpu().numpy().view(np.uint32)[:k] for o in out[:3]]
    return rc, err.value.decode(), res[0], res[1], res[2], n


def test_pair_counts_noalloc_matches_pair_counts(ctx, fsq, synth):
    import torch
    torch.cuda.set_device(0)
    lib = ctx.lib
    rows = synth.synth_rows(700, 6000, seed=41, founders=20, gaps=True)
    with ctx.msa_from_rows(rows) as msa, ctx.segment(msa, 12) as seg:
        assert seg.count >= 4
        pairs = [0, 1, seg.count // 2, seg.count - 2]
        # two rounds over the same pairs: every call after the first finds the table as the call before left it
        for round_no in range(2):
            for s in pairs:
                wl, wr, wc = seg.pair_counts(s)
                off_l, _, cls_l = seg.segment_sets(s)
                off_r, _, cls_r = seg.segment_sets(s + 1)
                runs_l, runs_r = len(off_l) - 1, len(off_r) - 1
                rc, msg, l, r, c, n = _call(lib, torch, cls_l, cls_r, runs_l, runs_r)
                assert rc == 0, msg
                assert n == len(wl), "round %d, pair %d: %d pairs, fsq_pair_counts has %d" % (round_no, s, n, len(wl))
                assert np.array_equal(l, wl) and np.array_equal(r, wr) and np.array_equal(c, wc), "round %d, pair %d" % (round_no, s)
                assert int(c.sum()) == rows.shape[0]


def test_pair_counts_noalloc_clears_what_the_last_call_left(ctx, fsq):
    """A large dense table first, a small sparse one right behind it: nothing of the first may show in the second."""
    import torch
    torch.cuda.set_device(0)
    lib = ctx.lib
    rng = np.random.default_rng(3)
    m = 50000
    big_l, big_r = rng.integers(0, 1000, m).astype(np.uint32), rng.integers(0, 1000, m).astype(np.uint32)
    small_l, small_r = (np.arange(m) % 7).astype(np.uint32), (np.arange(m) % 5).astype(np.uint32)

    def expect(a, b, rr):
        code, cnt = np.unique(a.astype(np.uint64) * rr + b, return_counts=True)
        return (code // rr).astype(np.uint32), (code % rr).astype(np.uint32), cnt.astype(np.uint32)

    for cls_l, cls_r, rl, rr in ((big_l, big_r, 1000, 1000), (small_l, small_r, 7, 5), (big_l, big_r, 1024, 1024), (small_l, small_r, 7, 5)):
        rc, msg, l, r, c, n = _call(lib, torch, cls_l, cls_r, rl, rr)
        assert rc == 0, msg
        el, er, ec = expect(cls_l, cls_r, rr)
        assert n == len(el)
        assert np.array_equal(l, el) and np.array_equal(r, er) and np.array_equal(c, ec)


def test_pair_counts_noalloc_refuses_what_does_not_fit(ctx, fsq):
    import torch
    torch.cuda.set_device(0)
    lib = ctx.lib
    cls = np.zeros(64, np.uint32)
    rc, msg, *_ = _call(lib, torch, cls, cls, 1025, 1024)
    assert rc == fsq.ERR_UNSUPPORTED and "FSQ_PAIR_COUNTS_NOALLOC_MAX_CELLS" in msg
    rc, msg, *_ = _call(lib, torch, cls, cls, 0, 4)
    assert rc == fsq.ERR_ARG and msg
    err = C.create_string_buffer(128)
    assert lib.fsq_pair_counts_noalloc(None, None, 4, 2, 2, None, None, None, None, None, err, len(err)) == fsq.ERR_ARG
    # a class index outside the stated number of runs: flagged on the device, the table is not written outside
    bad = np.array([0, 1, 2, 9], np.uint32)
    rc, msg, l, r, c, n = _call(lib, torch, bad, bad, 3, 3)
    assert rc == 0 and n == 0xffffffff
    # ... and the next call is clean again
    ok = np.array([0, 1, 2, 2], np.uint32)
    rc, msg, l, r, c, n = _call(lib, torch, ok, ok, 3, 3)
    assert rc == 0 and n == 3 and list(c) == [1, 1, 2]
