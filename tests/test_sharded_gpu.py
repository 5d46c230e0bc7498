"""Column-sharded long path on the GPU: several shards of one alignment driven through one device (ranks as threads, the
exchanges as device copies), against the oracle's unsharded answer.  The same driver runs under torchrun with NCCL
(tests/sharded_torchrun.py, bench.py --gpus N)."""
import threading

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def load_sharded():
    import importlib.util, os, sys
    from __graft_entry__ import load_package, PKG_DIR
    load_package()
    name = "founder_sequences_b200.sharded"
    if name in sys.modules:
        return sys.modules[name]
    spec = importlib.util.spec_from_file_location(name, os.path.join(PKG_DIR, "sharded.py"))
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


def run_sharded(fsq, rows, L, parts, options=None, cuts=None):
    sh = load_sharded()
    m, n = rows.shape
    ranges = cuts or sh.split_columns(n, parts)
    shared = sh.ThreadComm.Shared(len(ranges))
    results, errors = [None] * len(ranges), [None] * len(ranges)

    def work(rank):
        try:
            c0, c1 = ranges[rank]
            ctx = fsq.Context(0)
            for k, v in (options or {}).items():
                ctx.set_option(k, v)
            comp = np.zeros(256, np.uint8)
            for i, ch in enumerate(sorted(set(rows.tobytes()) | {0})):
                comp[ch] = i
            msa = ctx.msa_from_rows(np.ascontiguousarray(rows[:, c0:c1]), comp)
            eng = sh.GpuShard(ctx, msa, c0, n, L)
            try:
                results[rank] = sh.segment_sharded(eng, sh.ThreadComm(shared, rank))
            finally:
                eng.close()
                msa.free()
                ctx.close()
        except BaseException as e:      # noqa: a dead rank must not leave the others waiting on the barrier
            errors[rank] = e
            shared.barrier.abort()

    threads = [threading.Thread(target=work, args=(r,)) for r in range(len(ranges))]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    for e in errors:
        if e is not None and not isinstance(e, threading.BrokenBarrierError):
            raise e
    for e in errors:
        if e is not None:
            raise e
    return results


def check_against_oracle(oracle, rows, L, results):
    with oracle.open(rows) as o:
        assert o.segment(L) == 0
        olb, orb, _, osz = o.reduced()
        oa, od = o.reduced_states()
        tlb, trb, tmx, tsz = o.traceback()
        K = o.K()
    a = np.zeros_like(oa)
    d = np.zeros_like(od)
    covered = np.zeros(len(olb), bool)
    for res in results:
        assert res.K == K
        for x, y in zip(res.traceback, (tlb, trb, tmx, tsz)):
            assert np.array_equal(x, y)
        assert np.array_equal(res.lb, olb) and np.array_equal(res.rb, orb) and np.array_equal(res.size, osz)
        f, l = res.own
        assert not covered[f:l].any()
        covered[f:l] = True
        a[f:l], d[f:l] = res.a, res.d
    assert covered.all()
    assert np.array_equal(a, oa) and np.array_equal(d, od)


@pytest.mark.parametrize("parts", [1, 2, 3, 5])
@pytest.mark.parametrize("m,n,L,C", [(48, 3000, 10, 96), (200, 2600, 20, 100), (1000, 1800, 8, 0), (2500, 1500, 30, 0)])
def test_sharded_matches_oracle(fsq, oracle, synth, parts, m, n, L, C):
    rows = synth.synth_rows(m, n, seed=m + L + parts, founders=7, gaps=True)
    opts = {"block_columns": C} if C else {}
    check_against_oracle(oracle, rows, L, run_sharded(fsq, rows, L, parts, opts))


def test_sharded_uneven_cuts_and_tiny_shards(fsq, oracle, synth):
    """Shards shorter than L, a shard boundary inside the first 2L columns, and a one-column last shard."""
    m, n, L = 64, 1500, 12
    rows = synth.synth_rows(m, n, seed=5, founders=6, gaps=False)
    cuts = [(0, 7), (7, 20), (20, 700), (700, 1499), (1499, 1500)]
    check_against_oracle(oracle, rows, L, run_sharded(fsq, rows, L, len(cuts), {"block_columns": 50}, cuts))


def test_sharded_small_cap_retries_and_sequential_fallback(fsq, oracle, synth):
    rows = synth.synth_rows(128, 1500, seed=9, founders=40, recomb_block=50, mu=0.01)
    res = run_sharded(fsq, rows, 6, 3, {"block_columns": 100, "cap": 2})
    assert res[0].cap_retries >= 1
    check_against_oracle(oracle, rows, 6, res)
    sh = load_sharded()
    # max_rounds = 2 forces the rank-by-rank fallback
    orig = sh.segment_sharded
    try:
        sh.segment_sharded = lambda e, c, **kw: orig(e, c, max_rounds=2, guess_rounds=0)
        res = run_sharded(fsq, rows, 6, 3, {"block_columns": 100, "dp_chunks": 1000})
    finally:
        sh.segment_sharded = orig
    check_against_oracle(oracle, rows, 6, res)


def test_sharded_under_torchrun_nccl():
    """One rank per GPU over NCCL, when the box has more than one GPU."""
    import os, subprocess, sys, torch
    n_gpus = torch.cuda.device_count()
    if n_gpus < 2:
        pytest.skip("one GPU: the NCCL run needs at least two")
    here = os.path.dirname(os.path.abspath(__file__))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(min(n_gpus, 8)),
           "--master-addr", "127.0.0.1", "--master-port", "29533", os.path.join(here, "sharded_torchrun.py")]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:]
    sys.path.insert(0, here)
    from sharded_torchrun import CASES
    assert r.stdout.count("sharded ok:") == len(CASES), r.stdout[-3000:]
    assert "sharded all ok %d" % len(CASES) in r.stdout, r.stdout[-3000:]
