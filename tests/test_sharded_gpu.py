"""Column-sharded long path on the GPU (fsq_segment_sharded): several shards of one alignment driven through one device
(ranks as threads, the library's ranks-as-threads communicator), against the oracle's unsharded answer.  The same
entry point runs under torchrun with the library's NCCL communicator (tests/sharded_torchrun.py, bench.py --gpus N)."""
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


class ShardResult:
    pass


def run_sharded(fsq, rows, L, parts, options=None):
    """`parts` ranks as threads of this process, each with its own context on device 0, through fsq_segment_sharded with
    the library's ranks-as-threads communicator."""
    sh = load_sharded()
    lib = fsq.load_library()
    m, n = rows.shape
    ranges = sh.shard_columns(n, parts)
    group = sh.LocalGroup(lib, parts)
    results, errors = [None] * parts, [None] * parts
    comp = np.zeros(256, np.uint8)
    for i, ch in enumerate(sorted(set(rows.tobytes()) | {0})):
        comp[ch] = i

    def work(rank):
        try:
            c0, c1 = ranges[rank]
            ctx = fsq.Context(0)
            for k, v in (options or {}).items():
                ctx.set_option(k, v)
            msa = ctx.msa_from_rows(np.ascontiguousarray(rows[:, c0:c1]), comp)
            comm = group.comm(ctx, rank)
            try:
                seg = sh.segment_sharded(ctx, msa, comm, c0, n, L)
                res = ShardResult()
                res.K, res.count, res.own = seg.max_segment_size, seg.count, seg.own
                res.traceback = seg.traceback()
                res.lb, res.rb, res.size = seg.reduced_traceback()
                res.a, res.d = seg.reduced_pbwt_samples()
                res.cap_retries = ctx.stat("cap_retries")
                res.dp_fallbacks = ctx.stat("dp_fallbacks")
                res.windows = ctx.stat("windows")
                res.tail_entries = ctx.stat("exchange_tail_entries")
                res.window_restarts = ctx.stat("dp_window_restarts")
                seg.free()
                results[rank] = res
            finally:
                comm.free()
                msa.free()
                ctx.close()
        except BaseException as e:      # noqa
            errors[rank] = e

    threads = [threading.Thread(target=work, args=(r,)) for r in range(parts)]
    for t in threads:
        t.start()
    for t in threads:
        t.join(timeout=300)
    group.free()
    for e in errors:
        if e is not None:
            raise e
    assert all(r is not None for r in results), "a rank did not finish"
    return results


def check_against_oracle(oracle, rows, L, results):
    sh = load_sharded()
    with oracle.open(rows) as o:
        assert o.segment(L) == 0
        olb, orb, _, osz = o.reduced()
        oa, od = o.reduced_states()
        tlb, trb, tmx, tsz = o.traceback()
        K = o.K()
    for res in results:
        assert res.K == K
        for x, y in zip(res.traceback, (tlb, trb, tmx, tsz)):
            assert np.array_equal(x, y)
        assert np.array_equal(res.lb, olb) and np.array_equal(res.rb, orb) and np.array_equal(res.size, osz)
    a, d = sh.assemble_states([(r.own, r.a, r.d) for r in results], len(olb), rows.shape[0])
    assert np.array_equal(a, oa) and np.array_equal(d, od)


@pytest.mark.parametrize("parts", [1, 2, 3, 5])
@pytest.mark.parametrize("m,n,L,C", [(48, 3000, 10, 96), (200, 2600, 20, 100), (1000, 1800, 8, 0), (2500, 1500, 30, 0)])
def test_sharded_matches_oracle(fsq, oracle, synth, parts, m, n, L, C):
    rows = synth.synth_rows(m, n, seed=m + L + parts, founders=7, gaps=True)
    opts = {"block_columns": C} if C else {}
    check_against_oracle(oracle, rows, L, run_sharded(fsq, rows, L, parts, opts))


def test_sharded_tiny_shards(fsq, oracle, synth):
    """Shards of 32 columns: shorter than 2L, so that shard boundaries fall inside the first 2L columns and inside the
    last L, and several ranks own no DP entry or no traceback position at all."""
    m, n, L = 64, 250, 24
    rows = synth.synth_rows(m, n, seed=5, founders=6, gaps=False)
    check_against_oracle(oracle, rows, L, run_sharded(fsq, rows, L, 8, {"block_columns": 16}))


def test_sharded_rejects_too_many_ranks(fsq, synth):
    rows = synth.synth_rows(16, 100, seed=1, founders=3)
    with pytest.raises(ValueError):
        load_sharded().shard_columns(100, 8)
    lib = fsq.load_library()
    import ctypes as C
    c0, c1 = C.c_uint64(), C.c_uint64()
    assert lib.fsq_shard_columns(100, 8, 7, C.byref(c0), C.byref(c1)) == fsq.ERR_ARG
    assert lib.fsq_shard_columns(100, 2, 1, C.byref(c0), C.byref(c1)) == 0 and (c0.value, c1.value) == (64, 100)


def test_sharded_small_cap_retries_and_sequential_fallback(fsq, oracle, synth):
    rows = synth.synth_rows(128, 1500, seed=9, founders=40, recomb_block=50, mu=0.01)
    res = run_sharded(fsq, rows, 6, 3, {"block_columns": 100, "cap": 2})
    assert res[0].cap_retries >= 1
    check_against_oracle(oracle, rows, 6, res)
    # two rounds at most: the ranks then take turns, each running its entries as one chain
    res = run_sharded(fsq, rows, 6, 3, {"block_columns": 100, "dp_chunks": 1000, "dp_max_rounds": 2, "dp_guess_rounds": 0})
    assert res[0].dp_fallbacks >= 1
    check_against_oracle(oracle, rows, 6, res)


def test_sharded_windows_and_short_halo(fsq, oracle, synth):
    """Top lists in windows of two blocks over one buffer (every round re-scans them; the reduction then runs from the
    divergence arrays of a re-scan), and a halo of one row, which the greedy runs outgrow."""
    rows = synth.synth_rows(96, 2400, seed=21, founders=5, gaps=True)
    res = run_sharded(fsq, rows, 8, 3, {"block_columns": 64, "window_blocks": 2, "halo": 1})
    assert res[0].windows > 1
    check_against_oracle(oracle, rows, 8, res)
    res = run_sharded(fsq, rows, 8, 3, {"block_columns": 64, "halo": 1})
    check_against_oracle(oracle, rows, 8, res)


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


def test_sharded_traceback_fallback_when_jumps_outgrow_the_window(fsq, oracle, synth):
    """The ranks trace the path through their own entries and export only the last `trace_window` of them as places where
    it can enter; a window of one entry is outgrown at once, and the run falls back to the traceback over the whole
    (gathered) table."""
    rows = synth.synth_rows(80, 2000, seed=33, founders=6, gaps=True)
    lib = fsq.load_library()
    res = run_sharded(fsq, rows, 9, 4, {"block_columns": 90, "trace_window": 1})
    check_against_oracle(oracle, rows, 9, res)
    res = run_sharded(fsq, rows, 9, 4, {"block_columns": 90})
    check_against_oracle(oracle, rows, 9, res)


@pytest.mark.parametrize("parts,back", [(2, 1024), (3, 1024), (5, 1024), (3, 64)])
def test_sharded_exchange_of_tails(fsq, oracle, synth, parts, back):
    """Per round the ranks exchange the last dp_back_bound + 64 entries of their slices (what the rank behind can look
    back at) and the window minima of their chains instead of whole slices; a look-back bound that the data outgrow
    (64 entries here) is noticed on the device and the run starts over with whole slices."""
    m, n, L = 120, 9000, 6
    rows = synth.synth_rows(m, n, seed=5 + parts, founders=10, gaps=True, recomb_block=300, mu=0.004)
    opts = {"block_columns": 250, "dp_chunks": 40, "dp_back_bound": back}
    res = run_sharded(fsq, rows, L, parts, opts)
    check_against_oracle(oracle, rows, L, res)
    seen = [(r.tail_entries, r.window_restarts) for r in res]
    if back >= 128:
        assert all(x == (back + 64, 0) for x in seen), seen
    else:
        assert all(x == (0, 1) for x in seen), seen
    res = run_sharded(fsq, rows, L, parts, dict(opts, exchange_tails=0))
    check_against_oracle(oracle, rows, L, res)
    assert all(r.tail_entries == 0 for r in res)
    # the top lists in windows of two blocks: every round re-scans them, the exchange follows the last window
    res = run_sharded(fsq, rows, L, parts, dict(opts, window_blocks=2))
    check_against_oracle(oracle, rows, L, res)
    assert res[0].windows > 1
    # several rounds without the guess: every round's exchange matters
    res = run_sharded(fsq, rows, L, parts, dict(opts, dp_guess_rounds=0, dp_max_rounds=200))
    check_against_oracle(oracle, rows, L, res)


@pytest.mark.parametrize("B", [1, 2, 3, 4, 7, 8, 11])
def test_sharded_start_states_for_every_shape_of_the_sweep(fsq, oracle, synth, B):
    """A rank behind the first takes the summary of its whole shard out of the up-sweep (a side chain over the complete
    nodes of the binary decomposition of its B blocks; the root itself when B is a power of two) and seeds the down-sweep
    with its true start state."""
    C, L = 64, 5
    n = 3 * B * C
    rows = synth.synth_rows(60, n, seed=40 + B, founders=6, gaps=True, recomb_block=53, mu=0.01)
    check_against_oracle(oracle, rows, L, run_sharded(fsq, rows, L, 3, {"block_columns": C}))
