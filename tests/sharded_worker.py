"""Worker for tests/test_sharded_gloo.py: one rank of a world_size-N gloo group driving the column-sharded protocol
(founder-sequences_b200/sharded.py) with the CPU stand-in engine, checked against the oracle's unsharded answer."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def load_sharded():
    import importlib.util
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


def run_rank(rank, world, port, case, out_dir):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import oraclelib
        from __graft_entry__ import load_synth
        from cpu_shard import CpuShard
        sh = load_sharded()
        m, n, L, seed, cuts = case
        rows = load_synth().synth_rows(m, n, seed=seed, founders=4, gaps=True, recomb_block=23, mu=0.01, gap=0.01)
        chk = oraclelib.oracle()
        with chk.open(rows) as o:
            comp = o.alphabet()
            status = o.segment(L)
            ranges = cuts or sh.split_columns(n, world)
            c0, c1 = ranges[rank]
            eng = CpuShard(rows[:, c0:c1], comp, c0, n, L, chk)
            comm = sh.TorchComm()
            try:
                res = sh.segment_sharded(eng, comm)
            except sh.NotReducible:
                assert status == 2
                np.save(os.path.join(out_dir, "ok%d.npy" % rank), np.array([1]))
                return
            assert status == 0
            olb, orb, _, osz = o.reduced()
            oa, od = o.reduced_states()
            for x, y in zip(res.traceback, o.traceback()):
                assert np.array_equal(x, y), (x, y)
            assert res.K == o.K()
            assert np.array_equal(res.lb, olb) and np.array_equal(res.rb, orb) and np.array_equal(res.size, osz)
            f, l = res.own
            assert np.array_equal(res.a, oa[f:l]) and np.array_equal(res.d, od[f:l])
            owns = comm.gather_objects(res.own)
            assert owns[0][0] == 0 and owns[-1][1] == len(olb) and all(owns[i][1] == owns[i + 1][0] for i in range(world - 1))
            np.save(os.path.join(out_dir, "ok%d.npy" % rank), np.array([res.dp_rounds]))
    finally:
        dist.destroy_process_group()
