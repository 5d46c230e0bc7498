"""Parity of the column-sharded path under torchrun + NCCL (one rank per GPU) against the oracle's unsharded answer.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tests/sharded_torchrun.py

tests/test_sharded_gpu.py launches this when the box has more than one GPU; with one GPU the same driver is covered with
ranks as threads."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


CASES = [(64, 4000, 10, 96), (1000, 6000, 20, 0), (300, 2500, 7, 50), (2500, 4000, 30, 0), (5000, 3200, 50, 0)]


def main():
    import torch
    import torch.distributed as dist
    from __graft_entry__ import load_package, load_sharded, load_synth
    import oraclelib
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    fsq, sh, synth = load_package(), load_sharded(), load_synth()
    for m, n, L, C in CASES:
        rows = synth.synth_rows(m, n, seed=m + L, founders=8, gaps=True)
        c0, c1 = sh.shard_columns(n, world, rank)
        ctx = fsq.Context(local)
        if C:
            ctx.set_option("block_columns", C)
        comm = sh.nccl_comm(ctx, dist)
        comp = np.zeros(256, np.uint8)
        for i, ch in enumerate(sorted(set(rows.tobytes()) | {0})):
            comp[ch] = i
        msa = ctx.msa_from_rows(np.ascontiguousarray(rows[:, c0:c1]), comp)
        seg = sh.segment_sharded(ctx, msa, comm, c0, n, L)
        K, own = seg.max_segment_size, seg.own
        tb = seg.traceback()
        lb, rb, sz = seg.reduced_traceback()
        a, d = seg.reduced_pbwt_samples()
        rounds = ctx.stat("dp_rounds")
        seg.free(); comm.free(); msa.free(); ctx.close()
        with oraclelib.oracle().open(rows) as o:
            assert o.segment(L) == 0
            olb, orb, _, osz = o.reduced()
            oa, od = o.reduced_states()
            assert K == o.K()
            for x, y in zip(tb, o.traceback()):
                assert np.array_equal(x, y)
        assert np.array_equal(lb, olb) and np.array_equal(rb, orb) and np.array_equal(sz, osz)
        f, l = own
        assert np.array_equal(a, oa[f:l]) and np.array_equal(d, od[f:l])
        owns = [None] * world
        dist.all_gather_object(owns, own)
        assert owns[0][0] == 0 and owns[-1][1] == len(olb) and all(owns[i][1] == owns[i + 1][0] for i in range(world - 1))
        if rank == 0:
            print("sharded ok: %d x %d, L = %d over %d ranks: K = %d, %d segments, %d DP rounds" % (m, n, L, world, K, len(olb), rounds))
    if rank == 0:
        print("sharded all ok %d" % len(CASES))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
