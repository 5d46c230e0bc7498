"""Worker for tests/test_sharded_gloo.py: one rank of a world_size-N gloo group running the host side of a column-sharded
call (founder-sequences_b200/sharded.py) on CPU.  The sharded run itself is CUDA code behind fsq_segment_sharded and has
no CPU path; what runs here is everything a launcher does around it -- the split rule (the library's own
fsq_shard_columns, a plain host function, against its Python mirror), the hand-over of the communicator's unique id
through the process group, and the assembly of the ranks' results -- with the oracle's unsharded answer cut the way the
library cuts it standing in for the GPU results."""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def run_rank(rank, world, port, case, out_dir):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import oraclelib
        from __graft_entry__ import load_package, load_sharded, load_synth
        fsq, sh = load_package(), load_sharded()
        lib = fsq.load_library()
        m, n, L, seed = case
        rows = load_synth().synth_rows(m, n, seed=seed, founders=4, gaps=True, recomb_block=23, mu=0.01, gap=0.01)

        # 1. the split rule: the library's and the mirror's agree, the ranges tile [0, n)
        c0, c1 = C.c_uint64(), C.c_uint64()
        assert lib.fsq_shard_columns(n, world, rank, C.byref(c0), C.byref(c1)) == 0
        assert (c0.value, c1.value) == sh.shard_columns(n, world, rank)
        ranges = [None] * world
        dist.all_gather_object(ranges, (c0.value, c1.value))
        assert ranges[0][0] == 0 and ranges[-1][1] == n and all(ranges[i][1] == ranges[i + 1][0] for i in range(world - 1))

        # 2. the unique id travels from rank 0 to everyone through the group
        def make_id(ident):
            ident[:] = np.arange(128, dtype=np.uint8) ^ 0x5A

        def broadcast(ident):
            t = torch.from_numpy(ident)
            dist.broadcast(t, src=0)

        ident = sh.exchange_unique_id(make_id, rank, broadcast)
        assert np.array_equal(ident, np.arange(128, dtype=np.uint8) ^ 0x5A)

        # 3. every rank holds the states of the reduced segments whose right bound lies in its columns; together they tile
        with oraclelib.oracle().open(rows) as o:
            assert o.segment(L) == 0
            olb, orb, _, osz = o.reduced()
            oa, od = o.reduced_states()
        f = int(np.searchsorted(orb, c0.value, side="right"))
        l = int(np.searchsorted(orb, c1.value, side="right"))
        parts = [None] * world
        dist.all_gather_object(parts, ((f, l), oa[f:l], od[f:l]))
        a, d = sh.assemble_states(parts, len(olb), m)
        assert np.array_equal(a, oa) and np.array_equal(d, od)
        try:
            sh.assemble_states(parts[:-1], len(olb), m)
            raise AssertionError("a missing rank went unnoticed")
        except ValueError:
            pass
        np.save(os.path.join(out_dir, "ok%d.npy" % rank), np.array([f, l]))
    finally:
        dist.destroy_process_group()
