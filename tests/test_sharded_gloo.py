"""The host side of a column-sharded call on CPU: world_size 2 and 3 over gloo (tests/sharded_worker.py) -- the split rule,
the hand-over of the communicator id through the process group, the assembly of the ranks' results.  The sharded run
itself is CUDA code (fsq_segment_sharded; tests/test_sharded_gpu.py drives it with ranks as threads on one GPU,
tests/sharded_torchrun.py with one rank per GPU over NCCL).  Also: the boundary stitch the sharding
rests on, as a plain-Python model against straight scans."""
import os
import socket

import numpy as np
import pytest
import torch.multiprocessing as mp

from sharded_worker import run_rank


def free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("world,case", [
    (2, (10, 260, 6, 1)),
    (2, (14, 300, 9, 2)),
    (3, (12, 330, 5, 3)),
    (3, (9, 200, 4, 4)),
])
def test_sharded_host_side_over_gloo(tmp_path, world, case):
    port = free_port()
    mp.spawn(run_rank, args=(world, port, case, str(tmp_path)), nprocs=world, join=True)
    for r in range(world):
        assert os.path.exists(tmp_path / ("ok%d.npy" % r))


def test_stitch_composition_is_associative():
    """compose() of cpu_shard.py against straight scans, for random cut points."""
    from cpu_shard import compose, pbwt_step
    rng = np.random.default_rng(7)
    for _ in range(20):
        m, n = int(rng.integers(2, 9)), int(rng.integers(3, 30))
        sym = rng.integers(0, 3, size=(m, n))
        cuts = sorted(set([0, n] + [int(x) for x in rng.integers(0, n, size=3)]))
        a, d = np.arange(m), np.zeros(m, np.int64)
        truth = {0: (a, d)}
        for k in range(n):
            a, d = pbwt_step(a, d, sym[:, k], k + 1)
            truth[k + 1] = (a, d)
        summaries = []
        for s, e in zip(cuts[:-1], cuts[1:]):
            a, d = np.arange(m), np.full(m, s)
            for k in range(s, e):
                a, d = pbwt_step(a, d, sym[:, k], k + 1)
            summaries.append((s, a, d))
        state = truth[0]
        for s, pi, delta in summaries:
            state = compose(state[0], state[1], pi, delta, s)
        assert np.array_equal(state[0], truth[n][0]) and np.array_equal(state[1], truth[n][1])
        # right-to-left grouping gives the same state
        acc = summaries[-1]
        for s, pi, delta in reversed(summaries[:-1]):
            acc = (s,) + compose(pi, delta, acc[1], acc[2], acc[0])
        state = compose(truth[0][0], truth[0][1], acc[1], acc[2], 0)
        assert np.array_equal(state[0], truth[n][0]) and np.array_equal(state[1], truth[n][1])
