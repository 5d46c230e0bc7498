"""Development aid: segmentation parity against the oracle for many-row shapes (multi-warp scan CTAs), repeated."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import numpy as np
from __graft_entry__ import load_package, load_synth
import oraclelib
fsq, synth = load_package(), load_synth()
orc = oraclelib.oracle()
ctx = fsq.Context(0)
for (m, n, L, F, seed) in [(2500, 6000, 5, 24, 4), (2500, 6000, 50, 24, 4), (5000, 4000, 50, 32, 3), (1500, 8000, 20, 16, 7), (8000, 1500, 30, 40, 9), (2500, 3000, 50, 24, 5), (20000, 1200, 100, 64, 11)]:
    rows = synth.synth_rows(m, n, seed=seed, founders=F, gaps=True)
    with orc.open(rows) as o:
        assert o.segment(L) == 0
        olb, orb, _, osz = o.reduced()
        oa, od = o.reduced_states()
        odp = o.dp()
        K = o.K()
    bad = 0
    for rep in range(4):
        ctx.set_option("keep_dp", 1)
        with ctx.msa_from_rows(rows) as msa, ctx.segment(msa, L) as seg:
            lb, rb, sz = seg.reduced_traceback()
            a, d = seg.reduced_pbwt_samples()
            dp = seg.dp_table(n - L + 1)
            ok = seg.max_segment_size == K and np.array_equal(lb, olb) and np.array_equal(rb, orb) and np.array_equal(sz, osz) and np.array_equal(a, oa) and np.array_equal(d, od)
            ok = ok and all(np.array_equal(x, y) for x, y in zip(dp, odp))
            bad += 0 if ok else 1
    print(m, n, L, "K", K, "segments", len(olb), "warps", ctx.stat("warps"), "chunk", ctx.stat("chunk"), "mismatching runs", bad, flush=True)
