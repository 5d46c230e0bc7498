"""Development aid: where a many-row segmentation departs from the oracle, under several option sets."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import numpy as np
from __graft_entry__ import load_package, load_synth
import oraclelib
fsq, synth = load_package(), load_synth()
orc = oraclelib.oracle()
m, n, L, F, seed = 1500, 8000, 20, 16, 7
rows = synth.synth_rows(m, n, seed=seed, founders=F, gaps=True)
with orc.open(rows) as o:
    assert o.segment(L) == 0
    olb, orb, _, osz = o.reduced()
    oa, od = o.reduced_states()
    odp = o.dp()
    otb = o.traceback()
    K = o.K()
for opts in ({}, {"dp_chunks": 1}, {"kernel": 2}, {"kernel": 1}, {"store_states": 0}, {"block_columns": 8000}, {"dp_max_rounds": 2}, {"cap": 1500}):
    ctx = fsq.Context(0)
    ctx.set_option("keep_dp", 1)
    for k, v in opts.items():
        ctx.set_option(k, v)
    with ctx.msa_from_rows(rows) as msa, ctx.segment(msa, L) as seg:
        lb, rb, sz = seg.reduced_traceback()
        a, d = seg.reduced_pbwt_samples()
        dp = seg.dp_table(n - L + 1)
        tb = seg.traceback()
        msg = []
        if seg.max_segment_size != K: msg.append("K %d vs %d" % (seg.max_segment_size, K))
        for name, x, y in zip(("dp_lb", "dp_rb", "dp_max", "dp_size"), dp, odp):
            if not np.array_equal(x, y):
                i = int(np.nonzero(x != y)[0][0]); msg.append("%s first at %d of %d (%d vs %d), %d differ" % (name, i, len(x), x[i], y[i], int((x != y).sum())))
        if len(tb[0]) != len(otb[0]) or not all(np.array_equal(x, y) for x, y in zip(tb, otb)): msg.append("traceback (%d vs %d)" % (len(tb[0]), len(otb[0])))
        if len(lb) != len(olb) or not (np.array_equal(lb, olb) and np.array_equal(rb, orb) and np.array_equal(sz, osz)): msg.append("reduced (%d vs %d)" % (len(lb), len(olb)))
        elif not (np.array_equal(a, oa) and np.array_equal(d, od)): msg.append("states")
        print(opts, "rounds", ctx.stat("dp_rounds"), "->", "; ".join(msg) or "ok", flush=True)
    ctx.close()
