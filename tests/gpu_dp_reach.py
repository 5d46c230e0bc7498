import sys, os
sys.path.insert(0, os.getcwd())
from __graft_entry__ import load_package
fsq = load_package()
ctx = fsq.Context(0)
for name, m, n, L, F, seed, gaps in [("c1", 64, 100000, 10, 8, 1, False), ("c2", 1000, 1000000, 20, 16, 2, True), ("c3/10", 5000, 1000000, 50, 32, 3, True),
                                     ("c4L5", 2500, 1000000, 5, 24, 4, True), ("c4L10", 2500, 1000000, 10, 24, 4, True), ("c4L50", 2500, 1000000, 50, 24, 4, True), ("c4L200", 2500, 1000000, 200, 24, 4, True)]:
    for i in range(2):
        ctx.reset_stats()
        with ctx.msa_synth(m, n, seed, F, gaps) as msa, ctx.segment(msa, L) as seg:
            K = seg.max_segment_size
    print(name, "K", K, "reach", ctx.stat("dp_reach"), "rounds", ctx.stat("dp_rounds"), "chains", ctx.stat("dp_chunks"), "cap", ctx.stat("cap"), "ms_dp", ctx.stat("ms_dp"), flush=True)
