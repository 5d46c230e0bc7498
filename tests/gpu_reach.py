import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package
fsq = load_package()
ctx = fsq.Context(0)
for (m, n, L, F, seed) in [(1000, 1000000, 20, 16, 2), (5000, 400000, 50, 32, 3), (2500, 500000, 5, 24, 4), (2500, 500000, 200, 24, 4), (64, 100000, 10, 8, 1)]:
    msa = ctx.msa_synth(m, n, seed, F, True)
    ctx.reset_stats()
    seg = ctx.segment(msa, L)
    print(m, n, L, "K", seg.max_segment_size, "segments", seg.count, "reach", ctx.stat("dp_reach"), "rounds", ctx.stat("dp_rounds"), "cap", ctx.stat("cap"), "retries", ctx.stat("cap_retries"),
          "ms: a %.2f st %.2f b %.2f dp %.2f" % (ctx.stat("ms_pass_a"), ctx.stat("ms_stitch"), ctx.stat("ms_pass_b"), ctx.stat("ms_dp")))
    seg.free(); msa.free()
