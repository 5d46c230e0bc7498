"""Development aid: DP rounds and time against the chaining options on a many-row alignment (large K)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package
fsq = load_package()
ctx = fsq.Context(0)
m, n, L, F, seed = [int(x) for x in (sys.argv[1:6] if len(sys.argv) >= 6 else (20000, 100000, 100, 64, 5))]
msa = ctx.msa_synth(m, n, seed, F, True)
defaults = {"dp_chunks": 0, "dp_max_rounds": 0, "dp_guess_rounds": -1}
for opts in ({}, {"dp_max_rounds": 16, "dp_guess_rounds": 8}, {"dp_guess_rounds": 8}, {"dp_chunks": 16}, {"dp_chunks": 40}, {"dp_chunks": 1}):
    for k, v in defaults.items():
        ctx.set_option(k, opts.get(k, v))
    for rep in range(2):
        r0, d0, f0 = ctx.stat("dp_rounds"), ctx.stat("ms_dp"), ctx.stat("dp_fallbacks")
        with ctx.segment(msa, L) as seg:
            K = seg.max_segment_size
    print(opts, "K", K, "cap", ctx.stat("cap"), "chains", ctx.stat("dp_chunks"), "rounds", ctx.stat("dp_rounds") - r0, "fallbacks", ctx.stat("dp_fallbacks") - f0,
          "dp ms %.2f" % (ctx.stat("ms_dp") - d0), flush=True)
