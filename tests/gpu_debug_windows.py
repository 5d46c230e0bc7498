"""Debug driver (run under compute-sanitizer): windowed top lists, one rank and three ranks as threads."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
from __graft_entry__ import load_package, load_synth
import oraclelib
fsq, synth = load_package(), load_synth()
which = sys.argv[1] if len(sys.argv) > 1 else "single"
if which == "single":
    rows = synth.synth_rows(96, 3000, seed=3096, founders=7, gaps=True)
    with fsq.Context(0) as ctx:
        ctx.set_option("block_columns", 64); ctx.set_option("window_blocks", 3)
        with ctx.msa_from_rows(rows) as msa, ctx.segment(msa, 8) as seg:
            print("single ok", seg.max_segment_size, seg.count, ctx.stat("windows"))
else:
    import test_sharded_gpu as t
    rows = synth.synth_rows(96, 2400, seed=21, founders=5, gaps=True)
    res = t.run_sharded(fsq, rows, 8, 3, {"block_columns": 64, "window_blocks": 2, "halo": 1})
    print("multi ok", res[0].K, res[0].count)
