"""Development aid: the three cases of tests/sharded_torchrun.py with 8 ranks as threads on one GPU."""
import os, sys, traceback
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from __graft_entry__ import load_package, load_synth
import oraclelib
import test_sharded_gpu as T
fsq, synth = load_package(), load_synth()
for parts in (8, 7):
    for m, n, L, C in [(64, 4000, 10, 96), (1000, 6000, 20, 0), (300, 2500, 7, 50)]:
        rows = synth.synth_rows(m, n, seed=m + L, founders=8, gaps=True)
        try:
            res = T.run_sharded(fsq, rows, L, parts, {"block_columns": C} if C else None)
            T.check_against_oracle(oraclelib.oracle(), rows, L, res)
            print("ok", parts, m, n, L, C, "rounds", res[0].dp_rounds)
        except BaseException:
            print("FAILED", parts, m, n, L, C)
            traceback.print_exc()
