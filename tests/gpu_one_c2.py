"""Two config-2 calls (1,000 x 1M, L = 20) and nothing else: the target of ncu captures of single kernels."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package

fsq = load_package()
ctx = fsq.Context(0)
for i in range(2):
    with ctx.msa_synth(1000, 1000000, 2, 16, True) as msa, ctx.segment(msa, 20) as seg:
        print("K", seg.max_segment_size, "segments", seg.count, flush=True)
