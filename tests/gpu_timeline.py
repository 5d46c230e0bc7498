import sys,os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package
fsq=load_package()
ctx=fsq.Context(0)
msa=ctx.msa_synth(1000,1000000,2,16,True)
import time
for i in range(3):
    seg=ctx.segment(msa,20); seg.free()
ctx.reset_stats()
t0=time.perf_counter(); seg=ctx.segment(msa,20); seg.free(); t1=time.perf_counter()
print("wall ms",1e3*(t1-t0))
for k in ("pass_a","stitch","prep","pass_b","dp","traceback","pass_c","reduce","gather"):
    print("%-10s at %7.3f end %7.3f dur %6.3f"%(k,ctx.stat("at_"+k),ctx.stat("end_"+k),ctx.stat("ms_"+k)))
