"""Development aid: how long the NVML queries of bench.py's clock sampler take, idle and while the library runs."""
import os, sys, time, threading
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import pynvml
pynvml.nvmlInit()
h = pynvml.nvmlDeviceGetHandleByIndex(0)
calls = {
    "clock_sm": lambda: pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM),
    "max_clock_sm": lambda: pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM),
    "event_reasons": lambda: pynvml.nvmlDeviceGetCurrentClocksEventReasons(h),
    "power": lambda: pynvml.nvmlDeviceGetPowerUsage(h),
}
def bench(tag):
    for k, f in calls.items():
        f()
        t0 = time.perf_counter()
        for _ in range(5):
            f()
        print("%-6s %-14s %8.3f ms per call" % (tag, k, 1e3 * (time.perf_counter() - t0) / 5), flush=True)
bench("idle")
from __graft_entry__ import load_package
fsq = load_package()
ctx = fsq.Context(0)
msa = ctx.msa_synth(1000, 1000000, 2, 16, True)
stop = False
def load():
    while not stop:
        seg = ctx.segment(msa, 20); seg.free()
t = threading.Thread(target=load); t.start()
time.sleep(0.3)
bench("load")
stop = True; t.join()
