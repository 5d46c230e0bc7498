#!/usr/bin/env python
"""bench.py — PBWT + segmentation throughput (MSA cells/s) of libfsq_b200 on B200, with the kernel roofline and
the CPU reference timed beside it.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c1|c2|...]

One "step" = one pass of the hot path (fsq_segment: local scans, stitch, exact scan + top lists, DP, traceback,
re-scan, greedy reduction) over one synthetic founder-mosaic alignment.  `value` times it with the packed
alignment already resident in HBM; `e2e` times the same through the C ABI from host rows (pinned), with the
host->device copy, the transpose/pack and the device->host copy of the result inside the timed region.

N > 1 (torchrun, one rank per GPU): ONE alignment of N x n columns, column-sharded -- rank r holds columns
[r n, (r + 1) n) -- through the library's fsq_segment_sharded (its own NCCL communicator: boundary summaries,
DP table slices and reduction rows as all-gathers on the library's stream).  Per-GPU work is fixed as N grows ->
"scaling": "weak".  Before anything is timed the sharded result is checked against the unsharded one of rank 0
("sharded_parity").  The same run also times BASELINE config 3 (5,000 x 10M, L = 50) cut over the N GPUs -- total
work fixed as N grows -- as "config3_strong".

--impl reference times the reference's own CPU code for the path (oracle/_ref when it was built, else the C
restatement in oracle/), rank 0 only, on a bounded column sample of the same workload; it does not load the
product library or touch the GPU.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "MSA cells/sec (m x n) PBWT+segmentation"
BYTES_PER_CELL = 16.5     # read a,d (8) + write a,d (8) + one 4-bit symbol (0.5)

WORKLOADS = {
    # name: (m, n, L, founders, gaps, seed, description)
    "c1": (64, 100_000, 10, 8, False, 1, "synthetic 64 x 100k, sigma=4, L=10"),
    "c2": (1000, 1_000_000, 20, 16, True, 2, "synthetic 1,000 x 1M per GPU, sigma=5 (ACGT+gap), L=20"),
    "c2s": (1000, 200_000, 20, 16, True, 2, "synthetic 1,000 x 200k (1/5 of config 2), sigma=5, L=20"),
    "c5": (20000, 6_250_000, 100, 64, True, 5, "synthetic 20,000 x 6.25M per GPU (config 5 = 20,000 x 50M at 8 GPUs), L=100"),
    "c3s": (5000, 1_000_000, 50, 32, True, 3, "synthetic 5,000 x 1M (1/10 of config 3), sigma=5, L=50"),
    "c4s": (2500, 1_000_000, 50, 24, True, 4, "synthetic 2,500 x 1M (1/5 of config 4), L=50"),
    "c3": (5000, 1_250_000, 50, 32, True, 3, "synthetic 5,000 x 1.25M per GPU (config 3 = 5,000 x 10M at 8 GPUs), sigma=5, L=50"),
    "c3_n4": (5000, 2_500_000, 50, 32, True, 3, "synthetic 5,000 x 2.5M per GPU (config 3 = 5,000 x 10M at 4 GPUs), sigma=5, L=50"),
    "c3_n2": (5000, 5_000_000, 50, 32, True, 3, "synthetic 5,000 x 5M per GPU (config 3 = 5,000 x 10M at 2 GPUs), sigma=5, L=50"),
    "c3_whole": (5000, 10_000_000, 50, 32, True, 3, "synthetic 5,000 x 10M (config 3 whole on one GPU), sigma=5, L=50"),
    "c4": (2500, 5_000_000, 50, 24, True, 4, "synthetic 2,500 x 5M (config 4), L=50"),
    "c4_L5": (2500, 5_000_000, 5, 24, True, 4, "synthetic 2,500 x 5M (config 4), L=5"),
    "c4_L10": (2500, 5_000_000, 10, 24, True, 4, "synthetic 2,500 x 5M (config 4), L=10"),
    "c4_L200": (2500, 5_000_000, 200, 24, True, 4, "synthetic 2,500 x 5M (config 4), L=200"),
    "c5s": (20000, 250_000, 100, 64, True, 5, "synthetic 20,000 x 250k (1/200 of config 5), L=100"),
}


class ClockSampler:
    """SM clock and clock-event (throttle) reasons under load (B200_PROFILING.md's clocks line): `nvidia-smi -lms`
    beside the run (`smi`), or an in-process NVML thread (`nvml`).  Both disturb the run they watch -- see run()."""

    REASONS = [("hw_slowdown", 0x8), ("hw_thermal_slowdown", 0x40), ("sw_thermal_slowdown", 0x20), ("sw_power_cap", 0x4)]

    def __init__(self, index, kind="nvml", period=0.2):
        self.index, self.kind, self.period = index, kind, period
        self.proc = None
        self.nvml = None
        self.lines = []          # (time, sm_mhz, max_mhz, [reasons])
        self.stop_flag = threading.Event()
        self.thread = None

    def start(self):
        if self.kind == "none":
            return
        if self.kind == "nvml":
            try:
                import pynvml
                pynvml.nvmlInit()
                self.nvml = pynvml
                self.handle = self._nvml_handle(pynvml)
                self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM))
                self._sample_nvml()
                self.thread = threading.Thread(target=self._loop_nvml, daemon=True)
                self.thread.start()
                return
            except Exception:
                self.nvml = None
        q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q, "--format=csv,noheader,nounits", "-lms", str(int(1000 * self.period))],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read_smi, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _nvml_handle(self, pynvml):
        # the visible device `index` of this process: CUDA_VISIBLE_DEVICES may renumber
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        if vis:
            ids = [x.strip() for x in vis.split(",") if x.strip()]
            if self.index < len(ids):
                if ids[self.index].isdigit():
                    return pynvml.nvmlDeviceGetHandleByIndex(int(ids[self.index]))
                return pynvml.nvmlDeviceGetHandleByUUID(ids[self.index])
        return pynvml.nvmlDeviceGetHandleByIndex(self.index)

    def _sample_nvml(self):
        p = self.nvml
        sm = float(p.nvmlDeviceGetClockInfo(self.handle, p.NVML_CLOCK_SM))
        try:
            mask = int(p.nvmlDeviceGetCurrentClocksEventReasons(self.handle))
        except Exception:
            mask = int(p.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle))
        self.lines.append((time.perf_counter(), sm, self.max_mhz, [n for n, bit in self.REASONS if mask & bit]))

    def _loop_nvml(self):
        while not self.stop_flag.wait(self.period):
            try:
                self._sample_nvml()
            except Exception:
                return

    def _read_smi(self):
        for line in self.proc.stdout:
            f = [x.strip() for x in line.split(",")]
            if len(f) < 6:
                continue
            try:
                sm, mx = float(f[0]), float(f[1])
            except ValueError:
                continue
            names = [n for n, _ in self.REASONS]
            self.lines.append((time.perf_counter(), sm, mx, [n for n, v in zip(names, f[2:6]) if v.lower().startswith("active")]))

    def wait_ready(self, timeout=5.0):
        """nvidia-smi needs about a second before its first line."""
        t_end = time.perf_counter() + timeout
        while self.kind != "none" and not self.lines and time.perf_counter() < t_end:
            time.sleep(0.01)

    def window(self, t0, t1):
        """Keep the samples taken inside the timed region (the sampler is started before the warm-up so that its own
        start-up does not fall into the region); if the region was shorter than the sampling period, the nearest ones."""
        inside = [ln for ln in self.lines if t0 <= ln[0] <= t1]
        if not inside and self.lines:
            mid = 0.5 * (t0 + t1)
            inside = [min(self.lines, key=lambda x: abs(x[0] - mid))]
        self.lines = inside

    def stop(self):
        self.stop_flag.set()
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except subprocess.TimeoutExpired:
                self.proc.kill()
        if not self.lines:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no clock samples (%s)" % self.kind]}
        sm = sorted(ln[1] for ln in self.lines)
        reasons = set()
        for ln in self.lines:
            reasons.update(ln[3])
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": self.lines[-1][2], "reasons": sorted(reasons), "samples": len(sm),
                "source": "nvml" if self.nvml else "nvidia-smi"}


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md: 6.65 TB/s)"


# The reference arm times a bounded sample: the first REF_TARGET_CELLS / m columns of the workload, first pass only
# (PBWT + divergence histogram + DP per column -- single-threaded in the reference).  Both arms state it.
REF_TARGET_CELLS = 150e6


def sample_columns(m, n):
    return int(max(2000, min(n, 400_000, REF_TARGET_CELLS / m)))


def shared_config(args, m, n, L, desc):
    """The `config` both arms print, word for word: the workload, and the one place that says how the reference arm's
    bounded sample relates to it."""
    return {"workload": desc,
            "reference_arm": "reference timed on a %d-column prefix of the workload, first pass only (PBWT + divergence "
                             "histogram + DP per column; the sample re-scan and the reduction are not in it); the GPU arm "
                             "runs the whole path on all columns" % sample_columns(m, n)}


def cpu_baseline(rows_sample, L, want_ref=True):
    """Reference first pass on a bounded column sample.  The only place besides tests/ and smoke() that touches oracle/."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oraclelib
    m, nc = rows_sample.shape
    if want_ref and oraclelib.have_ref():
        chk, kind = oraclelib.ref(), "reference"
    else:
        chk, kind = oraclelib.oracle(), "port"
    with chk.open(rows_sample) as s:
        secs = s.time_first_pass(L, nc)
    cells = m * (nc - L)
    return {"value": cells / secs, "unit": "cells/s", "cores": 1, "kind": kind,
            "sample": "first pass (PBWT + divergence histogram + DP) over the first %d columns of the workload, %.1f s" % (nc, secs)}


def host_rows_sample(m, nc, seed, founders, gaps):
    """The sample on the host (numpy; the generator is a pure function of (seed, row, column)), cached on disk: the
    reference arm neither imports the product package nor touches the GPU."""
    import numpy as np
    cache = os.path.join(os.environ.get("TMPDIR", "/tmp"), "fsq_sample_%d_%d_%d_%d_%d.npy" % (m, nc, seed, founders, int(gaps)))
    if os.path.exists(cache):
        try:
            rows = np.load(cache)
            if rows.shape == (m, nc):
                return rows
        except Exception:
            pass
    from __graft_entry__ import load_synth
    rows = load_synth().synth_rows(m, nc, seed, founders, gaps)
    try:
        np.save(cache, rows)
    except OSError:
        pass
    return rows


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-parity", action="store_true", help="skip the sharded-vs-unsharded check of N > 1 runs")
    ap.add_argument("--no-strong", action="store_true", help="skip the config 3 strong-scaling leg")
    ap.add_argument("--block-columns", type=int, default=0)
    ap.add_argument("--warps", type=int, default=0)
    ap.add_argument("--cap", type=int, default=0)
    ap.add_argument("--clock-sampler", default="smi", choices=["nvml", "smi", "none"])
    ap.add_argument("--opt", action="append", default=[], help="library option key=value (experiments)")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    m, n, L, founders, gaps, seed, desc = WORKLOADS[args.workload]

    if args.impl == "reference":
        if rank != 0:
            return 0        # rank 0 alone runs and prints the reference arm
        return run_reference(args, m, n, L, founders, gaps, seed, desc)

    import numpy as np
    import torch
    from __graft_entry__ import load_package
    fsq = load_package()
    return run(args, rank, local_rank, world, m, n, L, founders, gaps, seed, desc, np, torch, fsq)


def run_reference(args, m, n, L, founders, gaps, seed, desc):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oraclelib
    use_ref = oraclelib.have_ref()
    nc = sample_columns(m, n)
    rows = host_rows_sample(m, nc, seed, founders, gaps)
    vals = []
    for i in range(args.warmup + args.steps):
        # each step is a bounded sample; warm-up steps use a shorter one
        cols = nc if i >= args.warmup else max(2 * L + 1, nc // 8)
        b = cpu_baseline(rows[:, :cols].copy(), L, use_ref)
        if i >= args.warmup:
            vals.append(b)
    v = sum(x["value"] for x in vals) / len(vals)
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": "cells/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * (m * (nc - L)) / v, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u32", "data": "synthetic",
        "config": shared_config(args, m, n, L, desc),
        "cpu_baseline": {"value": v, "unit": "cells/s", "cores": 1, "kind": vals[0]["kind"], "sample": vals[0]["sample"]},
        "e2e": {"value": v, "unit": "cells/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))
    return 0


def ncu_profile(workload, phase):
    """What the committed `ncu --set full` capture says about the dominant scan launch (profiles/roofline_traffic.json;
    per launch, same workload): DRAM bytes and warp instructions, or None when no capture covers this case."""
    try:
        with open(os.path.join(ROOT, "profiles", "roofline_traffic.json")) as f:
            t = json.load(f)
        return t[workload][phase]
    except (OSError, KeyError, ValueError):
        return None


def crc_rows(np, a, d):
    import zlib
    return np.array([zlib.crc32(d[i].tobytes(), zlib.crc32(a[i].tobytes())) for i in range(a.shape[0])], np.uint32)


def sharded_parity(np, torch, dist, fsq, sharded, ctx, comm, rank, world, dev, m, n, L, founders, gaps, seed):
    """Rank 0 runs the UNSHARDED fsq_segment over the same world x n columns; every rank then compares K, the reduced
    (lb, rb, size) and the crc32 of each of its own states with it.  Returns (verdict, digest)."""
    import zlib
    n_total = n * world
    ref = [None]
    if rank == 0:
        with fsq.Context(dev) as c0:
            with c0.msa_synth(m, n_total, seed, founders, gaps) as full, c0.segment(full, L) as seg:
                lb, rb, sz = seg.reduced_traceback()
                a, d = seg.reduced_pbwt_samples_view()
                ref[0] = (seg.max_segment_size, lb, rb, sz, crc_rows(np, a, d))
    dist.broadcast_object_list(ref, src=0)
    K, lb, rb, sz, crcs = ref[0]
    c0, c1 = sharded.shard_columns(n_total, world, rank)
    with ctx.msa_synth(m, c1 - c0, seed, founders, gaps, col0=c0) as msa, sharded.segment_sharded(ctx, msa, comm, c0, n_total, L) as seg:
        mlb, mrb, msz = seg.reduced_traceback()
        a, d = seg.reduced_pbwt_samples_view()
        f, l = seg.own
        ok = (seg.max_segment_size == K and np.array_equal(mlb, lb) and np.array_equal(mrb, rb) and np.array_equal(msz, sz)
              and np.array_equal(crc_rows(np, a, d), crcs[f:l]))
        own = (f, l)
    flags = [None] * world
    dist.all_gather_object(flags, (bool(ok), own))
    tiles = flags[0][1][0] == 0 and flags[-1][1][1] == len(lb) and all(flags[i][1][1] == flags[i + 1][1][0] for i in range(world - 1))
    verdict = "ok" if all(f[0] for f in flags) and tiles else "MISMATCH on ranks %s%s" % ([i for i, f in enumerate(flags) if not f[0]], "" if tiles else "; own ranges do not tile")
    digest = "%08x" % zlib.crc32(sz.tobytes(), zlib.crc32(rb.tobytes(), zlib.crc32(lb.tobytes(), K)))
    return verdict, {"K": int(K), "segments": int(len(lb)), "crc32_bounds": digest, "crc32_states": "%08x" % zlib.crc32(crcs.tobytes()),
                     "against": "fsq_segment of rank 0 over the same %d x %d alignment, unsharded" % (m, n_total)}


def run(args, rank, local_rank, world, m, n, L, founders, gaps, seed, desc, np, torch, fsq):
    cells = m * n
    if not torch.cuda.is_available():
        print(json.dumps({"error": "no CUDA device: libfsq_b200 has no CPU fallback"}))
        return 1

    dist = None
    saved_stdout = None
    if world > 1:
        # stdout carries the one JSON line: whatever the libraries print while the ranks run (the NCCL banner) goes to stderr
        sys.stdout.flush()
        saved_stdout = os.dup(1)
        os.dup2(2, 1)
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"      # the version banner goes to stdout, which carries the one JSON line
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    dev = local_rank
    torch.cuda.set_device(dev)

    def new_context():
        c = fsq.Context(dev)
        if args.block_columns:
            c.set_option("block_columns", args.block_columns)
        if args.warps:
            c.set_option("warps", args.warps)
        if args.cap:
            c.set_option("cap", args.cap)
        for kv in args.opt:
            k, v = kv.split("=")
            c.set_option(k, int(v))
        return c

    ctx = new_context()
    sharded = None
    comm = None
    n_total = n * world
    parity = None
    if world > 1:
        from __graft_entry__ import load_sharded
        sharded = load_sharded()
        comm = sharded.nccl_comm(ctx, dist)
        # the unsharded run of rank 0 holds the whole alignment (4 bits per cell) and the states of every column it can keep
        whole_bytes = m * n_total // 2
        if not args.no_parity and whole_bytes > 60e9:
            parity = ("skipped: the unsharded alignment (%.0f GB packed) does not fit one GPU; the same shape on a slice is checked by tests/gpu_c5_shape_check.py" % (whole_bytes / 1e9), None)
        elif not args.no_parity:
            verdict, digest = sharded_parity(np, torch, dist, fsq, sharded, ctx, comm, rank, world, dev, m, n, L, founders, gaps, seed)
            parity = (verdict, digest)
            if verdict != "ok":
                sys.stderr.write("sharded parity: %s\n" % verdict)
                if rank == 0:
                    if saved_stdout is not None:
                        sys.stdout.flush()
                        os.dup2(saved_stdout, 1)
                    print(json.dumps({"error": "sharded result differs from the unsharded one", "sharded_parity": verdict, "sharded_parity_digest": digest}))
                dist.destroy_process_group()
                return 2
            ctx.close()                 # the timed steps start from a context that has learned nothing yet (first_call_ms)
            comm.free()
            ctx = new_context()
            comm = sharded.nccl_comm(ctx, dist)
    # every rank holds its columns of the same alignment (the library's split rule: equal shares rounded up to 32 columns):
    # the generator is a pure function of (seed, row, column)
    my0, my1 = (0, n) if world == 1 else sharded.shard_columns(n_total, world, rank)
    my_n = my1 - my0
    msa = ctx.msa_synth(m, my_n, seed, founders, gaps, col0=my0)

    def barrier():
        torch.cuda.synchronize(dev)
        if dist is not None:
            dist.barrier()
            torch.cuda.synchronize(dev)

    def step(c=None, a=None, cm=None, first=None, total=None):
        c, a = c or ctx, a or msa
        if sharded is None:
            seg = c.segment(a, L if total is None else total[1])
        else:
            cm = cm or comm
            seg = sharded.segment_sharded(c, a, cm, my0 if first is None else first, n_total if total is None else total[0], L if total is None else total[1])
        info = (seg.max_segment_size, seg.count)
        seg.free()
        return info

    # the first call on a fresh context pays what later calls have learned (top-list bound retries, buffer allocation)
    barrier()
    t0 = time.perf_counter()
    info = step()
    barrier()
    first_call_ms = 1e3 * (time.perf_counter() - t0)
    first_call_retries = int(ctx.stat("cap_retries"))
    for _ in range(max(0, args.warmup - 1)):
        info = step()

    # Clocks and clock-event reasons under this very load: a pass of the same steps with the sampler beside it, right
    # before the timed pass.  Sampling inside the timed pass itself is measurably intrusive on this box (an nvidia-smi
    # sample stalls the process's launches for ~9 ms, an in-process NVML thread far longer: tests/micro/nvml_cost.py and
    # profiles/README.md), which on a 9 ms step is an error of 5-80 %, so the two are kept apart and both are reported.
    sampler = ClockSampler(dev, args.clock_sampler, period=0.1)
    sampler.start()
    sampler.wait_ready()
    barrier()
    s0 = time.perf_counter()
    n_sampled = max(args.steps, 40) if args.clock_sampler != "none" else 0
    if n_sampled and cells * world > 2e10:
        n_sampled = max(3, min(n_sampled, int(40 * 1e9 / cells)))
    for _ in range(n_sampled):
        info = step()
    barrier()
    s1 = time.perf_counter()
    sampler.window(s0, s1)
    clocks = sampler.stop()
    if n_sampled:
        clocks["sampled"] = "a pass of %d identical steps immediately before the timed one, sampler beside it" % n_sampled
        clocks["ms_per_step_while_sampling"] = 1e3 * (s1 - s0) / n_sampled

    if n_sampled:
        time.sleep(0.05)             # the sampler process is gone; two more untimed steps before the timed ones
        for _ in range(2):
            info = step()
    ctx.reset_stats()
    barrier()
    ctx.set_option("timer", 1)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        info = step()
    ctx.set_option("timer", 0)
    barrier()
    t1 = time.perf_counter()
    wall = t1 - t0
    secs = ctx.stat("ms_timer") * 1e-3        # CUDA events on the stream the kernels are launched on
    if dist is not None:
        t = torch.tensor([secs, wall], device="cuda:%d" % dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        secs, wall = float(t[0].item()), float(t[1].item())
    phases = {}
    for k in ("pass_a", "stitch", "prep", "pass_b", "dp", "exchange", "traceback", "reduce", "pass_c"):
        phases[k] = ctx.stat("ms_" + k) / args.steps
    launches = ctx.stat("kernel_launches")
    cap = ctx.stat("cap")
    geom = {"block_columns": int(ctx.stat("block_columns")), "blocks": int(ctx.stat("blocks")), "warps_per_block": int(ctx.stat("warps")), "cap": int(cap), "cap_retries": int(ctx.stat("cap_retries")),
            "chunk": int(ctx.stat("chunk")), "dp_chains": int(ctx.stat("dp_chunks")), "dp_rounds_per_step": ctx.stat("dp_rounds") / args.steps,
            "states_kept_by_pass_b": bool(ctx.stat("store_states")), "list_windows": int(ctx.stat("windows")),
            "host_syncs_per_step": ctx.stat("host_syncs") / args.steps,
            "host_ms_enqueueing_per_step": ctx.stat("hms_enqueue") / args.steps, "host_ms_waiting_per_step": ctx.stat("hms_wait") / args.steps,
            "first_call_ms": first_call_ms, "first_call_cap_retries": first_call_retries}

    # ---- end to end: host rows -> C ABI -> host results ----
    e2e = None
    e2e_cold = None
    if not args.no_e2e:
        rows_dev = torch.empty((m, my_n), dtype=torch.uint8, device="cuda:%d" % dev)
        ctx.synth_rows_device(rows_dev.data_ptr(), my_n, m, my_n, seed, founders, gaps, col0=my0)
        rows_host = torch.empty((m, my_n), dtype=torch.uint8, pin_memory=True)
        rows_host.copy_(rows_dev)
        del rows_dev
        torch.cuda.synchronize(dev)
        rows_np = rows_host.numpy()
        comp = np.zeros(256, np.uint8)
        for i, ch in enumerate(sorted(b"-ACGT" if gaps else b"ACGT")):
            comp[ch] = i + 1

        def e2e_step(c, cm):
            with c.msa_from_rows(rows_np, comp) as mm:
                if sharded is None:
                    seg = c.segment(mm, L)
                else:
                    seg = sharded.segment_sharded(c, mm, cm, my0, n_total, L)
                with seg:
                    lb, rb, sz = seg.reduced_traceback()
                    a, d = seg.reduced_pbwt_samples_view()
                    return seg.max_segment_size, len(lb), lb.nbytes + rb.nbytes + sz.nbytes + a.nbytes + d.nbytes

        e_steps = max(1, min(args.steps, 3))
        e2e_step(ctx, comm)
        barrier()
        t0 = time.perf_counter()
        for _ in range(e_steps):
            einfo = e2e_step(ctx, comm)
        barrier()
        esecs = time.perf_counter() - t0
        if dist is not None:
            t = torch.tensor([esecs], device="cuda:%d" % dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            esecs = float(t.item())
        assert einfo[:2] == info, (einfo, info)
        e2e = {"value": world * cells * e_steps / esecs, "unit": "cells/s", "h2d_bytes_per_step": m * n + 256,
               "d2h_bytes_per_step": int(einfo[2]) + 64, "steps": e_steps,
               "timing": "host clock around the C-ABI calls (they block), barrier + device synchronise on both sides, max over ranks"}
        if world == 1:
            # cold: what INTEGRATION.md's shim does once per run -- a new context, upload, the path, results, teardown
            torch.cuda.synchronize(dev)
            t0 = time.perf_counter()
            c2 = new_context()
            cinfo = e2e_step(c2, None)
            c2.close()
            torch.cuda.synchronize(dev)
            csecs = time.perf_counter() - t0
            assert cinfo[:2] == info, (cinfo, info)
            e2e_cold = {"value": cells / csecs, "unit": "cells/s", "ms": 1e3 * csecs,
                        "what": "fsq_create, fsq_msa_from_rows, fsq_segment, results to the host, fsq_destroy: nothing cached, nothing learned"}
        del rows_host, rows_np

    # ---- BASELINE config 3 (5,000 x 10M, L = 50) cut over the ranks: total work fixed as N grows ----
    strong = None
    if not args.no_strong and args.workload == "c2" and world in (1, 2, 4, 8):
        msa.free()
        msa = None
        sm, sn, sL, sf, sg, ss, sdesc = WORKLOADS["c3_whole"]
        c0, c1 = (0, sn) if world == 1 else sharded.shard_columns(sn, world, rank)
        c3 = new_context()
        cm3 = sharded.nccl_comm(c3, dist) if world > 1 else None
        a3 = c3.msa_synth(sm, c1 - c0, ss, sf, sg, col0=c0)
        barrier()
        t0 = time.perf_counter()
        sinfo = step(c3, a3, cm3, first=c0, total=(sn, sL))
        barrier()
        s_first = 1e3 * (time.perf_counter() - t0)
        s_steps = 3
        step(c3, a3, cm3, first=c0, total=(sn, sL))
        c3.reset_stats()
        barrier()
        c3.set_option("timer", 1)
        for _ in range(s_steps):
            sinfo = step(c3, a3, cm3, first=c0, total=(sn, sL))
        c3.set_option("timer", 0)
        barrier()
        ssecs = c3.stat("ms_timer") * 1e-3
        if dist is not None:
            t = torch.tensor([ssecs], device="cuda:%d" % dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ssecs = float(t.item())
        strong = {"workload": "BASELINE config 3 whole: 5,000 x 10M, L = 50, cut over %d GPU%s (total work fixed: strong scaling)" % (world, "" if world == 1 else "s"),
                  "value": sm * sn * s_steps / ssecs, "unit": "cells/s", "ms_per_step": 1e3 * ssecs / s_steps, "steps": s_steps,
                  "K": sinfo[0], "segments": sinfo[1], "first_call_ms": s_first, "cap": int(c3.stat("cap")),
                  "list_windows": int(c3.stat("windows")), "states_kept_by_pass_b": bool(c3.stat("store_states")),
                  "phases_ms": {k: c3.stat("ms_" + k) / s_steps for k in ("pass_a", "stitch", "prep", "pass_b", "dp", "exchange", "traceback", "reduce", "pass_c")}}
        a3.free()
        if cm3:
            cm3.free()
        c3.close()

    if rank != 0:
        if msa:
            msa.free()
        if comm:
            comm.free()
        ctx.close()
        dist.destroy_process_group()
        return 0

    peak, peak_src = measured_peak()
    dom = max(("pass_a", "pass_b", "pass_c"), key=lambda k: phases[k])
    ach = cells * BYTES_PER_CELL / (phases[dom] * 1e-3) / 1e9 if phases[dom] > 0 else 0.0
    prof = ncu_profile(args.workload, dom) or {}
    sm_count, sm_ghz = 148, (clocks.get("sm_mhz") or 1965.0) / 1e3
    issue_peak = sm_count * 4 * sm_ghz * 1e9                 # warp instructions per second: one per cycle per SM sub-partition
    roofline = {"bound": "hbm", "kernel": "pbwt_cblock_kernel (%s)" % dom, "achieved": ach, "peak": peak, "unit": "GB/s",
                "frac": ach / peak, "traffic": prof.get("dram_bytes"), "traffic_source": prof.get("source"), "peak_source": peak_src,
                "algorithmic_bytes_per_launch": cells * BYTES_PER_CELL, "launch_ms": phases[dom],
                "note": "frac is the algorithmic model (16.5 B per cell: the state read and written once per column) against the HBM peak; "
                        "the kernel keeps the state in shared memory, so its real DRAM traffic is `traffic` and what binds it is instruction issue: see `issue`"}
    if prof.get("dram_bytes") and phases[dom] > 0:
        roofline["dram_frac"] = prof["dram_bytes"] / (phases[dom] * 1e-3) / 1e9 / peak
    if prof.get("warp_instructions") and phases[dom] > 0:
        wi = prof["warp_instructions"]
        roofline["issue"] = {"warp_instructions_per_launch": wi, "per_32_cells": wi / (cells / 32.0), "peak_per_s": issue_peak,
                             "frac": wi / (phases[dom] * 1e-3) / issue_peak,
                             "peak_source": "148 SMs x 4 sub-partitions x one warp instruction per cycle at the sampled SM clock"}
    config = shared_config(args, m, n, L, desc)
    config.update({"K": info[0], "segments": info[1], "inputs": "packed alignment %.0f MB per rank (> 126 MB L2), regenerated state per step" % (m / 2 * n / 1e6),
                   "parallelism": "1 GPU" if world == 1 else "one alignment of %d columns, column-sharded over %d ranks (fsq_segment_sharded: NCCL all-gathers on the library's stream)" % (n_total, world),
                   "timing": "CUDA events on the library's stream around the timed steps, max over ranks; wall %.3f ms/step" % (1e3 * wall / args.steps), **geom})
    line = {
        "metric": METRIC, "value": world * cells * args.steps / secs, "unit": "cells/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * secs / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u32", "data": "synthetic",
        "config": config,
        "phases_ms": phases, "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline,
    }
    if parity:
        line["sharded_parity"] = parity[0]
        line["sharded_parity_digest"] = parity[1]
    if e2e:
        line["e2e"] = e2e
    if e2e_cold:
        line["e2e_cold"] = e2e_cold
    if strong:
        line["config3_strong"] = strong
    if not args.no_cpu_baseline:
        nc = sample_columns(m, n)
        line["cpu_baseline"] = cpu_baseline(host_rows_sample(m, nc, seed, founders, gaps), L, True)
    if saved_stdout is not None:
        sys.stdout.flush()
        os.dup2(saved_stdout, 1)
    print(json.dumps(line))
    sys.stdout.flush()
    if msa:
        msa.free()
    if comm:
        comm.free()
    ctx.close()
    if dist is not None:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
