#!/usr/bin/env python
"""bench.py — PBWT + segmentation throughput (MSA cells/s) of libfsq_b200 on B200, with the kernel roofline and
the CPU reference timed beside it.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c1|c2|...]

One "step" = one pass of the hot path (fsq_segment: local scans, stitch, exact scan + top lists, DP, traceback,
re-scan, greedy reduction) over one synthetic founder-mosaic alignment.  `value` times it with the packed
alignment already resident in HBM; `e2e` times the same through the C ABI from host rows (pinned), with the
host->device copy, the transpose/pack and the device->host copy of the result inside the timed region.

N > 1 (torchrun, one rank per GPU): ONE alignment of N x n columns, column-sharded -- rank r holds columns
[r n, (r + 1) n) -- through founder-sequences_b200/sharded.py (boundary summaries all-gathered, DP table slices
broadcast per round, reduction lists exchanged; NCCL).  Per-GPU work is fixed as N grows -> "scaling": "weak".

--impl reference times the reference's own CPU code for the path (oracle/_ref when it was built, else the C
restatement in oracle/), rank 0 only, on a bounded column sample of the same workload.
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


def cpu_baseline(rows_sample, L, want_ref=True):
    """Reference first pass (PBWT + histogram + DP per column; single-threaded in the reference) on a
    bounded column sample.  The only place besides tests/ and smoke() that touches oracle/."""
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


def sample_columns(m, kind_ref=True):
    # about 10-20 s of CPU work: the compiled reference does ~10-13 Mcells/s on one core, the C port ~3x that
    target_cells = 150e6 if kind_ref else 400e6
    return int(max(2000, min(400_000, target_cells / m)))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
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

    import numpy as np
    import torch
    from __graft_entry__ import load_package
    fsq = load_package()

    if args.impl == "reference" and rank != 0:
        return 0        # rank 0 alone runs and prints the reference arm
    return run(args, rank, local_rank, world, m, n, L, founders, gaps, seed, desc, np, torch, fsq)


def ncu_traffic(workload, phase):
    """dram__bytes_read.sum + dram__bytes_write.sum of the dominant scan launch from the committed `ncu --set full`
    capture (profiles/roofline_traffic.json; per launch, same workload), or None when no capture covers this case."""
    try:
        with open(os.path.join(ROOT, "profiles", "roofline_traffic.json")) as f:
            t = json.load(f)
        e = t[workload][phase]
        return e["dram_bytes"], e["source"]
    except (OSError, KeyError, ValueError):
        return None, None


def make_rows_sample(np, torch, fsq, m, nc, seed, founders, gaps, device_ok):
    if device_ok:
        with fsq.Context(0) as ctx:
            buf = torch.empty((m, nc), dtype=torch.uint8, device="cuda:0")
            ctx.synth_rows_device(buf.data_ptr(), nc, m, nc, seed, founders, gaps)
            return buf.cpu().numpy()
    from __graft_entry__ import load_synth
    return load_synth().synth_rows(m, nc, seed, founders, gaps)


def run(args, rank, local_rank, world, m, n, L, founders, gaps, seed, desc, np, torch, fsq):
    cells = m * n
    have_gpu = torch.cuda.is_available()

    if args.impl == "reference":
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import oraclelib
        use_ref = oraclelib.have_ref()
        nc = sample_columns(m, use_ref)
        rows = make_rows_sample(np, torch, fsq, m, nc, seed, founders, gaps, have_gpu)
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
            "config": {"workload": desc, "sample_columns": nc},
            "cpu_baseline": {"value": v, "unit": "cells/s", "cores": 1, "kind": vals[0]["kind"], "sample": vals[0]["sample"]},
            "e2e": {"value": v, "unit": "cells/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        }
        print(json.dumps(line))
        return 0

    if not have_gpu:
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
    ctx = fsq.Context(dev)
    if args.block_columns:
        ctx.set_option("block_columns", args.block_columns)
    if args.warps:
        ctx.set_option("warps", args.warps)
    if args.cap:
        ctx.set_option("cap", args.cap)
    for kv in args.opt:
        k, v = kv.split("=")
        ctx.set_option(k, int(v))
    sharded = None
    comm = None
    n_total = n * world
    if world > 1:
        from __graft_entry__ import load_sharded
        sharded = load_sharded()
        comm = sharded.TorchComm()
    # every rank holds its columns of the same alignment: the generator is a pure function of (seed, row, column)
    msa = ctx.msa_synth(m, n, seed, founders, gaps, col0=rank * n)

    def barrier():
        torch.cuda.synchronize(dev)
        if dist is not None:
            dist.barrier()
            torch.cuda.synchronize(dev)

    def step():
        if sharded is None:
            seg = ctx.segment(msa, L)
            info = (seg.max_segment_size, seg.count)
            seg.free()
            return info
        eng = sharded.GpuShard(ctx, msa, rank * n, n_total, L)
        try:
            res = sharded.segment_sharded(eng, comm, want_states=False, want_traceback=False, want_bounds=False)
        finally:
            eng.close()
        if getattr(res, "trace", None) and rank in (0, world - 1):
            sys.stderr.write("shard trace (rank %d, ms): " % rank + ", ".join("%s %.3f" % kv for kv in res.trace) + "\n")
            sys.stderr.write("host ms (accumulated): " + ", ".join("%s %.3f" % (k, ctx.stat("hms_" + k)) for k in ("traceback", "traceback_copy", "reduce", "reduce_walk", "reduced_states")) + "\n")
        return (res.K, res.count)

    for _ in range(args.warmup):
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
    for k in ("pass_a", "stitch", "prep", "pass_b", "dp", "traceback", "reduce", "pass_c"):
        phases[k] = ctx.stat("ms_" + k) / args.steps
    launches = ctx.stat("kernel_launches")
    cap = ctx.stat("cap")
    geom = {"block_columns": int(ctx.stat("block_columns")), "blocks": int(ctx.stat("blocks")), "warps_per_block": int(ctx.stat("warps")), "cap": int(cap), "cap_retries": int(ctx.stat("cap_retries")),
            "chunk": int(ctx.stat("chunk")), "dp_chains": int(ctx.stat("dp_chunks")), "dp_rounds_per_step": ctx.stat("dp_rounds") / args.steps,
            "states_kept_by_pass_b": bool(ctx.stat("store_states"))}

    # ---- end to end: host rows -> C ABI -> host results ----
    e2e = None
    if not args.no_e2e:
        rows_dev = torch.empty((m, n), dtype=torch.uint8, device="cuda:%d" % dev)
        ctx.synth_rows_device(rows_dev.data_ptr(), n, m, n, seed, founders, gaps, col0=rank * n)
        rows_host = torch.empty((m, n), dtype=torch.uint8, pin_memory=True)
        rows_host.copy_(rows_dev)
        del rows_dev
        torch.cuda.synchronize(dev)
        rows_np = rows_host.numpy()
        comp = np.zeros(256, np.uint8)
        for i, ch in enumerate(sorted(b"-ACGT" if gaps else b"ACGT")):
            comp[ch] = i + 1

        def e2e_step():
            if sharded is None:
                with ctx.msa_from_rows(rows_np, comp) as mm, ctx.segment(mm, L) as seg:
                    lb, rb, sz = seg.reduced_traceback()
                    a, d = seg.reduced_pbwt_samples_view()
                    return seg.max_segment_size, len(lb), lb.nbytes + rb.nbytes + sz.nbytes + a.nbytes + d.nbytes
            with ctx.msa_from_rows(rows_np, comp) as mm:
                eng = sharded.GpuShard(ctx, mm, rank * n, n_total, L)
                try:
                    res = sharded.segment_sharded(eng, comm, want_states=True)
                finally:
                    eng.close()
                return res.K, len(res.lb), res.lb.nbytes + res.rb.nbytes + res.size.nbytes + res.a.nbytes + res.d.nbytes

        e_steps = max(1, min(args.steps, 3))
        e2e_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(e_steps):
            einfo = e2e_step()
        barrier()
        esecs = time.perf_counter() - t0
        if dist is not None:
            t = torch.tensor([esecs], device="cuda:%d" % dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            esecs = float(t.item())
        assert einfo[:2] == info, (einfo, info)
        tb_bytes = 20 * einfo[1]            # the traceback read back for the reduction (an upper bound uses the segment count)
        e2e = {"value": world * cells * e_steps / esecs, "unit": "cells/s", "h2d_bytes_per_step": m * n + 256,
               "d2h_bytes_per_step": int(einfo[2]) + tb_bytes, "steps": e_steps,
               "timing": "host clock around the C-ABI calls (they block), barrier + device synchronise on both sides, max over ranks"}

    if rank != 0:
        msa.free()
        ctx.close()
        dist.destroy_process_group()
        return 0

    peak, peak_src = measured_peak()
    dom = max(("pass_a", "pass_b", "pass_c"), key=lambda k: phases[k])
    ach = cells * BYTES_PER_CELL / (phases[dom] * 1e-3) / 1e9 if phases[dom] > 0 else 0.0
    traffic, traffic_src = ncu_traffic(args.workload, dom)
    roofline = {"bound": "hbm", "kernel": "pbwt_cblock_kernel (%s)" % dom, "achieved": ach, "peak": peak, "unit": "GB/s",
                "frac": ach / peak, "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": cells * BYTES_PER_CELL, "launch_ms": phases[dom]}
    line = {
        "metric": METRIC, "value": world * cells * args.steps / secs, "unit": "cells/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * secs / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u32", "data": "synthetic",
        "config": {"workload": desc, "K": info[0], "segments": info[1], "inputs": "packed alignment %.0f MB per rank (> 126 MB L2), regenerated state per step" % (msa.m / 2 * n / 1e6),
                   "parallelism": "1 GPU" if world == 1 else "one alignment of %d columns, column-sharded over %d ranks (NCCL: summary all-gather, DP table broadcasts, reduction lists)" % (n_total, world),
                   "timing": "CUDA events on the library's stream around the timed steps, max over ranks; wall %.3f ms/step" % (1e3 * wall / args.steps), **geom},
        "phases_ms": phases, "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline,
    }
    if e2e:
        line["e2e"] = e2e
    if not args.no_cpu_baseline:
        use_ref = True
        nc = sample_columns(m, True)
        rows = make_rows_sample(np, torch, fsq, m, nc, seed, founders, gaps, True)
        line["cpu_baseline"] = cpu_baseline(rows, L, use_ref)
    if saved_stdout is not None:
        sys.stdout.flush()
        os.dup2(saved_stdout, 1)
    print(json.dumps(line))
    sys.stdout.flush()
    msa.free()
    ctx.close()
    if dist is not None:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
