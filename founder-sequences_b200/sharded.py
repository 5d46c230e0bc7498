"""Column-sharded run of the long path: one rank per GPU, each holding a contiguous column range of ONE alignment.

The reference runs the path in a single process; this is the driver for the fsq_shard_* stages of
include/fsq.h.  It is written SPMD-style against two small interfaces:

  * an *engine* -- the per-rank stages.  `GpuShard` is the product (ctypes over libfsq_b200.so, device buffers wrapped as
    torch tensors for the collectives).  The tests have a CPU engine built on the oracle for the gloo runs.
  * a *communicator* -- `TorchComm` (torch.distributed: nccl on GPUs, gloo on CPU) or `ThreadComm` (ranks as threads
    of one process, for driving several shards through one GPU).

Exchanges, all small next to the scan: one all-gather of the rank summaries (8 m bytes each), per DP round one
broadcast per rank of its slice of the table (4 bytes per column) plus two scalar max-reductions, the three table
columns once at the end, and K + 1 (value, count) pairs per traceback position for the reduction.
"""
import ctypes as C
import os
import threading
import time

import numpy as np
import torch


class NotReducible(RuntimeError):
    pass


# ---- communicators ---------------------------------------------------------------------------------------------------

class TorchComm:
    """torch.distributed process group (already initialised by the caller)."""

    def __init__(self, group=None):
        import torch.distributed as dist
        self.dist = dist
        self.group = group
        self.rank = dist.get_rank(group)
        self.size = dist.get_world_size(group)

    def broadcast(self, tensor, src):
        if tensor.numel():
            self.dist.broadcast(tensor, src=src, group=self.group)

    def all_gather(self, tensor):
        out = [torch.empty_like(tensor) for _ in range(self.size)]
        self.dist.all_gather(out, tensor, group=self.group)
        return out

    def all_gather_flat(self, tensor):
        """[size, len] tensor of every rank's 1-D tensor (same length everywhere)."""
        out = torch.empty((self.size, tensor.numel()), dtype=tensor.dtype, device=tensor.device)
        if tensor.is_cuda:
            self.dist.all_gather_into_tensor(out.view(-1), tensor, group=self.group)
        else:
            parts = [out[r] for r in range(self.size)]
            self.dist.all_gather(parts, tensor, group=self.group)
        return out

    def fence(self, device):
        """The engine launches on its own stream: what a collective wrote must have landed before the next stage."""
        if getattr(device, "type", "cpu") == "cuda":
            torch.cuda.synchronize(device)

    def max(self, value, device):
        t = torch.tensor([int(value)], dtype=torch.int64, device=device)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX, group=self.group)
        return int(t.item())

    def max_many(self, values, device):
        """Element-wise maximum over the ranks of a few non-negative integers: one collective, one read-back."""
        t = torch.tensor([int(v) for v in values], dtype=torch.int64, device=device)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX, group=self.group)
        return [int(v) for v in t.tolist()]

    def gather_ints(self, values, device):
        """Every rank's tuple of integers, in rank order (one all-gather of a small tensor; no pickling)."""
        t = torch.tensor([int(v) for v in values], dtype=torch.int64, device=device)
        out = [torch.empty_like(t) for _ in range(self.size)]
        self.dist.all_gather(out, t, group=self.group)
        return [tuple(r) for r in torch.stack(out).tolist()]

    def gather_objects(self, obj):
        out = [None] * self.size
        self.dist.all_gather_object(out, obj, group=self.group)
        return out

    def barrier(self):
        self.dist.barrier(group=self.group)


class ThreadComm:
    """Ranks as threads of one process; collectives through a shared slot table and a barrier."""

    class Shared:
        def __init__(self, size):
            self.size = size
            self.barrier = threading.Barrier(size)
            self.slots = [None] * size

    def __init__(self, shared, rank):
        self.sh = shared
        self.rank = rank
        self.size = shared.size

    def _exchange(self, value):
        self.sh.slots[self.rank] = value
        self.sh.barrier.wait()
        vals = list(self.sh.slots)
        self.sh.barrier.wait()
        return vals

    def broadcast(self, tensor, src):
        vals = self._exchange(tensor if self.rank == src else None)
        if self.rank != src and tensor.numel():
            tensor.copy_(vals[src])
            if tensor.is_cuda:
                torch.cuda.synchronize(tensor.device)
        self.sh.barrier.wait()          # the source may not touch its tensor before everyone has copied

    def all_gather(self, tensor):
        vals = self._exchange(tensor)
        out = [v.clone() for v in vals]
        if tensor.is_cuda:
            torch.cuda.synchronize(tensor.device)
        self.sh.barrier.wait()
        return out

    def all_gather_flat(self, tensor):
        return torch.stack(self.all_gather(tensor))

    def fence(self, device):
        pass        # broadcast / all_gather above synchronise the device themselves

    def max(self, value, device):
        return max(int(v) for v in self._exchange(int(value)))

    def max_many(self, values, device):
        vals = self._exchange([int(v) for v in values])
        return [max(col) for col in zip(*vals)]

    def gather_ints(self, values, device):
        return [tuple(v) for v in self._exchange([int(v) for v in values])]

    def gather_objects(self, obj):
        return self._exchange(obj)

    def barrier(self):
        self.sh.barrier.wait()


# ---- the product engine ----------------------------------------------------------------------------------------------

class _DevView:
    """A device pointer as a torch tensor, through __cuda_array_interface__ (no copy)."""

    def __init__(self, ptr, shape, typestr):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False), "version": 2}


def _view(ptr, count, device, typestr="<i4"):
    if count == 0:
        return torch.empty(0, dtype=torch.int32, device=device)
    return torch.as_tensor(_DevView(ptr, (count,), typestr), device=device)


class GpuShard:
    """The fsq_shard_* stages of libfsq_b200.so for the columns held by `msa` (whose first column is its col0)."""

    def __init__(self, ctx, msa, first_column, n_total, L):
        self.ctx, self.lib, self.msa = ctx, ctx.lib, msa
        self.m, self.c0, self.c1, self.n, self.L = msa.m, first_column, first_column + msa.n, n_total, L
        self.device = torch.device("cuda", ctx.device)
        ctx.check(self.lib.fsq_msa_set_first_column(msa.h, first_column))
        h = C.c_void_p()
        ctx.check(self.lib.fsq_shard_create(ctx.h, msa.h, n_total, L, C.byref(h)))
        self.h = h

    def close(self):
        if self.h:
            self.lib.fsq_shard_free(self.h)
            self.h = None

    def local_summary(self):
        pi, delta = C.c_void_p(), C.c_void_p()
        self.ctx.check(self.lib.fsq_shard_local_summary(self.h, C.byref(pi), C.byref(delta)))
        return torch.stack([_view(pi.value, self.m, self.device), _view(delta.value, self.m, self.device)])

    def set_start(self, summaries, first_columns):
        """summaries: list of (2, m) int32 device tensors of the ranks before this one, in rank order."""
        count = len(summaries)
        if count == 0:
            self.ctx.check(self.lib.fsq_shard_set_start(self.h, None, None, None, 0))
            return
        pis = torch.stack([s[0] for s in summaries]).contiguous()
        deltas = torch.stack([s[1] for s in summaries]).contiguous()
        torch.cuda.synchronize(self.device)
        cols = np.ascontiguousarray(first_columns, np.uint64)
        self.ctx.check(self.lib.fsq_shard_set_start(self.h, C.c_void_p(pis.data_ptr()), C.c_void_p(deltas.data_ptr()),
                                                    cols.ctypes.data_as(C.c_void_p), count))

    def scan(self):
        self.ctx.check(self.lib.fsq_shard_scan(self.h))

    def dp_table(self):
        pm, pl, ps = C.c_void_p(), C.c_void_p(), C.c_void_p()
        n, f, l = C.c_uint64(), C.c_uint64(), C.c_uint64()
        self.ctx.check(self.lib.fsq_shard_dp_table(self.h, C.byref(pm), C.byref(pl), C.byref(ps), C.byref(n), C.byref(f), C.byref(l)))
        tabs = [_view(p.value, n.value, self.device) for p in (pm, pl, ps)]
        return tabs, (f.value, l.value)

    def dp_round(self):
        ch, rc = C.c_uint32(), C.c_uint32()
        self.ctx.check(self.lib.fsq_shard_dp_round(self.h, C.byref(ch), C.byref(rc)))
        return ch.value, rc.value

    def dp_sync(self, guess, reach):
        self.ctx.check(self.lib.fsq_shard_dp_sync(self.h, 1 if guess else 0, reach))

    def dp_single_chain(self):
        self.ctx.check(self.lib.fsq_shard_dp_single_chain(self.h))

    def dp_final(self):
        ov = C.c_uint32()
        self.ctx.check(self.lib.fsq_shard_dp_final(self.h, C.byref(ov)))
        return ov.value

    def grow_cap(self):
        self.ctx.check(self.lib.fsq_shard_grow_cap(self.h))

    def traceback(self):
        t, k = C.c_uint32(), C.c_uint32()
        self.ctx.check(self.lib.fsq_shard_traceback(self.h, C.byref(t), C.byref(k)))
        return t.value, k.value

    def get_traceback(self, count):
        lb, rb = np.zeros(count, np.uint64), np.zeros(count, np.uint64)
        mx, sz = np.zeros(count, np.uint32), np.zeros(count, np.uint32)
        p = lambda a: a.ctypes.data_as(C.c_void_p)
        self.ctx.check(self.lib.fsq_shard_get_traceback(self.h, p(lb), p(rb), p(mx), p(sz)))
        return lb, rb, mx, sz

    def position_range(self, c0, c1):
        f, l = C.c_uint32(), C.c_uint32()
        self.ctx.check(self.lib.fsq_shard_position_range(self.h, int(c0), int(c1), C.byref(f), C.byref(l)))
        return f.value, l.value

    def reduce_lists(self, count):
        pl = C.c_void_p()
        w, f, l = C.c_uint32(), C.c_uint32(), C.c_uint32()
        rc = self.lib.fsq_shard_reduce_lists(self.h, C.byref(pl), C.byref(w), C.byref(f), C.byref(l))
        if rc == 4:
            raise NotReducible(self.lib.fsq_last_error(self.ctx.h).decode())
        self.ctx.check(rc)
        row = 2 * w.value
        return _view(pl.value, count * row, self.device), row, (f.value, l.value)

    def reduce(self):
        n, f, l = C.c_uint32(), C.c_uint32(), C.c_uint32()
        self.ctx.check(self.lib.fsq_shard_reduce(self.h, C.byref(n), C.byref(f), C.byref(l)))
        return n.value, (f.value, l.value)

    def bounds(self, count):
        lb, rb, sz = np.zeros(count, np.uint64), np.zeros(count, np.uint64), np.zeros(count, np.uint32)
        p = lambda a: a.ctypes.data_as(C.c_void_p)
        self.ctx.check(self.lib.fsq_shard_bounds(self.h, p(lb), p(rb), p(sz)))
        return lb, rb, sz

    def states(self, own):
        cnt = own[1] - own[0]
        a, d = np.zeros((cnt, self.m), np.uint32), np.zeros((cnt, self.m), np.uint32)
        if cnt:
            self.ctx.check(self.lib.fsq_shard_states(self.h, a.ctypes.data_as(C.c_void_p), d.ctypes.data_as(C.c_void_p)))
        return a, d


# ---- the driver --------------------------------------------------------------------------------------------------------

class ShardedResult:
    """What every rank knows afterwards: the reduced segmentation (replicated), and the states of the reduced segments
    whose right bound lies in its own columns."""

    def __init__(self):
        self.K = 0
        self.count = 0                 # reduced segments
        self.traceback = None          # (lb, rb, max, size), replicated
        self.lb = self.rb = self.size = None
        self.own = (0, 0)              # reduced segments [first, last) held here
        self.a = self.d = None         # their states
        self.dp_rounds = 0
        self.cap_retries = 0


def _exchange_slices(comm, tensors, ranges, extras=None):
    """Every rank sends its own slice [first, last) of each tensor to all others (ranges: per rank).  ONE all-gather for
    all tensors (same dtype), the slices padded to the longest one and laid side by side -- a broadcast per rank costs a
    collective launch per rank and adds up the ranks' skew -- then the received slices are copied into place.
    extras: a few non-negative integers that ride along; returns their element-wise maximum over the ranks (one
    read-back instead of a second collective)."""
    width = max(l - f for f, l in ranges)
    extras = [int(v) for v in (extras or [])]
    if not tensors:
        return comm.max_many(extras, torch.device("cpu")) if extras else []
    t0 = tensors[0]
    k, e = len(tensors), len(extras)
    send = torch.zeros(k * width + e, dtype=t0.dtype, device=t0.device)
    f, l = ranges[comm.rank]
    for i, t in enumerate(tensors):
        send[i * width: i * width + (l - f)] = t[f:l]
    if e:
        send[k * width:] = torch.tensor(extras, dtype=t0.dtype)
    got = comm.all_gather_flat(send)                    # [size, k * width + e]
    for src, (f, l) in enumerate(ranges):
        if src != comm.rank and l > f:
            for i, t in enumerate(tensors):
                t[f:l] = got[src, i * width: i * width + (l - f)]
    if e:
        out = [int(v) for v in got[:, k * width:].max(dim=0).values.tolist()]      # the read-back also drains the copies above
    else:
        out = []
    comm.fence(t0.device)
    return out


def entry_range(c0, c1, n, L):
    """The DP entries computed by the rank that holds columns [c0, c1) of n: entry t belongs to column t + L - 1, and
    the rank with the last column also owns the final entry (fsq_shard_dp_table; shard_init in fsq_capi.cu)."""
    n_entries, dp_size = n - 2 * L + 1, n - L + 1
    first = min(max(0, c0 + 1 - L), n_entries)
    last = dp_size if c1 == n else min(max(0, c1 + 1 - L), n_entries)
    return first, last


def segment_sharded(engine, comm, max_rounds=64, guess_rounds=None, want_states=True, want_traceback=True, want_bounds=True):
    """Run the long path over the column shards held by the ranks of `comm`; `engine` is this rank's shard.
    Raises NotReducible when K == m (generate_context.cc:192-200)."""
    res = ShardedResult()
    if guess_rounds is None:
        guess_rounds = max_rounds           # a round is cheap, the take-turns fallback is not (fsq_segment's rule)
    dev = getattr(engine, "device", torch.device("cpu"))
    trace = [] if os.environ.get("FSQ_SHARD_TRACE") else None
    t_prev = [time.perf_counter()]

    def mark(name):
        # host-side stage clock (FSQ_SHARD_TRACE=1): the stages block, so no device synchronise is needed here
        if trace is not None:
            now = time.perf_counter()
            trace.append((name, 1e3 * (now - t_prev[0])))
            t_prev[0] = now

    # 1. local scans, then the boundary states: rank r composes the summaries of the ranks before it
    summary = engine.local_summary()
    mark("local_summary")
    # the column range of every rank travels with its summary: no separate exchange
    m = summary.shape[1]
    tail = torch.tensor([engine.c0, engine.c1], dtype=summary.dtype, device=summary.device)
    packed = comm.all_gather(torch.cat([summary.reshape(-1), tail]))
    summaries = [p[: 2 * m].view(2, m) for p in packed]
    cols = torch.stack([p[2 * m:] for p in packed]).cpu().tolist()
    firsts = [int(c[0]) for c in cols]
    mark("all_gather summaries")
    engine.set_start(summaries[: comm.rank], firsts[: comm.rank])
    mark("set_start")

    # 2. exact scan + DP until the table is a fixed point everywhere; the top-list bound grows if the DP says so
    while True:
        engine.scan()
        mark("scan")
        tabs, own = engine.dp_table()
        ranges = [entry_range(int(c[0]), int(c[1]), engine.n, engine.L) for c in cols]
        assert tuple(ranges[comm.rank]) == tuple(own), (ranges[comm.rank], own)
        rounds = 0
        sequential = False
        while True:
            if sequential:
                # fallback: the ranks take turns, each running its entries as one chain behind the complete prefix
                for r in range(comm.size):
                    if r == comm.rank:
                        engine.dp_round()
                    comm.broadcast(tabs[0][ranges[r][0]: ranges[r][1]], r)
                    comm.fence(dev)
                    engine.dp_sync(False, 0)
                rounds += 1
                break
            changed, reach = engine.dp_round()
            mark("dp_round")
            rounds += 1
            changed, reach = _exchange_slices(comm, tabs[:1], ranges, extras=(changed, reach))
            mark("dp exchange + flags")
            if comm.size == 1 and rounds > 1 and not changed:
                break
            if rounds > 1 and not changed:
                break
            if rounds >= max_rounds:
                engine.dp_single_chain()
                sequential = True
                engine.dp_sync(False, reach)
                continue
            engine.dp_sync(rounds <= guess_rounds, reach)
            mark("dp_sync")
        overflow = engine.dp_final()
        mark("dp_final")
        overflow, = _exchange_slices(comm, tabs, ranges, extras=(overflow,))
        mark("table exchange")
        res.dp_rounds += rounds
        if not overflow:
            break
        engine.grow_cap()
        res.cap_retries += 1

    # 3. traceback (replicated), reduction lists (exchanged), greedy reduction (replicated), own states
    T, K = engine.traceback()
    mark("traceback")
    res.K = K
    if want_traceback:
        res.traceback = engine.get_traceback(T)
        mark("get_traceback")
    if not K < engine.m:
        raise NotReducible("the maximum segment size %d is equal to the number of input sequences" % K)
    lists, row, own_pos = engine.reduce_lists(T)
    mark("reduce_lists")
    # the traceback is replicated: every rank can tell which rows every other rank fills
    pos_ranges = [engine.position_range(int(c[0]), int(c[1])) for c in cols]
    assert tuple(pos_ranges[comm.rank]) == tuple(own_pos), (pos_ranges[comm.rank], own_pos)
    _exchange_slices(comm, [lists], [(f * row, l * row) for f, l in pos_ranges])
    mark("lists exchange")
    count, own_seg = engine.reduce()
    mark("reduce")
    res.count = count
    if want_bounds:
        res.lb, res.rb, res.size = engine.bounds(count)
    res.own = own_seg
    if want_states:
        res.a, res.d = engine.states(own_seg)
    mark("bounds/states")
    res.trace = trace
    return res


def split_columns(n, parts):
    """Contiguous column ranges [c0, c1) for `parts` ranks, multiples of 32 columns except the last."""
    per = (n + parts - 1) // parts
    per = (per + 31) // 32 * 32
    cuts = [min(n, r * per) for r in range(parts + 1)]
    cuts[-1] = n
    return [(cuts[r], cuts[r + 1]) for r in range(parts)]
