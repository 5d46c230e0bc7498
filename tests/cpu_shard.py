"""TEST INFRASTRUCTURE: a CPU stand-in for the per-rank engine of founder-sequences_b200/sharded.py, so that the
column-sharded protocol (summary composition, DP table exchange and fixed point, reduction lists) can run under gloo
on a machine without a GPU.  Plain Python/numpy restatements for tiny inputs, each following SURVEY.md Appendix A
(A.1 step, A.3/DP(k), A.4 via the oracle's rmq, A.6 reduction, A.10 stitch); the product never imports this."""
import numpy as np
import torch


def pbwt_step(a, d, sym_col, marker):
    """A.1 (lib/libbio/include/libbio/pbwt.hh:66-142): one column; marker = k + 1."""
    m = len(a)
    order = sorted(range(m), key=lambda i: sym_col[a[i]])          # stable: buckets in symbol order, a-order inside
    na, nd = np.empty(m, np.int64), np.empty(m, np.int64)
    last = {}
    dst_of = {i: dst for dst, i in enumerate(order)}
    for i in range(m):
        c = sym_col[a[i]]
        dst = dst_of[i]
        na[dst] = a[i]
        if c not in last:
            nd[dst] = marker
        else:
            nd[dst] = max(d[last[c] + 1: i + 1])
        last[c] = i
    return na, nd


def compose(in_a, in_d, pi, delta, s_mark):
    """A.10: state at e = (state at s) o (local summary of [s, e)); also composes two summaries."""
    m = len(pi)
    cls = np.empty(m, np.int64)
    start_d = []
    c = -1
    for i in range(m):
        if i == 0 or delta[i] > s_mark:
            c += 1
            start_d.append(delta[i])
        cls[pi[i]] = c
    counts = np.bincount(cls, minlength=c + 1)
    pos = np.concatenate([[0], np.cumsum(counts)[:-1]])
    out_a, out_d = np.empty(m, np.int64), np.empty(m, np.int64)
    prev = {}
    for i, x in enumerate(in_a):
        k = cls[x]
        dst = pos[k]
        pos[k] += 1
        out_a[dst] = x
        out_d[dst] = start_d[k] if k not in prev else max(in_d[prev[k] + 1: i + 1])
        prev[k] = i
    return out_a, out_d


class CpuShard:
    def __init__(self, rows_local, comp, c0, n_total, L, checker):
        self.sym = comp[rows_local].astype(np.int64)          # (m, n_local)
        self.m, self.nl = self.sym.shape
        self.c0, self.c1, self.n, self.L = c0, c0 + self.nl, n_total, L
        self.chk = checker                                     # oraclelib.Checker: rmq.hh emulation
        self.device = torch.device("cpu")
        self.dp_size = n_total - L + 1
        self.n_entries = n_total - 2 * L + 1
        lo, hi = max(0, c0 + 1 - L), max(0, self.c1 + 1 - L)
        self.t_first, self.t_last = min(lo, self.n_entries), min(hi, self.n_entries)

    def close(self):
        pass

    # ---- scans ----
    def local_summary(self):
        a, d = np.arange(self.m), np.full(self.m, self.c0)
        for j in range(self.nl):
            a, d = pbwt_step(a, d, self.sym[:, j], self.c0 + j + 1)
        return torch.from_numpy(np.stack([a, d]).astype(np.int32))

    def set_start(self, summaries, first_columns):
        if not summaries:
            assert self.c0 == 0
            self.start = (np.arange(self.m), np.zeros(self.m, np.int64))
            return
        a, d = (x.numpy().astype(np.int64) for x in summaries[0])
        for s, first in zip(summaries[1:], first_columns[1:]):
            pi, delta = (x.numpy().astype(np.int64) for x in s)
            a, d = compose(a, d, pi, delta, first)
        self.start = (a, d)

    def scan(self):
        a, d = self.start
        self.kept = []                                       # state after local column j, i.e. at right bound c0 + j + 1
        for j in range(self.nl):
            a, d = pbwt_step(a, d, self.sym[:, j], self.c0 + j + 1)
            self.kept.append((a, d))
        self.tb_max = np.full(self.dp_size, 0xffffffff, np.uint32)
        self.tb_max[: self.n_entries] = 0
        self.tb_lb = np.zeros(self.dp_size, np.uint32)
        self.tb_size = np.full(self.dp_size, 0xffffffff, np.uint32)
        self.reach = 0

    def dp_table(self):
        tabs = [torch.from_numpy(x.view(np.int32)) for x in (self.tb_max, self.tb_lb, self.tb_size)]
        last = self.dp_size if self.c1 == self.n else self.t_last
        return tabs, (self.t_first, last)

    # ---- DP(k), SURVEY.md A.3 / segmentation_lp_context.cc:393-481 ----
    def _dp_column(self, k):
        L, m = self.L, self.m
        d = self.kept[k - self.c0][1]
        vals, cnts = np.unique(d, return_counts=True)          # ascending (v, count) pairs
        best = (0, m, m)                                       # lb, max, size
        r, diff = 0, int(cnts[0])
        dp_rb = int(vals[0])
        if dp_rb == 0:
            if m - diff < best[1]:
                best = (0, m - diff, m - diff)
            r = 1
            if r < len(vals):
                dp_rb = int(vals[1])
                diff += int(cnts[1])
        r += 1
        ranges = []
        while r < len(vals):
            dp_lb, dp_rb = dp_rb, int(vals[r])
            hi = min(dp_rb, k + 2 - L)
            ok = True
            if dp_lb < L:
                if L < hi:
                    dp_lb = L
                else:
                    ok = False
            if ok and dp_lb < hi:
                ranges.append((dp_lb - L, hi - L, m - diff))
            diff += int(cnts[r])
            r += 1
        if ranges:
            idx = self.chk.rmq_queries(self.tb_max[: self.n_entries], [x[0] for x in ranges], [x[1] for x in ranges])
            for (beg, _end, size), i in zip(ranges, idx):
                self.reach = max(self.reach, (k + 1 - L) - beg)
                cand = max(int(self.tb_max[i]), size)
                if cand < best[1]:
                    best = (int(i) + L, cand, size)
        return best

    def dp_round(self):
        changed = 0
        for t in range(self.t_first, self.t_last):
            lb, mx, sz = self._dp_column(t + self.L - 1)
            if self.tb_max[t] != mx:
                changed = 1
            self.tb_max[t], self.tb_lb[t], self.tb_size[t] = mx, lb, sz
        return changed, self.reach

    def dp_sync(self, guess, reach):
        pass                                                    # rmq answers are computed from the table on demand; no guess

    def dp_single_chain(self):
        pass                                                    # the own entries are one chain already

    def dp_final(self):
        if self.c1 == self.n:
            t = self.n - self.L
            lb, mx, sz = self._dp_column(self.n - 1)
            self.tb_max[t], self.tb_lb[t], self.tb_size[t] = mx, lb, sz
        return 0

    def grow_cap(self):
        raise AssertionError("the CPU engine keeps complete lists")

    # ---- traceback, reduction ----
    def traceback(self):
        t = self.n - self.L
        path = [t]
        while self.tb_lb[t] != 0:
            t = int(self.tb_lb[t]) - self.L
            path.append(t)
        path.reverse()
        self.tb = (np.array([self.tb_lb[t] for t in path], np.uint64), np.array([t + self.L for t in path], np.uint64),
                   np.array([self.tb_max[t] for t in path], np.uint32), np.array([self.tb_size[t] for t in path], np.uint32))
        self.K = int(self.tb[2][-1])
        return len(path), self.K

    def get_traceback(self, count):
        return self.tb

    def position_range(self, c0, c1):
        rbs = self.tb[1]
        return int(np.searchsorted(rbs, c0, side="right")), int(np.searchsorted(rbs, c1, side="right"))

    def reduce_lists(self, count):
        from founder_sequences_b200.sharded import NotReducible
        if not self.K < self.m:
            raise NotReducible("K == m")
        width = self.K + 1
        self.lists = np.full((count, width, 2), 0xffffffff, np.uint32)
        rbs = self.tb[1]
        f = int(np.searchsorted(rbs, self.c0, side="right"))
        l = int(np.searchsorted(rbs, self.c1, side="right"))
        for p in range(f, l):
            d = self.kept[int(rbs[p]) - 1 - self.c0][1]
            vals = np.unique(d)[::-1]
            i = 0
            for v in vals:
                s = int((d > v).sum())
                if s <= self.K and i < width:
                    self.lists[p, i] = (v, s)
                    i += 1
        return torch.from_numpy(self.lists.reshape(-1).view(np.int32)), 2 * width, (f, l)

    def reduce(self):
        rbs, sizes = self.tb[1], self.tb[3]
        T = len(rbs)

        def count(t, lb):
            for v, s in self.lists[t]:
                if v <= lb:
                    return int(s)
            return self.m + 1

        self.red = []
        self.red_pos = []
        cur_lb, prev, prev_size = 0, 0, int(sizes[0])
        for t in range(1, T):                                   # A.6, segmentation_lp_context.cc:335-390
            s = count(t, cur_lb)
            if s <= self.K:
                prev_size = s
            else:
                self.red.append((cur_lb, int(rbs[prev]), prev_size))
                self.red_pos.append(prev)
                prev_size = int(sizes[t])
                cur_lb = int(rbs[prev])
            prev = t
        self.red.append((cur_lb, int(rbs[prev]), prev_size))
        self.red_pos.append(prev)
        red_rb = np.array([x[1] for x in self.red])
        self.own = (int(np.searchsorted(red_rb, self.c0, side="right")), int(np.searchsorted(red_rb, self.c1, side="right")))
        return len(self.red), self.own

    def bounds(self, count):
        return (np.array([x[0] for x in self.red], np.uint64), np.array([x[1] for x in self.red], np.uint64),
                np.array([x[2] for x in self.red], np.uint32))

    def states_of(self, rb):
        return self.kept[rb - 1 - self.c0]

    def states(self, own):
        a = np.zeros((own[1] - own[0], self.m), np.uint32)
        d = np.zeros((own[1] - own[0], self.m), np.uint32)
        for i, s in enumerate(range(*own)):
            a[i], d[i] = self.states_of(self.red[s][1])
        return a, d
