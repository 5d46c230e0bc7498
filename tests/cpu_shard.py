"""TEST INFRASTRUCTURE: plain-Python models of the PBWT step and of the boundary stitch (A.10) the
column sharding rests on, for tiny inputs (tests/test_sharded_gloo.py); the product never imports this."""
import numpy as np


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
