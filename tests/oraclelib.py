"""ctypes wrappers around the two checkers (TEST INFRASTRUCTURE):

    oracle/libfsq_oracle.so      plain-C restatement            (prefix fsqo_)
    oracle/_ref/libfsq_ref.so    the reference compiled in place (prefix fsqref_), when it was built

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs import this module.
"""
import ctypes as C
import os
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_SO = os.path.join(ROOT, "oracle", "libfsq_oracle.so")
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libfsq_ref.so")

_vp = C.c_void_p
_u32 = C.c_uint32
_u64 = C.c_uint64


def _ptr(a):
    return a.ctypes.data_as(_vp) if a is not None else None


class Checker:
    """Common interface over both libraries."""

    def __init__(self, path, prefix, is_ref):
        self.lib = C.CDLL(path)
        self.p = prefix
        self.is_ref = is_ref
        f = self._f
        f("create", _vp, [_vp, _u32, _u64])
        f("destroy", None, [_vp])
        f("alphabet", None, [_vp, _vp])
        f("scan_dump", C.c_longlong, [_vp, _u64, _vp, _vp, _vp, _vp, _vp, _u64])
        if is_ref:
            f("time_first_pass", C.c_double, [_vp, _u64, _u64, C.c_double])
            f("segment", C.c_int, [_vp, _u64, C.c_double, C.c_int])
            f("get_segment_texts", None, [_vp, _u64, _vp, _vp, _vp])
            f("get_pair_weights", None, [_vp, _u64, _vp])
        else:
            f("time_first_pass", C.c_double, [_vp, _u64, _u64])
            f("segment", C.c_int, [_vp, _u64])
            f("segment_sets", _u32, [_vp, _u64, _vp, _vp, _vp])
            f("pair_weights", None, [_vp, _u64, _vp])
        f("first_pass_seconds", C.c_double, [_vp])
        f("K", _u32, [_vp])
        for nm in ("dp_size", "traceback_size", "reduced_size"):
            f(nm, _u64, [_vp])
        for nm in ("get_dp", "get_traceback", "get_reduced"):
            f(nm, None, [_vp, _vp, _vp, _vp, _vp])
        for nm in ("get_traceback_states", "get_reduced_states"):
            f(nm, None, [_vp, _vp, _vp])
        f("short_path", _u64, [_vp, _vp, _vp])
        f("rmq_queries", None, [_vp, _u64, _vp, _vp, _u64, _vp])
        if not is_ref:
            f("match_founders", _u64, [_vp, _u32, _vp, _u32, _u64, _u64, _vp, _vp, _vp, _vp, _u64])

    def _f(self, name, restype, argtypes):
        fn = getattr(self.lib, self.p + name)
        fn.restype = restype
        fn.argtypes = argtypes
        setattr(self, "_" + name, fn)

    # -- handle-less helpers ------------------------------------------------------------------
    def rmq_queries(self, vals, beg, end):
        vals = np.ascontiguousarray(vals, np.uint32)
        beg = np.ascontiguousarray(beg, np.uint64)
        end = np.ascontiguousarray(end, np.uint64)
        out = np.zeros(len(beg), np.uint64)
        self._rmq_queries(_ptr(vals), len(vals), _ptr(beg), _ptr(end), len(beg), _ptr(out))
        return out

    def match_founders(self, seqs, founders, min_len=0):
        """match_founder_sequences.cc:127-214: (seq, lb, rb, founder bit sets [count, ceil(K / 32)])."""
        seqs = np.ascontiguousarray(seqs, np.uint8)
        founders = np.ascontiguousarray(founders, np.uint8)
        m, n = seqs.shape
        K = founders.shape[0]
        words = (K + 31) // 32
        cap = 1024
        while True:
            sq, lb, rb = np.zeros(cap, np.uint32), np.zeros(cap, np.uint64), np.zeros(cap, np.uint64)
            bits = np.zeros((cap, words), np.uint32)
            cnt = self._match_founders(_ptr(seqs), m, _ptr(founders), K, n, min_len, _ptr(sq), _ptr(lb), _ptr(rb), _ptr(bits), cap)
            if cnt != 0xffffffffffffffff:
                return sq[:cnt], lb[:cnt], rb[:cnt], bits[:cnt]
            cap *= 8

    def open(self, rows):
        return Session(self, rows)


class Session:
    def __init__(self, chk, rows):
        rows = np.ascontiguousarray(rows, np.uint8)
        assert rows.ndim == 2
        self.c = chk
        self.m, self.n = rows.shape
        self.h = chk._create(_ptr(rows), self.m, self.n)

    def close(self):
        if self.h:
            self.c._destroy(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def alphabet(self):
        comp = np.zeros(256, np.uint8)
        self.c._alphabet(self.h, _ptr(comp))
        return comp

    def scan_dump(self, ncols, hist=True):
        m = self.m
        a = np.zeros((ncols, m), np.uint32)
        d = np.zeros((ncols, m), np.uint32)
        if hist:
            cap = ncols * (m + 1)
            off = np.zeros(ncols + 1, np.uint64)
            hv = np.zeros(cap, np.uint32)
            hc = np.zeros(cap, np.uint32)
            r = self.c._scan_dump(self.h, ncols, _ptr(a), _ptr(d), _ptr(off), _ptr(hv), _ptr(hc), cap)
            assert r >= 0
            return a, d, off, hv[:r], hc[:r]
        self.c._scan_dump(self.h, ncols, _ptr(a), _ptr(d), None, None, None, 0)
        return a, d

    def time_first_pass(self, L, ncols):
        if self.c.is_ref:
            return self.c._time_first_pass(self.h, L, ncols, 4.0)
        return self.c._time_first_pass(self.h, L, ncols)

    def segment(self, L, with_texts=False):
        """Returns the status code (0 ok, 1 short path, 2 K == m)."""
        if self.c.is_ref:
            return self.c._segment(self.h, L, 4.0, 1 if with_texts else 0)
        return self.c._segment(self.h, L)

    def _args(self, getter, count):
        lb = np.zeros(count, np.uint64)
        rb = np.zeros(count, np.uint64)
        mx = np.zeros(count, np.uint32)
        sz = np.zeros(count, np.uint32)
        getter(self.h, _ptr(lb), _ptr(rb), _ptr(mx), _ptr(sz))
        return lb, rb, mx, sz

    def K(self):
        return self.c._K(self.h)

    def first_pass_seconds(self):
        return self.c._first_pass_seconds(self.h)

    def dp(self):
        return self._args(self.c._get_dp, self.c._dp_size(self.h))

    def traceback(self):
        return self._args(self.c._get_traceback, self.c._traceback_size(self.h))

    def reduced(self):
        return self._args(self.c._get_reduced, self.c._reduced_size(self.h))

    def _states(self, getter, count):
        a = np.zeros((count, self.m), np.uint32)
        d = np.zeros((count, self.m), np.uint32)
        getter(self.h, _ptr(a), _ptr(d))
        return a, d

    def traceback_states(self):
        return self._states(self.c._get_traceback_states, self.c._traceback_size(self.h))

    def reduced_states(self):
        return self._states(self.c._get_reduced_states, self.c._reduced_size(self.h))

    def short_path(self):
        row = np.zeros(self.m, np.uint32)
        ln = np.zeros(self.m, np.uint32)
        cnt = self.c._short_path(self.h, _ptr(row), _ptr(ln))
        return row[:cnt], ln[:cnt]

    # oracle only
    def segment_sets(self, s):
        off = np.zeros(self.m + 1, np.uint32)
        ids = np.zeros(self.m, np.uint32)
        cls = np.zeros(self.m, np.uint32)
        runs = self.c._segment_sets(self.h, s, _ptr(off), _ptr(ids), _ptr(cls))
        return off[: runs + 1], ids, cls

    def pair_weights(self, s, runs_l, runs_r):
        W = np.zeros((runs_l, runs_r), np.int32)
        self.c._pair_weights(self.h, s, _ptr(W))
        return W

    # ref only
    def segment_texts(self, s):
        K = self.K()
        off = np.zeros(K + 1, np.uint64)
        ids = np.zeros(self.m, np.uint64)
        cp = np.zeros(K, np.uint64)
        self.c._get_segment_texts(self.h, s, _ptr(off), _ptr(ids), _ptr(cp))
        return off, ids[: off[K]], cp

    def ref_pair_weights(self, s):
        K = self.K()
        W = np.zeros((K, K), np.int32)
        self.c._get_pair_weights(self.h, s, _ptr(W))
        return W


def oracle():
    if not os.path.exists(ORACLE_SO):
        raise RuntimeError("oracle/libfsq_oracle.so is missing: run `make -C oracle` or __graft_entry__.build()")
    return Checker(ORACLE_SO, "fsqo_", False)


def have_ref():
    return os.path.exists(REF_SO)


def ref():
    return Checker(REF_SO, "fsqref_", True)
