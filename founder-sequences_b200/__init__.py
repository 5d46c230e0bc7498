"""founder-sequences_b200 — Python host side over the C ABI of libfsq_b200.so (include/fsq.h).

The reference's host is C++ (its mirror lives in host/); this module is the thin ctypes binding that the
tests and bench.py drive.  It contains no arithmetic: every result comes from the CUDA library, and loading
fails loudly when the library has not been built (there is no CPU fallback).

The directory name has a hyphen, so import it by path::

    import importlib.util, sys
    spec = importlib.util.spec_from_file_location("founder_sequences_b200", ".../founder-sequences_b200/__init__.py",
                                                  submodule_search_locations=[".../founder-sequences_b200"])
    fsq = importlib.util.module_from_spec(spec); sys.modules[spec.name] = fsq; spec.loader.exec_module(fsq)

(`__graft_entry__.load_package()` does exactly this.)
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "lib", "libfsq_b200.so")

OK, ERR_ARG, ERR_CUDA, ERR_SHORT, ERR_NOT_REDUCIBLE, ERR_NOMEM, ERR_UNSUPPORTED, ERR_INTERNAL = range(8)

# Every symbol include/fsq.h declares; tests check that the built library exports all of them.
EXPORTS = [
    "fsq_create", "fsq_destroy", "fsq_last_error", "fsq_version", "fsq_set_option", "fsq_get_stat",
    "fsq_msa_from_rows", "fsq_msa_from_device_rows", "fsq_msa_synth", "fsq_synth_rows_device", "fsq_msa_free",
    "fsq_msa_rows", "fsq_msa_columns", "fsq_identity_columns", "fsq_pbwt_states", "fsq_segment", "fsq_segmentation_free",
    "fsq_segmentation_count", "fsq_segmentation_max_segment_size", "fsq_segmentation_bounds",
    "fsq_segmentation_states", "fsq_segmentation_states_range", "fsq_segmentation_states_view", "fsq_segmentation_traceback_count", "fsq_segmentation_traceback",
    "fsq_segmentation_dp_table", "fsq_distinct_rows", "fsq_segment_sets", "fsq_pair_weights", "fsq_pair_counts", "fsq_pair_counts_noalloc", "fsq_match_founders",
    "fsq_msa_set_first_column", "fsq_shard_columns", "fsq_comm_unique_id", "fsq_comm_create", "fsq_comm_adopt_nccl",
    "fsq_local_group_create", "fsq_local_group_free", "fsq_comm_create_local", "fsq_comm_free", "fsq_comm_rank", "fsq_comm_size",
    "fsq_segment_sharded", "fsq_segmentation_own_range",
]


class FsqError(RuntimeError):
    def __init__(self, code, message):
        super().__init__("fsq error %d: %s" % (code, message))
        self.code = code


_lib = None


def load_library():
    """dlopen libfsq_b200.so and declare the prototypes.  Raises if the library is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            "%s is missing: build it with `python founder-sequences_b200/build.py` "
            "(or __graft_entry__.build()); there is no CPU fallback" % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    vp, u32, u64, i64, ci = C.c_void_p, C.c_uint32, C.c_uint64, C.c_int64, C.c_int
    pp = C.POINTER(vp)

    def proto(name, res, args):
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args

    proto("fsq_create", ci, [ci, pp])
    proto("fsq_destroy", None, [vp])
    proto("fsq_last_error", C.c_char_p, [vp])
    proto("fsq_version", C.c_char_p, [])
    proto("fsq_set_option", ci, [vp, C.c_char_p, i64])
    proto("fsq_get_stat", ci, [vp, C.c_char_p, C.POINTER(C.c_double)])
    proto("fsq_msa_from_rows", ci, [vp, vp, u32, u64, vp, pp])
    proto("fsq_msa_from_device_rows", ci, [vp, vp, u64, u32, u64, vp, pp])
    proto("fsq_msa_synth", ci, [vp, u32, u64, u64, u64, u32, ci, pp])
    proto("fsq_synth_rows_device", ci, [vp, vp, u64, u32, u64, u64, u64, u32, ci])
    proto("fsq_msa_free", None, [vp])
    proto("fsq_msa_rows", u32, [vp])
    proto("fsq_msa_columns", u64, [vp])
    proto("fsq_identity_columns", ci, [vp, vp, vp, C.POINTER(u64), pp])
    proto("fsq_pbwt_states", ci, [vp, vp, vp, u32, vp, vp])
    proto("fsq_segment", ci, [vp, vp, u64, pp])
    proto("fsq_segmentation_free", None, [vp])
    proto("fsq_segmentation_count", u32, [vp])
    proto("fsq_segmentation_max_segment_size", u32, [vp])
    proto("fsq_segmentation_bounds", ci, [vp, vp, vp, vp])
    proto("fsq_segmentation_states", ci, [vp, vp, vp])
    proto("fsq_segmentation_states_range", ci, [vp, u32, u32, vp, vp])
    proto("fsq_segmentation_states_view", ci, [vp, pp, pp])
    proto("fsq_segmentation_traceback_count", u32, [vp])
    proto("fsq_segmentation_traceback", ci, [vp, vp, vp, vp, vp])
    proto("fsq_segmentation_dp_table", ci, [vp, vp, vp, vp, vp])
    proto("fsq_distinct_rows", ci, [vp, vp, vp, vp, C.POINTER(u32)])
    proto("fsq_segment_sets", ci, [vp, vp, u32, vp, vp, vp, C.POINTER(u32)])
    proto("fsq_pair_weights", ci, [vp, vp, u32, vp, C.POINTER(u32), C.POINTER(u32)])
    proto("fsq_pair_counts", ci, [vp, vp, u32, vp, vp, vp, C.POINTER(u32), C.POINTER(u32), C.POINTER(u32)])
    proto("fsq_pair_counts_noalloc", ci, [vp, vp, u32, u32, u32, vp, vp, vp, vp, vp, C.c_char_p, C.c_size_t])
    proto("fsq_match_founders", ci, [vp, vp, vp, u64, vp, vp, vp, vp, u64, C.POINTER(u64)])
    pu32, pu64 = C.POINTER(u32), C.POINTER(u64)
    proto("fsq_msa_set_first_column", ci, [vp, u64])
    proto("fsq_shard_columns", ci, [u64, ci, ci, pu64, pu64])
    proto("fsq_comm_unique_id", ci, [vp])
    proto("fsq_comm_create", ci, [vp, vp, ci, ci, pp])
    proto("fsq_comm_adopt_nccl", ci, [vp, vp, ci, ci, pp])
    proto("fsq_local_group_create", ci, [ci, pp])
    proto("fsq_local_group_free", None, [vp])
    proto("fsq_comm_create_local", ci, [vp, vp, ci, pp])
    proto("fsq_comm_free", None, [vp])
    proto("fsq_comm_rank", ci, [vp])
    proto("fsq_comm_size", ci, [vp])
    proto("fsq_segment_sharded", ci, [vp, vp, vp, u64, u64, pp])
    proto("fsq_segmentation_own_range", ci, [vp, pu32, pu32])
    _lib = lib
    return lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


class Context:
    """One per device; mirrors the role of the reference's generate_context as the owner of the run."""

    def __init__(self, device=0):
        self.lib = load_library()
        self.device = device
        h = C.c_void_p()
        rc = self.lib.fsq_create(device, C.byref(h))
        if rc:
            raise FsqError(rc, self.lib.fsq_last_error(None).decode())
        self.h = h

    def close(self):
        if self.h:
            self.lib.fsq_destroy(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def check(self, rc):
        if rc:
            raise FsqError(rc, self.lib.fsq_last_error(self.h).decode())

    def set_option(self, key, value):
        self.check(self.lib.fsq_set_option(self.h, key.encode(), int(value)))

    def stat(self, key):
        v = C.c_double()
        self.check(self.lib.fsq_get_stat(self.h, key.encode(), C.byref(v)))
        return v.value

    def reset_stats(self):
        self.stat("reset")

    # ---- alignments ----
    def msa_from_rows(self, rows, comp=None):
        """rows: (m, n) uint8 array (host).  The row pointers handed to the library mirror the reference's
        sequence_vector of spans."""
        rows = np.ascontiguousarray(rows, np.uint8)
        m, n = rows.shape
        ptrs = (C.c_void_p * m)(*[rows.ctypes.data + i * rows.strides[0] for i in range(m)])
        h = C.c_void_p()
        cp = _p(np.ascontiguousarray(comp, np.uint8)) if comp is not None else None
        self.check(self.lib.fsq_msa_from_rows(self.h, ptrs, m, n, cp, C.byref(h)))
        return Msa(self, h, m, n)

    def msa_from_device_rows(self, dptr, row_stride, m, n, comp):
        h = C.c_void_p()
        comp = np.ascontiguousarray(comp, np.uint8)
        self.check(self.lib.fsq_msa_from_device_rows(self.h, C.c_void_p(dptr), row_stride, m, n, _p(comp), C.byref(h)))
        return Msa(self, h, m, n)

    def msa_synth(self, m, n, seed, founders, gaps=False, col0=0):
        h = C.c_void_p()
        self.check(self.lib.fsq_msa_synth(self.h, m, n, col0, seed, founders, 1 if gaps else 0, C.byref(h)))
        return Msa(self, h, m, n)

    def synth_rows_device(self, dptr, row_stride, m, n, seed, founders, gaps=False, col0=0):
        self.check(self.lib.fsq_synth_rows_device(self.h, C.c_void_p(dptr), row_stride, m, n, col0, seed, founders, 1 if gaps else 0))

    def identity_columns(self, msa, want_msa=False):
        """remove_identity_columns' test: the 0/1 mask of the columns in which all rows agree, and (want_msa) the alignment
        without them."""
        mask = np.zeros(msa.n, np.uint8)
        cnt = C.c_uint64()
        h = C.c_void_p()
        self.check(self.lib.fsq_identity_columns(self.h, msa.h, _p(mask), C.byref(cnt), C.byref(h) if want_msa else None))
        if want_msa:
            return mask, Msa(self, h, msa.m, msa.n - cnt.value)
        return mask

    def match_founders(self, sequences, founders, min_segment_length=0):
        """match_founder_sequences: (seq, lb, rb, founder bit sets [count, ceil(K / 32)]) for the rows of `sequences`
        against the rows of `founders` (two alignments packed with the same comp table)."""
        cnt = C.c_uint64()
        self.check(self.lib.fsq_match_founders(self.h, sequences.h, founders.h, min_segment_length, None, None, None, None, 0, C.byref(cnt)))
        n, words = cnt.value, (founders.m + 31) // 32
        sq, lb, rb = np.zeros(n, np.uint32), np.zeros(n, np.uint64), np.zeros(n, np.uint64)
        bits = np.zeros((n, words), np.uint32)
        self.check(self.lib.fsq_match_founders(self.h, sequences.h, founders.h, min_segment_length, _p(sq), _p(lb), _p(rb), _p(bits), max(n, 1), C.byref(cnt)))
        return sq, lb, rb, bits

    # ---- the path ----
    def pbwt_states(self, msa, columns):
        """(a_c, d_c) for ascending columns c in [1, n]: the arrays pbwt_context::input_permutation() /
        input_divergence() hold once sequence_idx() == c."""
        cols = np.ascontiguousarray(columns, np.uint64)
        a = np.zeros((len(cols), msa.m), np.uint32)
        d = np.zeros((len(cols), msa.m), np.uint32)
        self.check(self.lib.fsq_pbwt_states(self.h, msa.h, _p(cols), len(cols), _p(a), _p(d)))
        return a, d

    def segment(self, msa, segment_length_bound, comm=None, n_total=None):
        """The long path.  With a communicator (sharded.Comm): msa holds this rank's columns of an alignment of n_total
        columns (fsq_segment_sharded); every rank of the communicator makes the same call."""
        h = C.c_void_p()
        if comm is None:
            rc = self.lib.fsq_segment(self.h, msa.h, segment_length_bound, C.byref(h))
        else:
            rc = self.lib.fsq_segment_sharded(self.h, msa.h, comm.h, n_total, segment_length_bound, C.byref(h))
        if rc == ERR_NOT_REDUCIBLE and h:
            seg = Segmentation(self, h, msa.m)
            err = FsqError(rc, self.lib.fsq_last_error(self.h).decode())
            err.segmentation = seg
            raise err
        self.check(rc)
        return Segmentation(self, h, msa.m)

    def distinct_rows(self, msa):
        row = np.zeros(msa.m, np.uint32)
        ln = np.zeros(msa.m, np.uint32)
        cnt = C.c_uint32()
        self.check(self.lib.fsq_distinct_rows(self.h, msa.h, _p(row), _p(ln), C.byref(cnt)))
        return row[: cnt.value], ln[: cnt.value]


class Msa:
    def __init__(self, ctx, h, m, n):
        self.ctx, self.h, self.m, self.n = ctx, h, m, n

    def free(self):
        if self.h:
            self.ctx.lib.fsq_msa_free(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.free()


class Segmentation:
    """Mirror of founder_sequences::segmentation_container (segmentation_container.hh:15-20)."""

    def __init__(self, ctx, h, m):
        self.ctx, self.h, self.m = ctx, h, m
        lib = ctx.lib
        self.max_segment_size = lib.fsq_segmentation_max_segment_size(h)
        self.count = lib.fsq_segmentation_count(h)
        f, l = C.c_uint32(), C.c_uint32()
        lib.fsq_segmentation_own_range(h, C.byref(f), C.byref(l))
        self.own = (f.value, l.value)       # the reduced segments whose states this handle holds (all, on one GPU)

    def free(self):
        if self.h:
            self.ctx.lib.fsq_segmentation_free(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.free()

    def reduced_traceback(self):
        lb = np.zeros(self.count, np.uint64)
        rb = np.zeros(self.count, np.uint64)
        sz = np.zeros(self.count, np.uint32)
        self.ctx.check(self.ctx.lib.fsq_segmentation_bounds(self.h, _p(lb), _p(rb), _p(sz)))
        return lb, rb, sz

    def reduced_pbwt_samples(self):
        cnt = self.own[1] - self.own[0]
        a = np.zeros((cnt, self.m), np.uint32)
        d = np.zeros((cnt, self.m), np.uint32)
        self.ctx.check(self.ctx.lib.fsq_segmentation_states(self.h, _p(a), _p(d)))
        return a, d

    def reduced_pbwt_samples_range(self, first, count):
        """States of segments [first, first + count), copied from the device on their own."""
        a = np.zeros((count, self.m), np.uint32)
        d = np.zeros((count, self.m), np.uint32)
        self.ctx.check(self.ctx.lib.fsq_segmentation_states_range(self.h, first, count, _p(a), _p(d)))
        return a, d

    def reduced_pbwt_samples_view(self):
        """Zero-copy views of the handle's pinned host buffers (valid until free())."""
        pa, pd = C.c_void_p(), C.c_void_p()
        self.ctx.check(self.ctx.lib.fsq_segmentation_states_view(self.h, C.byref(pa), C.byref(pd)))
        shape = (self.own[1] - self.own[0], self.m)
        if 0 == shape[0]:
            return np.zeros(shape, np.uint32), np.zeros(shape, np.uint32)
        a = np.ctypeslib.as_array(C.cast(pa, C.POINTER(C.c_uint32)), shape)
        d = np.ctypeslib.as_array(C.cast(pd, C.POINTER(C.c_uint32)), shape)
        return a, d

    def traceback(self):
        t = self.ctx.lib.fsq_segmentation_traceback_count(self.h)
        lb = np.zeros(t, np.uint64)
        rb = np.zeros(t, np.uint64)
        mx = np.zeros(t, np.uint32)
        sz = np.zeros(t, np.uint32)
        self.ctx.check(self.ctx.lib.fsq_segmentation_traceback(self.h, _p(lb), _p(rb), _p(mx), _p(sz)))
        return lb, rb, mx, sz

    def dp_table(self, size):
        lb = np.zeros(size, np.uint64)
        rb = np.zeros(size, np.uint64)
        mx = np.zeros(size, np.uint32)
        sz = np.zeros(size, np.uint32)
        self.ctx.check(self.ctx.lib.fsq_segmentation_dp_table(self.h, _p(lb), _p(rb), _p(mx), _p(sz)))
        return lb, rb, mx, sz

    def segment_sets(self, s):
        off = np.zeros(self.m + 1, np.uint32)
        ids = np.zeros(self.m, np.uint32)
        cls = np.zeros(self.m, np.uint32)
        runs = C.c_uint32()
        self.ctx.check(self.ctx.lib.fsq_segment_sets(self.ctx.h, self.h, s, _p(off), _p(ids), _p(cls), C.byref(runs)))
        return off[: runs.value + 1], ids, cls

    def pair_counts(self, s):
        """greedy_matcher.cc:293-343: the distinct (lhs class, rhs class) pairs of segments (s, s + 1), ascending, with counts."""
        l, r, c = np.zeros(self.m, np.uint32), np.zeros(self.m, np.uint32), np.zeros(self.m, np.uint32)
        n = C.c_uint32()
        self.ctx.check(self.ctx.lib.fsq_pair_counts(self.ctx.h, self.h, s, _p(l), _p(r), _p(c), C.byref(n), None, None))
        return l[: n.value], r[: n.value], c[: n.value]

    def pair_weights(self, s):
        k = max(1, self.max_segment_size)
        W = np.zeros(k * k, np.int32)   # a reduced segment has at most K runs
        rl, rr = C.c_uint32(), C.c_uint32()
        self.ctx.check(self.ctx.lib.fsq_pair_weights(self.ctx.h, self.h, s, _p(W), C.byref(rl), C.byref(rr)))
        return W[: rl.value * rr.value].reshape(rl.value, rr.value).copy()
