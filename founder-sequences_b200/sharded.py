"""Column-sharded run of the long path: one rank per GPU, each holding a contiguous column range of ONE alignment.

The reference runs the path in a single process.  The sharded run itself -- every stage, every exchange
(ncclAllGather on the library's own stream) -- lives in libfsq_b200.so behind fsq_segment_sharded() (include/fsq.h);
this module is the thin host side a launcher needs:

  * `nccl_comm(ctx, dist)`   the library's NCCL communicator for the ranks of a torch.distributed process group (the
                             128-byte unique id travels through the group, whatever its backend)
  * `LocalGroup`             ranks as threads of one process (several shards through one GPU: the tests)
  * `shard_columns`          the library's split rule (fsq_shard_columns)
  * `segment_sharded`        the collective call; returns the library's Segmentation handle
  * `assemble_states`        puts the ranks' own states side by side (what rank 0 of a host driver would do)
"""
import ctypes as C

import numpy as np


class Comm:
    """An fsq_comm handle."""

    def __init__(self, lib, h, rank, size, keep=None):
        self.lib, self.h, self.rank, self.size, self._keep = lib, h, rank, size, keep

    def free(self):
        if self.h:
            self.lib.fsq_comm_free(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.free()


def exchange_unique_id(make_id, rank, broadcast):
    """Rank 0 makes the 128-byte id, `broadcast(array, src=0)` hands it to the others (in place)."""
    ident = np.zeros(128, np.uint8)
    if rank == 0:
        make_id(ident)
    broadcast(ident)
    return ident


def nccl_comm(ctx, dist, group=None, device=None):
    """The library's own NCCL communicator over the ranks of a torch.distributed group (one rank per GPU)."""
    import torch
    rank, size = dist.get_rank(group), dist.get_world_size(group)
    lib = ctx.lib
    backend = dist.get_backend(group)
    dev = torch.device("cuda", ctx.device) if backend == "nccl" else torch.device("cpu")

    def make_id(ident):
        ctx.check(lib.fsq_comm_unique_id(ident.ctypes.data_as(C.c_void_p)))

    def broadcast(ident):
        t = torch.from_numpy(ident).to(dev)
        dist.broadcast(t, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
        ident[:] = t.cpu().numpy()

    ident = exchange_unique_id(make_id, rank, broadcast)
    h = C.c_void_p()
    ctx.check(lib.fsq_comm_create(ctx.h, ident.ctypes.data_as(C.c_void_p), rank, size, C.byref(h)))
    return Comm(lib, h, rank, size)


class LocalGroup:
    """Ranks as threads of one process (fsq_local_group): `comm(ctx, rank)` per thread."""

    def __init__(self, lib, size):
        self.lib, self.size = lib, size
        h = C.c_void_p()
        if lib.fsq_local_group_create(size, C.byref(h)):
            raise RuntimeError("fsq_local_group_create failed")
        self.h = h

    def comm(self, ctx, rank):
        h = C.c_void_p()
        ctx.check(self.lib.fsq_comm_create_local(ctx.h, self.h, rank, C.byref(h)))
        return Comm(self.lib, h, rank, self.size, keep=self)

    def free(self):
        if self.h:
            self.lib.fsq_local_group_free(self.h)
            self.h = None


def shard_columns(n, size, rank=None):
    """The library's split rule (fsq_shard_columns): equal shares of ceil(n / size) rounded up to 32 columns.  Without a
    rank: the list of all ranks' ranges; raises when a rank would be left without columns."""
    per = (n + size - 1) // size
    per = (per + 31) // 32 * 32 if size > 1 else n
    out = []
    for r in range(size):
        c0 = min(n, r * per)
        c1 = min(n, c0 + per)
        if c0 >= c1:
            raise ValueError("%d ranks are too many for %d columns: rank %d would hold none" % (size, n, r))
        out.append((c0, c1))
    return out if rank is None else out[rank]


def segment_sharded(ctx, msa, comm, first_column, n_total, L):
    """Collective: every rank calls it with its shard (columns [first_column, first_column + msa.n) of n_total).
    Returns the rank's Segmentation: the whole reduced segmentation, and the states of the segments in `.own`."""
    ctx.check(ctx.lib.fsq_msa_set_first_column(msa.h, first_column))
    return ctx.segment(msa, L, comm=comm, n_total=n_total)


def assemble_states(parts, count, m):
    """parts: (own, a, d) per rank -> the states of all `count` reduced segments; checks that the ranks' ranges tile."""
    a = np.zeros((count, m), np.uint32)
    d = np.zeros((count, m), np.uint32)
    covered = np.zeros(count, bool)
    for (f, l), pa, pd in parts:
        if covered[f:l].any():
            raise ValueError("reduced segments [%d, %d) are held by two ranks" % (f, l))
        covered[f:l] = True
        a[f:l], d[f:l] = pa, pd
    if not covered.all():
        raise ValueError("some reduced segments are held by no rank")
    return a, d
