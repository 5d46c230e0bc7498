"""Ad-hoc check (one GPU): config 5's shape on a slice -- 20,000 rows, chunk length 53, K in the thousands -- column-sharded
over ranks as threads with the top lists in windows, against the unsharded run of the same library."""
import os, sys, threading, zlib
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
from __graft_entry__ import load_package, load_sharded
fsq, sh = load_package(), load_sharded()
m, n, L, parts = 20000, int(sys.argv[1]) if len(sys.argv) > 1 else 48000, 100, int(sys.argv[2]) if len(sys.argv) > 2 else 2
wb = int(sys.argv[3]) if len(sys.argv) > 3 else 6
with fsq.Context(0) as c0:
    with c0.msa_synth(m, n, 5, 64, True) as full, c0.segment(full, L) as seg:
        ref = (seg.max_segment_size,) + seg.reduced_traceback()
        a, d = seg.reduced_pbwt_samples_view()
        ref_crc = np.array([zlib.crc32(d[i].tobytes(), zlib.crc32(a[i].tobytes())) for i in range(a.shape[0])], np.uint32)
        print("unsharded: K = %d, %d segments, cap %d, %d rounds, windows %d" % (ref[0], len(ref[1]), c0.stat("cap"), c0.stat("dp_rounds"), c0.stat("windows")), flush=True)
lib = fsq.load_library()
group = sh.LocalGroup(lib, parts)
out, err = [None] * parts, [None] * parts
def work(rank):
    try:
        c0_, c1_ = sh.shard_columns(n, parts, rank)
        ctx = fsq.Context(0)
        ctx.set_option("window_blocks", wb)
        msa = ctx.msa_synth(m, c1_ - c0_, 5, 64, True, col0=c0_)
        comm = group.comm(ctx, rank)
        seg = sh.segment_sharded(ctx, msa, comm, c0_, n, L)
        lb, rb, sz = seg.reduced_traceback()
        a, d = seg.reduced_pbwt_samples_view()
        crc = np.array([zlib.crc32(d[i].tobytes(), zlib.crc32(a[i].tobytes())) for i in range(a.shape[0])], np.uint32)
        out[rank] = (seg.max_segment_size, lb, rb, sz, seg.own, crc, ctx.stat("windows"), ctx.stat("cap"), ctx.stat("dp_rounds"), ctx.stat("cap_retries"))
        seg.free(); comm.free(); msa.free(); ctx.close()
    except BaseException as e:
        err[rank] = e
ts = [threading.Thread(target=work, args=(r,)) for r in range(parts)]
[t.start() for t in ts]; [t.join() for t in ts]
for e in err:
    if e is not None: raise e
for r, o in enumerate(out):
    ok = o[0] == ref[0] and all(np.array_equal(x, y) for x, y in zip(o[1:4], ref[1:4])) and np.array_equal(o[5], ref_crc[o[4][0]:o[4][1]])
    print("rank %d: %s  own %s windows %d cap %d rounds %d retries %d" % (r, "ok" if ok else "MISMATCH", o[4], o[6], o[7], o[8], o[9]), flush=True)
    assert ok
print("c5 shape check ok")
