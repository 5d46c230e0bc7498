"""Device timeline of one column-sharded config-2 step (1,000 x 1M per rank, L = 20) on every rank: each phase occurrence
with its start and end relative to the call's start (option "timeline").

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tests/gpu_timeline_sharded.py"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
NAMES = ["pass_a", "stitch", "prep", "pass_b", "dp", "exchange", "traceback", "reduce", "pass_c", "?"]


def main():
    import torch
    import torch.distributed as dist
    from __graft_entry__ import load_package, load_sharded
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    fsq, sh = load_package(), load_sharded()
    m, n, L = 1000, 1000000, 20
    ctx = fsq.Context(local)
    comm = sh.nccl_comm(ctx, dist)
    c0, c1 = sh.shard_columns(n * world, world, rank)
    msa = ctx.msa_synth(m, c1 - c0, 2, 16, True, col0=c0)
    for _ in range(4):
        sh.segment_sharded(ctx, msa, comm, c0, n * world, L).free()
    ctx.set_option("timeline", 1)
    ctx.reset_stats()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    sh.segment_sharded(ctx, msa, comm, c0, n * world, L).free()
    t1 = time.perf_counter()
    lines = ["rank %d wall %.3f ms, host: call %.3f = enqueue %.3f + wait %.3f + after the last read %.3f" % (rank, 1e3 * (t1 - t0), ctx.stat("hms_call"), ctx.stat("hms_enqueue"), ctx.stat("hms_wait"), ctx.stat("hms_after"))]
    last = 0.0
    for i in range(int(ctx.stat("tl_count"))):
        at, end, name = ctx.stat("tl_at_%d" % i), ctx.stat("tl_end_%d" % i), NAMES[int(ctx.stat("tl_id_%d" % i))]
        lines.append("  %-10s at %7.3f end %7.3f dur %6.3f  gap before %6.3f" % (name, at, end, end - at, at - last))
        last = end
    out = [None] * world
    dist.all_gather_object(out, "\n".join(lines))
    if rank == 0:
        for r in (0, world // 2, world - 1) if world > 3 else range(world):
            print(out[r])
    comm.free(); msa.free(); ctx.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
