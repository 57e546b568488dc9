"""world_size-2 gloo test of the multi-GPU plumbing (runs on CPU): round-robin frame
ownership is a partition, every rank derives the same descriptors for a frame regardless
of who owns it, and the aggregate is sum(units) / max(time)."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import harness as H

sys.path.insert(0, H.ROOT)


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from b200av1 import shard
    import bench
    mine = shard.frames_of_rank(7, rank, world)
    # the per-frame work is identical whoever owns the frame (seeded by frame id, not rank)
    blocks, samples = bench.make_descriptors(2, 1, [(64, 48), (32, 24), (32, 24)], 64 * 48 + 2 * 32 * 24, 5)
    digest = int(np.bitwise_xor.reduce(blocks["src_offset"].astype(np.int64) * 31 + blocks["tx_type"]))
    units, secs = float(len(mine)), 0.5 + rank  # rank 1 is slower
    agg, usum, tmax = shard.aggregate_throughput(units, secs, dist)
    dist.barrier()
    q.put((rank, mine, digest, agg, usum, tmax))
    dist.destroy_process_group()


def test_frame_sharding_gloo_world2():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    frames = res[0][1] + res[1][1]
    assert sorted(frames) == list(range(7)) and not set(res[0][1]) & set(res[1][1])
    assert res[0][2] == res[1][2]
    for r in res:
        assert r[4] == 7.0 and r[5] == 1.5 and abs(r[3] - 7.0 / 1.5) < 1e-12
