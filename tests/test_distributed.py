"""world_size-2 gloo test of the sweep's only collective: the final reduce of error counters.
Frames are sharded by global index; each rank counts its shard; all_reduce(sum) merges int64[nSNR][4]."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import polar_b200_loader
    sweep = polar_b200_loader.load().sweep
    total = 1000
    lo, hi = sweep.shard_range(total, rank, world)
    rng = np.random.default_rng(5)
    err = rng.integers(0, 3, (total, 2))  # deterministic "bit errors per frame" for 2 SNR points
    counts = torch.zeros((2, 4), dtype=torch.int64)
    for p in range(2):
        e = err[lo:hi, p]
        counts[p] = torch.tensor([hi - lo, int((e > 0).sum()), int(e.sum()), 0])
    merged = sweep.reduce_counts(counts)
    q.put((rank, lo, hi, merged.tolist()))
    dist.destroy_process_group()


def test_count_reduce_gloo():
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    res.sort()
    assert res[0][1] == 0 and res[0][2] == res[1][1] and res[1][2] == 1000  # contiguous, complete shards
    rng = np.random.default_rng(5)
    err = rng.integers(0, 3, (1000, 2))
    exp = [[1000, int((err[:, p] > 0).sum()), int(err[:, p].sum()), 0] for p in range(2)]
    assert res[0][3] == exp and res[1][3] == exp


def test_shard_range_properties():
    import polar_b200_loader
    sweep = polar_b200_loader.load().sweep
    for total in (0, 1, 7, 1000, 1 << 20):
        for world in (1, 2, 4, 8):
            spans = [sweep.shard_range(total, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == total
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            assert max(h - l for l, h in spans) - min(h - l for l, h in spans) <= 1
