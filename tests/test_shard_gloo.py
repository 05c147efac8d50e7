"""CPU, world_size 2 over gloo: the multi-GPU host logic (env ranges, seeds, end-of-rollout gather)."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from vectorizedmultiagentsimulator_b200.shard import gather_stats, shard_range, shard_seed


def test_shard_ranges_cover_the_batch():
    for n in (1, 7, 32768, 1000003):
        for w in (1, 2, 4, 8):
            r = [shard_range(n, k, w) for k in range(w)]
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(r, r[1:]))
            assert max(e - b for b, e in r) - min(e - b for b, e in r) <= 1
    assert shard_seed(5, 3) == 8


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    b, e = shard_range(11, rank, world)
    stats = {"reward": torch.arange(b, e, dtype=torch.float32) * 2, "done": (torch.arange(b, e) % 3 == 0)}
    out = gather_stats(stats, dst=0)
    if rank == 0:
        q.put({k: v.tolist() for k, v in out.items()})
    else:
        assert out is None
    dist.barrier()
    dist.destroy_process_group()


def test_gather_stats_gloo_world2():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = q.get(timeout=120)
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert got["reward"] == [2.0 * i for i in range(11)]
    assert got["done"] == [i % 3 == 0 for i in range(11)]
