"""Host-side logic of the multi-GPU path on the CPU: shard arithmetic and the world-size-2 gloo reduction."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from roma_b200.sharding import aggregate, shard_range


def test_shard_range_partitions_exactly():
    for n in (0, 1, 7, 8, 64, 65):
        for world in (1, 2, 3, 4, 8):
            seen = [i for r in range(world) for i in shard_range(n, world, r)]
            assert seen == list(range(n))
            sizes = [len(shard_range(n, world, r)) for r in range(world)]
            assert max(sizes) - min(s for s in sizes if s or True) <= (n + world - 1) // world
    assert list(shard_range(64, 8, 3)) == list(range(24, 32))
    with pytest.raises(ValueError):
        shard_range(8, 2, 2)


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        pairs = len(shard_range(13, world, rank))           # 7 + 6
        total, ms = aggregate(pairs, 10.0 + 5.0 * rank)     # rank 1 is the slow one
        q.put((rank, total, ms))
    finally:
        dist.destroy_process_group()


def test_aggregate_gloo_world2():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    out = sorted(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert out == [(0, 13, 15.0), (1, 13, 15.0)]            # sum of pairs, max of times on every rank


def test_aggregate_single_process_is_identity():
    assert aggregate(8, 12.5) == (8, 12.5)
