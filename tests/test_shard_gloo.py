"""N > 1 host logic on CPU: world_size-2 gloo processes shard the pair batch
without overlap or gaps and agree on the max-over-ranks interval."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from raft_stir_b200.shard import max_over_ranks, shard_pairs


def test_shard_pairs_partition():
    for total in (0, 1, 7, 64, 65):
        for world in (1, 2, 3, 8):
            seen = []
            for r in range(world):
                first, count = shard_pairs(total, world, r)
                seen.extend(range(first, first + count))
            assert seen == list(range(total))
            counts = [shard_pairs(total, world, r)[1] for r in range(world)]
            assert max(counts) - min(counts) <= 1
    with pytest.raises(ValueError):
        shard_pairs(4, 2, 2)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, total, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        first, count = shard_pairs(total, world, rank)
        mine = torch.zeros(total, dtype=torch.int64)
        mine[first:first + count] = 1
        dist.all_reduce(mine)                          # every pair owned exactly once
        slow = max_over_ranks(10.0 + 5.0 * rank)       # rank 1 is the slow one
        # each rank "processes" its pairs; per-rank results never need to meet
        g = torch.Generator().manual_seed(1000 + rank)
        local = torch.randn(count, 4, generator=g).sum().item()
        out.put((rank, mine.tolist(), slow, count, local))
    finally:
        dist.destroy_process_group()


def test_two_rank_gloo_sharding():
    world, total = 2, 7
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, total, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert [r[3] for r in res] == [4, 3]
    for _, owned, slow, _, _ in res:
        assert owned == [1] * total
        assert slow == 15.0
