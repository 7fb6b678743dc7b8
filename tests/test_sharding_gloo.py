"""Host-side logic of the multi-GPU path on CPU: frame sharding and max-over-ranks timing (gloo, world_size 2)."""
import os

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from rdp_b200.pipeline import frame_shard, max_over_ranks


def test_frame_shard_partitions_exactly():
    for total in (0, 1, 7, 64, 65, 65536):
        for world in (1, 2, 3, 8):
            seen = []
            for r in range(world):
                sh = frame_shard(total, world, r)
                seen.extend(sh)
                assert abs(len(sh) - total / world) < 1
            assert seen == list(range(total))
    with pytest.raises(ValueError):
        frame_shard(10, 2, 2)


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        shard = frame_shard(101, world, rank)
        # every rank "processes" its own frames; results are gathered on the host, no data-path collective
        local = [(i, float(i) * 0.5) for i in shard]
        gathered = [None] * world
        dist.all_gather_object(gathered, local)
        slowest = max_over_ranks(10.0 + rank, device="cpu")
        q.put((rank, sum(len(g) for g in gathered), sorted(i for g in gathered for i, _ in g) == list(range(101)), slowest))
    finally:
        dist.destroy_process_group()


def test_two_ranks_shard_and_reduce_with_gloo():
    world, port = 2, 29571
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, n, complete, slowest in res:
        assert n == 101 and complete
        assert slowest == 11.0          # the slowest rank's time, identical on every rank
