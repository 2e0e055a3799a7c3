"""CPU: the N>1 host path (clip sharding + job-level reductions) on a world_size-2 gloo group."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from timelapse_video_interpolation_b200 import sharding


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n_clips, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        mine = sharding.shard_clips(n_clips, rank, world)
        # every rank learns every rank's shard (test-only exchange; the product path has no collective)
        gathered = [None] * world
        dist.all_gather_object(gathered, mine)
        t_max = sharding.max_over_ranks(10.0 + rank)            # rank 1 is the slow one
        total = sharding.sum_over_ranks(len(mine) * 14)          # clip-frames at T=14
        q.put((rank, mine, gathered, t_max, total))
    finally:
        dist.destroy_process_group()


def test_two_rank_gloo_sharding_partitions_clips_and_reduces_times():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port, world, n_clips = _free_port(), 2, 7
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_clips, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = sorted(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, mine, gathered, t_max, total in results:
        assert mine == list(range(rank, n_clips, world))
        flat = sorted(i for shard in gathered for i in shard)
        assert flat == list(range(n_clips))                      # a partition: nothing lost, nothing twice
        assert t_max == 11.0                                     # max over ranks, identical on every rank
        assert total == n_clips * 14


def test_single_process_helpers_are_identity():
    assert sharding.shard_clips(5, 0, 1) == [0, 1, 2, 3, 4]
    assert sharding.max_over_ranks(3.5) == 3.5 and sharding.sum_over_ranks(2) == 2.0
    for bad in [(3, 2, 2), (3, -1, 2), (3, 0, 0)]:
        try:
            sharding.shard_clips(*bad)
            raise AssertionError("expected ValueError")
        except ValueError:
            pass
    assert torch.tensor(1).item() == 1
