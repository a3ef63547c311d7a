"""Host logic of the multi-GPU path on CPU: world_size-2 gloo."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from stereo_vo_b200 import sharding


def test_shard_partition_properties():
    for n in (0, 1, 7, 64, 65, 200):
        for world in (1, 2, 3, 4, 8):
            allids = []
            for r in range(world):
                ids = sharding.shard_sequences(n, world, r)
                allids += ids
                for s in ids:
                    assert sharding.owner_of(s, n, world) == r
            assert allids == list(range(n))
            sizes = [len(sharding.shard_sequences(n, world, r)) for r in range(world)]
            assert max(sizes) - min(sizes) <= 1
    assert sharding.weak_shard(64, 8, 3) == list(range(192, 256))
    with pytest.raises(ValueError):
        sharding.shard_sequences(4, 2, 2)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n_total, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    mine = sharding.shard_sequences(n_total, world, rank)
    # every rank "processes" its own sequences: no data-path collective is needed; the only
    # exchange is the timing / frame-count reduction
    ms, frames = sharding.reduce_time_and_frames(10.0 + 5.0 * rank, len(mine) * 3, dist)
    gathered = [None] * world
    dist.all_gather_object(gathered, mine)
    if rank == 0:
        out.put((ms, frames, gathered))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_gloo_sharding():
    world, n_total = 2, 65
    port = _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_total, q)) for r in range(world)]
    for p in procs:
        p.start()
    ms, frames, gathered = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert ms == 15.0                      # max over ranks
    assert frames == n_total * 3           # sum over ranks
    flat = [s for part in gathered for s in part]
    assert flat == list(range(n_total))    # a partition: every sequence exactly once
