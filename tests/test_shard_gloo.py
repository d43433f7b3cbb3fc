"""World-size-2 gloo test (CPU) of the multi-GPU host logic: per-rank shard plans, distinct Philox streams,
max-over-ranks timing and the whole-job rate.  The data path itself has no collective to test."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from augcl_b200 import shard
        from oracle import augcl_oracle as O
        p = shard.plan(100_000, 128, base_seed=7)
        assert (p.rank, p.world, p.global_batch) == (rank, world, 256)
        shard.barrier()
        elapsed = 0.010 * (rank + 1)                      # rank 1 is the slow one
        tmax = shard.max_over_ranks(elapsed)
        per_rank = shard.gather_over_ranks(elapsed)
        assert per_rank == [0.010, 0.020]
        idx = O.rng_randint(128, 0, p.capacity, p.draw_seed, 0)
        gathered = [None] * world
        dist.all_gather_object(gathered, (p.ring_seed, p.draw_seed, idx.tolist()))
        out[rank] = (tmax, shard.aggregate_rate(1000 * p.batch, world, tmax), gathered)
    finally:
        dist.destroy_process_group()


def test_two_rank_shards_and_max_timing():
    world, port = 2, _free_port()
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, port, out), nprocs=world, join=True)
    assert set(out.keys()) == {0, 1}
    for r in range(world):
        tmax, rate, gathered = out[r]
        assert abs(tmax - 0.020) < 1e-12                 # MAX over ranks, seen by every rank
        assert abs(rate - 2 * 128_000 / 0.020) < 1e-6
        (rs0, ds0, i0), (rs1, ds1, i1) = gathered
        assert rs0 != rs1 and ds0 != ds1                 # independent rings and draw streams
        assert i0 != i1 and max(i0 + i1) < 100_000 and min(i0 + i1) >= 0


def test_single_process_defaults():
    from augcl_b200 import shard
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):
        os.environ.pop(k, None)
    p = shard.plan(10, 4)
    assert (p.rank, p.world, p.global_batch) == (0, 1, 4)
    assert shard.max_over_ranks(1.5) == 1.5
    assert shard.gather_over_ranks(1.5) == [1.5]


def test_core_slices_are_disjoint_and_cover():
    from augcl_b200 import shard
    cores = list(range(32))
    for world in (1, 2, 4, 8):
        sl = [shard.core_slice(r, world, cores) for r in range(world)]
        assert all(len(s) == 32 // world for s in sl)
        flat = [c for s in sl for c in s]
        assert sorted(flat) == cores                      # disjoint, nothing left idle
    # more ranks than cores: one core each, wrapping
    assert [shard.core_slice(r, 4, [5, 9]) for r in range(4)] == [[5], [9], [5], [9]]
    assert shard.core_slice(0, 2, []) == []
    got = shard.pin_to_cores(0, 1)                        # whole allowed set: a no-op pin
    assert got is None or sorted(got) == sorted(os.sched_getaffinity(0))
