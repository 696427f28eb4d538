"""Host-side multi-GPU logic (foundationstereo_b200/shard.py) on CPU: partitioning properties and
a world_size-2 run over the gloo backend (one process per rank, rendezvous on 127.0.0.1), with a
CPU stand-in for the per-batch step -- the data path itself has no collective and no CPU
implementation, so what is checked here is exactly what crosses ranks: who owns which pairs and
the max-time / sum-pairs reduction bench.py reports."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from foundationstereo_b200 import shard


@pytest.mark.parametrize("n,world", [(64, 1), (64, 2), (64, 4), (64, 8), (7, 4), (3, 8), (0, 2)])
def test_shard_range_partitions_exactly(n, world):
    ranges = [shard.shard_range(n, r, world) for r in range(world)]
    assert ranges[0][0] == 0 and ranges[-1][1] == n
    for (a0, a1), (b0, b1) in zip(ranges, ranges[1:]):
        assert a1 == b0 and a0 <= a1
    sizes = [b - a for a, b in ranges]
    assert sum(sizes) == n and max(sizes) - min(sizes) <= 1


def test_micro_batches_and_slices():
    assert list(shard.micro_batches(8, 16, 3)) == [(8, 11), (11, 14), (14, 16)]
    assert list(shard.micro_batches(5, 5, 4)) == []
    with pytest.raises(ValueError):
        list(shard.micro_batches(0, 4, 0))
    with pytest.raises(ValueError):
        shard.shard_range(4, 2, 2)
    x = {"a": torch.arange(10).reshape(10, 1), "l": [torch.arange(10), torch.arange(10) * 2]}
    y = shard.slice_batch(x, 2, 5)
    assert y["a"].flatten().tolist() == [2, 3, 4] and y["l"][1].tolist() == [4, 6, 8]


def test_reduce_throughput_single_rank():
    pairs, ms, rate = shard.reduce_throughput(8, 200.0)
    assert (pairs, ms) == (8, 200.0) and abs(rate - 40.0) < 1e-9


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, n_pairs, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        inputs = {"disp": torch.arange(n_pairs, dtype=torch.float32).reshape(n_pairs, 1, 1, 1),
                  "net": [torch.arange(n_pairs, dtype=torch.float32).reshape(n_pairs, 1)]}
        step = lambda d: d["disp"] * 2 + d["net"][0].reshape(-1, 1, 1, 1)   # stand-in for StereoHotPath.run
        outs, rng = shard.run_sharded(step, inputs, n_pairs, rank, world, micro_batch=3)
        got = torch.cat(outs).flatten().tolist() if outs else []
        local_ms = 100.0 * (rank + 1)                # rank 1 is the slow one
        pairs, ms, rate = shard.reduce_throughput(rng[1] - rng[0], local_ms)
        ranges = shard.gather_ranges(rng)
        q.put((rank, rng, got, pairs, ms, rate, ranges))
    finally:
        dist.destroy_process_group()


def test_world_size_2_over_gloo():
    world, n_pairs = 2, 11
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_pairs, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = sorted(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    covered = []
    for rank, rng, got, pairs, ms, rate, ranges in results:
        assert got == [3.0 * i for i in range(rng[0], rng[1])]   # each rank computed exactly its own pairs
        covered += list(range(rng[0], rng[1]))
        assert pairs == n_pairs and ms == 200.0                      # SUM of pairs, MAX of times
        assert abs(rate - n_pairs / 0.2) < 1e-6
        assert ranges == [(0, 6), (6, 11)]
    assert covered == list(range(n_pairs))
