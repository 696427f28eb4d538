"""Multi-GPU sharding of the hot path: stereo pairs are independent and every
op on the path is per pair, so a batch of pairs is split into contiguous chunks, one chunk per
rank (one process per GPU, weights replicated), with NO collective on the data path.  The
only cross-rank traffic is the reduction of the timing / throughput scalars at the end
(`reduce_throughput`), which works on the NCCL backend (GPU tensors) and on gloo (CPU
tensors; that is what the world_size-2 CPU tests use).

The reference has no multi-GPU inference path (inference is one process, one GPU,
scripts/run_demo.py:112-167); BASELINE.json configs[3] (64 KITTI-shape pairs over 1/2/4/8
GPUs) is what this module serves.
"""
from __future__ import annotations

from typing import Callable, Dict, Iterator, List, Sequence, Tuple

import torch
import torch.distributed as dist

Tensor = torch.Tensor


def shard_range(n_items: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous [begin, end) of `n_items` owned by `rank`: the first n % world ranks get one
    extra item, every item belongs to exactly one rank, empty ranges when world > n_items."""
    if not (0 <= rank < world) or n_items < 0:
        raise ValueError(f"bad shard request n_items={n_items} rank={rank} world={world}")
    base, extra = divmod(n_items, world)
    begin = rank * base + min(rank, extra)
    return begin, begin + base + (1 if rank < extra else 0)


def micro_batches(begin: int, end: int, size: int) -> Iterator[Tuple[int, int]]:
    """[begin, end) cut into runs of at most `size` pairs (the per-launch batch on one GPU;
    bounded so the per-pair cost / geometry volumes of the largest config fit in HBM)."""
    if size <= 0:
        raise ValueError("micro-batch size must be positive")
    i = begin
    while i < end:
        j = min(i + size, end)
        yield i, j
        i = j


def slice_batch(x: Dict[str, object], lo: int, hi: int) -> Dict[str, object]:
    """Rows [lo, hi) of every tensor / list of tensors of a hot-path input dict."""
    out: Dict[str, object] = {}
    for k, v in x.items():
        out[k] = [t[lo:hi] for t in v] if isinstance(v, (list, tuple)) else v[lo:hi]
    return out


def run_sharded(step: Callable[[Dict[str, object]], Tensor], inputs: Dict[str, object], n_pairs: int, rank: int,
                world: int, micro_batch: int) -> Tuple[List[Tensor], Tuple[int, int]]:
    """Runs `step` (one pass of the hot path over a batch dict -> disparity) over this rank's
    chunk of the `n_pairs` pairs in `inputs`, micro-batch by micro-batch.  Returns the list of
    per-micro-batch results and the rank's [begin, end)."""
    begin, end = shard_range(n_pairs, rank, world)
    outs = []
    for lo, hi in micro_batches(begin, end, micro_batch):
        outs.append(step(slice_batch(inputs, lo, hi)))
    return outs, (begin, end)


def reduce_throughput(local_pairs: int, local_ms: float, device: torch.device | str = "cpu") -> Tuple[int, float, float]:
    """(total pairs, max-over-ranks milliseconds, pairs/s) of a job in which every rank
    processed `local_pairs` pairs in `local_ms` device milliseconds.  Without an initialised
    process group this is the single-rank identity."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return local_pairs, local_ms, (local_pairs / (local_ms * 1e-3) if local_ms > 0 else 0.0)
    t = torch.tensor([float(local_ms)], dtype=torch.float64, device=device)
    n = torch.tensor([float(local_pairs)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dist.all_reduce(n, op=dist.ReduceOp.SUM)
    ms, pairs = float(t.item()), int(round(n.item()))
    return pairs, ms, (pairs / (ms * 1e-3) if ms > 0 else 0.0)


def gather_ranges(rng: Tuple[int, int]) -> List[Tuple[int, int]]:
    """Every rank's [begin, end), in rank order (used by the driver to check coverage)."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return [tuple(rng)]
    out: List[object] = [None] * dist.get_world_size()
    dist.all_gather_object(out, tuple(rng))
    return [tuple(r) for r in out]
