"""Scan-sharded data parallelism: per-scan inference has no cross-GPU dependency, so the
batch/stream of scans is partitioned over the ranks of one box and nothing is exchanged on the data path.
The only collectives are the ones around a measurement (max of the per-rank device time, sum of the points)
and, for callers that want every rank to hold all results, an all_gather of per-scan outputs.
Works with any torch.distributed backend (nccl on the GPU box, gloo in the CPU tests)."""
from __future__ import annotations

from typing import List, Sequence, Tuple

import torch


def shard_scan_ids(num_scans: int, rank: int, world: int) -> List[int]:
    """Rank r takes scans r, r+world, r+2*world, ... (round robin keeps ranks within one scan of each other)."""
    if not (0 <= rank < world):
        raise ValueError("rank %d outside world %d" % (rank, world))
    return list(range(rank, num_scans, world))


def owner_of(scan_id: int, world: int) -> int:
    return scan_id % world


def reduce_measurement(ms: Sequence[float], counts: Sequence[float], dist=None, device="cpu") -> Tuple[List[float], List[float]]:
    """-> (max over ranks of every time, sum over ranks of every count)."""
    t = torch.tensor(list(ms), dtype=torch.float64, device=device)
    c = torch.tensor(list(counts), dtype=torch.float64, device=device)
    if dist is not None and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(c, op=dist.ReduceOp.SUM)
    return [float(v) for v in t.tolist()], [float(v) for v in c.tolist()]


def gather_predictions(local: dict, num_scans: int, dist=None) -> list:
    """local: {scan_id: int64 tensor of per-point labels} of this rank -> list over ALL scans (every rank)."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return [local[i] for i in range(num_scans)]
    world = dist.get_world_size()
    parts = [None] * world
    dist.all_gather_object(parts, {k: v.cpu() for k, v in local.items()})
    merged = {}
    for p in parts:
        merged.update(p)
    missing = [i for i in range(num_scans) if i not in merged]
    if missing:
        raise RuntimeError("scans %s were processed by no rank" % missing)
    return [merged[i] for i in range(num_scans)]
