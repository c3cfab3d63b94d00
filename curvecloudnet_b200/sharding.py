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


class GradReducer:
    """Training-step gradient all-reduce ("one gradient all-reduce per step, bucketed in reverse-layer
    order and overlapped with backward"; the reference itself trains on one GPU, src/main.py:56-60).

    Every parameter's `.grad` is a view into one flat fp32 buffer, cut into buckets in REVERSE parameter order (the
    order backward finishes them).  A post-accumulate hook counts a bucket's parameters; when the last one has its
    gradient, the bucket's all-reduce is issued asynchronously (NCCL runs it on its own stream over NVLink/NVSwitch
    while backward keeps going).  finish() waits for the buckets and leaves the MEAN over ranks in `.grad`.
    Parameters that got no gradient this step hold zeros (prepare() clears the buffer), so every rank issues the
    same collectives."""

    def __init__(self, params, dist=None, bucket_bytes: int = 16 << 20):
        self.params = [p for p in params if p.requires_grad]
        self.dist = dist if dist is not None and dist.is_initialized() and dist.get_world_size() > 1 else None
        self.world = self.dist.get_world_size() if self.dist is not None else 1
        total = sum(p.numel() for p in self.params)
        dev = self.params[0].device if self.params else "cpu"
        self.flat = torch.zeros(total, dtype=torch.float32, device=dev)
        self.buckets = []                 # [(start, end, n_params)]
        self._bucket_of = {}
        off, b_start, b_count = 0, 0, 0
        for p in reversed(self.params):
            n = p.numel()
            p.grad = self.flat[off:off + n].view_as(p)
            self._bucket_of[id(p)] = len(self.buckets)
            off += n
            b_count += 1
            if (off - b_start) * 4 >= bucket_bytes:
                self.buckets.append((b_start, off, b_count))
                b_start, b_count = off, 0
        if b_count:
            self.buckets.append((b_start, off, b_count))
        self._pending = [0] * len(self.buckets)
        self._work = []
        self._next = 0
        self._hooks = [p.register_post_accumulate_grad_hook(self._on_grad) for p in self.params]

    def prepare(self):
        """Call before every backward: zero the flat gradient, re-arm the buckets."""
        self.flat.zero_()
        lo, hi = self.flat.data_ptr(), self.flat.data_ptr() + self.flat.numel() * 4
        for p in self.params:             # an optimizer's zero_grad(set_to_none=True) would detach the views
            if p.grad is None or not (lo <= p.grad.data_ptr() < hi):
                raise RuntimeError("GradReducer: .grad was replaced; use optimizer.zero_grad(set_to_none=False) "
                                   "or just GradReducer.prepare()")
        self._pending = [b[2] for b in self.buckets]
        self._work = []
        self._next = 0                    # buckets are launched strictly in index order on every rank

    def _on_grad(self, p):
        b = self._bucket_of[id(p)]
        self._pending[b] -= 1
        self._launch_ready()

    def _launch_ready(self):
        """Issue the all-reduce of every complete bucket at the head of the queue.  A bucket only starts after all buckets
        before it have started, so that every rank issues the same collectives in the same ORDER even if a data-dependent
        branch makes the ranks finish their buckets in different orders (NCCL pairs collectives by issue order)."""
        while self._next < len(self.buckets) and self._pending[self._next] == 0:
            self._launch(self._next)
            self._next += 1

    def _launch(self, b):
        if self.dist is None:
            return
        s, e, _ = self.buckets[b]
        self._work.append(self.dist.all_reduce(self.flat[s:e], op=self.dist.ReduceOp.SUM, async_op=True))

    def finish(self):
        """Call after backward: flush, in order, the buckets that did not complete (parameters without a gradient this
        step hold zeros), wait, average."""
        while self._next < len(self.buckets):
            self._launch(self._next)
            self._next += 1
        for w in self._work:
            w.wait()
        if self.world > 1:
            self.flat.div_(self.world)
        self._work = []

    def remove(self):
        for h in self._hooks:
            h.remove()
