"""N>1 host logic on CPU: world-size-2 gloo group (scans shard, no data-path collective)."""
import os
import socket

import torch
import torch.multiprocessing as mp

from curvecloudnet_b200.sharding import gather_predictions, owner_of, reduce_measurement, shard_scan_ids


def test_round_robin_partition_is_exact():
    for world in (1, 2, 4, 8):
        for n in (0, 1, 7, 8, 33):
            seen = sorted(s for r in range(world) for s in shard_scan_ids(n, r, world))
            assert seen == list(range(n))
            assert all(owner_of(s, world) == r for r in range(world) for s in shard_scan_ids(n, r, world))
            sizes = [len(shard_scan_ids(n, r, world)) for r in range(world)]
            assert max(sizes) - min(sizes) <= 1


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        n_scans = 5
        mine = shard_scan_ids(n_scans, rank, world)
        # a stand-in for the per-scan result: deterministic function of the scan id
        local = {s: torch.arange(10 + s, dtype=torch.int64) * (s + 1) for s in mine}
        ms, cnt = reduce_measurement([10.0 + rank, 3.0 - rank], [sum(10 + s for s in mine), len(mine)], dist)
        allp = gather_predictions(local, n_scans, dist)
        ok = (ms == [10.0 + world - 1, 3.0] and cnt == [float(sum(10 + s for s in range(n_scans))), float(n_scans)]
              and all(torch.equal(allp[s], torch.arange(10 + s, dtype=torch.int64) * (s + 1)) for s in range(n_scans)))
        out.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


def test_world2_gloo_measurement_and_gather():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(60)
    assert res == [(0, True), (1, True)]
