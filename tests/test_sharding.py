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


def _grad_reducer_worker(rank, world, port, q):
    import os
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from curvecloudnet_b200.sharding import GradReducer
    torch.manual_seed(0)
    net = torch.nn.Sequential(torch.nn.Linear(8, 16), torch.nn.ReLU(), torch.nn.Linear(16, 4), torch.nn.Linear(4, 2))
    unused = torch.nn.Linear(3, 3)                       # gets no gradient: its bucket must still be reduced
    params = list(net.parameters()) + list(unused.parameters())
    red = GradReducer(params, dist, bucket_bytes=64)     # several small buckets
    assert len(red.buckets) > 2
    outs = []
    for step in range(2):
        torch.manual_seed(100 * step + rank)
        x = torch.randn(5, 8)
        red.prepare()
        net(x).pow(2).sum().backward()
        red.finish()
        outs.append([p.grad.clone() for p in params])
    # reference: the same two steps for every rank, averaged
    want = []
    for step in range(2):
        acc = None
        for r in range(world):
            torch.manual_seed(100 * step + r)
            x = torch.randn(5, 8)
            g = torch.autograd.grad(net(x).pow(2).sum(), list(net.parameters()))
            acc = [a + b for a, b in zip(acc, g)] if acc is not None else list(g)
        want.append([a / world for a in acc])
    ok = all(torch.allclose(o, w, atol=1e-6) for step in range(2) for o, w in zip(outs[step], want[step]))
    ok = ok and all(float(o.abs().max()) == 0.0 for o in outs[1][len(want[1]):])
    q.put((rank, bool(ok)))
    dist.destroy_process_group()


def test_grad_reducer_world2_gloo():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29641
    procs = [ctx.Process(target=_grad_reducer_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
    assert res == [(0, True), (1, True)]
