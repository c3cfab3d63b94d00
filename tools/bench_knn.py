"""Dev tool: exact 3-NN, brute force (ccn_knn_exact) against the hashed grid (ccn_knn_grid), on laser-scan levels.
usage: python tools/bench_knn.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from curvecloudnet_b200 import ops
from curvecloudnet_b200.synthetic import driving_scan_batch


def timed(fn):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(10):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


for kind, seeds in (("nuscenes", [0]), ("kitti", [2]), ("kitti", [2, 3])):
    d = driving_scan_batch(seeds, kind=kind).to("cuda")
    pos, batch = d.pos, d.batch
    for stride in (2, 6, 17):
        sub = torch.arange(0, pos.shape[0], stride, device="cuda")
        p2, b2 = pos[sub].contiguous(), batch[sub].contiguous()
        s2 = ops.sample_ptr(b2)
        a = ops.knn_exact(pos, batch, p2, s2, 3, use_grid=False)
        g = ops.knn_exact(pos, batch, p2, s2, 3, batch2=b2, use_grid=True)
        assert torch.equal(a[0], g[0]) and torch.equal(a[1], g[1])
        tb = timed(lambda: ops.knn_exact(pos, batch, p2, s2, 3, use_grid=False))
        tg = timed(lambda: ops.knn_exact(pos, batch, p2, s2, 3, batch2=b2, use_grid=True))
        print("%-8s B=%d queries %6d sources %6d: brute %8.1f us   grid %8.1f us" % (kind, len(seeds), pos.shape[0], p2.shape[0], tb, tg))
