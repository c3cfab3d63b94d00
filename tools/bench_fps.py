"""Dev tool: device time of ccn_fps at the level sizes of the driving configs.  usage: python tools/bench_fps.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from curvecloudnet_b200 import ops

for n, B in [(1933, 1), (580, 1), (1933, 6), (5000, 1), (8000, 2), (30000, 1)]:
    gen = torch.Generator().manual_seed(n)
    pos = (torch.rand(n * B, 3, generator=gen) * torch.tensor([7.0, 7.0, 0.5])).cuda()
    batch = torch.repeat_interleave(torch.arange(B), n).cuda()
    sptr = ops.sample_ptr(batch)
    ksel = ops.fps_count(sptr.lengths, 0.3)
    starts = [0] * B

    def run():
        return ops.fps(pos, sptr, ksel, starts)

    for _ in range(3):
        run()
    torch.cuda.synchronize()
    ts = []
    for _ in range(10):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); run(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts.sort()
    print("fps n=%6d x B=%d -> K=%5d: %8.1f us (%.3f us / round)" % (n, B, ksel[0], ts[len(ts) // 2], ts[len(ts) // 2] / max(ksel[0], 1)))
