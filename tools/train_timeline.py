"""Device timeline of one training step (ShapeNet-seg B = 16 x 2048 by default) from CUPTI activity records
(torch.profiler): span, busy / idle time of the stream, the distribution of the gaps between consecutive kernels and a
per-kernel table - tells whether the step waits for the host (gaps) or for the kernels (busy).
Usage: python tools/train_timeline.py [model] [batch]"""
import os
import sys
from collections import defaultdict

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    from torch.profiler import ProfilerActivity, profile
    name = sys.argv[1] if len(sys.argv) > 1 else "shapenet-seg"
    batch = int(sys.argv[2]) if len(sys.argv) > 2 else 16
    from curvecloudnet_b200 import autograd as A
    from curvecloudnet_b200.configs import build_model
    from curvecloudnet_b200.model import ModelBase
    from curvecloudnet_b200.synthetic import object_scan_batch, randomize_model_
    model = randomize_model_(build_model(ModelBase, name, precision="bf16x3"), 1234).cuda().train()
    opt = torch.optim.Adam(model.parameters(), lr=1e-3, fused=True)
    gen = torch.Generator().manual_seed(5)
    batches, targets = [], []
    for s in range(2):
        d = object_scan_batch(list(range(100 + s * batch, 100 + (s + 1) * batch)))
        batches.append(d.to("cuda"))
        targets.append(torch.randint(0, model.n_out, (d.pos.shape[0],), generator=gen).cuda())

    marks = []

    def step(i, mark=False):
        opt.zero_grad(set_to_none=True)
        out = model(batches[i % 2])
        loss = A.nll_loss(out, targets[i % 2])
        if mark:
            marks.append(torch.zeros(3, device="cuda"))      # a memset / fill between forward and backward
        loss.backward()
        opt.step()

    for i in range(6):
        step(i)
    torch.cuda.synchronize()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        for i in range(3):
            flush.zero_()
            torch.cuda.synchronize()
            step(i)
            torch.cuda.synchronize()
    ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
    ks = sorted(((e.time_range.start, e.time_range.end, e.name, getattr(e, "device_resource_id", 0)) for e in ev), key=lambda t: t[0])
    fills = [i for i, k in enumerate(ks) if "FillFunctor<unsigned char>" in k[2] and k[1] - k[0] > 20]
    if len(fills) < 3:
        print("could not find the flush kernels (%d)" % len(fills))
        return
    last = ks[fills[-1] + 1:]
    t0, t1 = last[0][0], max(k[1] for k in last)
    print("%s B=%d training step: %d device activities, span %.1f us" % (name, batch, len(last), t1 - t0))
    by_stream = defaultdict(list)
    for k in last:
        by_stream[k[3]].append(k)
    for s, lst in sorted(by_stream.items(), key=lambda kv: -len(kv[1])):
        busy = sum(k[1] - k[0] for k in lst)
        print("stream %s: %d activities, busy %.1f us, first %.1f last %.1f" % (s, len(lst), busy, lst[0][0] - t0, lst[-1][1] - t0))
    main_stream = max(by_stream, key=lambda s: len(by_stream[s]))
    lst = by_stream[main_stream]
    gaps = [(lst[i + 1][0] - max(k[1] for k in lst[:i + 1][-4:]), lst[i][0] - t0, lst[i][2], lst[i + 1][2]) for i in range(len(lst) - 1)]
    gaps = [(max(g, 0.0), at, a, b) for g, at, a, b in gaps]
    print("main stream idle %.1f us in %d gaps" % (sum(g[0] for g in gaps), len(gaps)))
    for lo, hi in ((0, 1), (1, 2), (2, 4), (4, 8), (8, 20), (20, 50), (50, 1e9)):
        sel = [g[0] for g in gaps if lo <= g[0] < hi]
        print("  gap %4g-%-4g us: %4d gaps, %8.1f us" % (lo, hi if hi < 1e9 else float("inf"), len(sel), sum(sel)))
    print("largest gaps (us): at, after -> before")
    for g, at, a, b in sorted(gaps, key=lambda g: -g[0])[:25]:
        print("  %7.1f  at %8.1f   %-50.50s -> %-50.50s" % (g, at, a, b))
    # windows of 1 ms: busy fraction along the step
    nb = int((t1 - t0) // 1000) + 1
    busy = [0.0] * nb
    for k in lst:
        a, b = k[0] - t0, k[1] - t0
        while a < b:
            w = int(a // 1000)
            e = min(b, (w + 1) * 1000.0)
            busy[w] += e - a
            a = e
    print("busy fraction per 1 ms window: " + " ".join("%.2f" % (x / 1000.0) for x in busy))
    tab = defaultdict(lambda: [0, 0.0])
    for k in last:
        tab[k[2]][0] += 1
        tab[k[2]][1] += k[1] - k[0]
    print("per kernel (all streams):")
    for nme, (cnt, tot) in sorted(tab.items(), key=lambda kv: -kv[1][1])[:45]:
        print("  %-90.90s %4d %9.1f us  avg %7.2f" % (nme, cnt, tot, tot / cnt))


if __name__ == "__main__":
    main()
