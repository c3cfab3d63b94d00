"""Device timeline of one headline forward (B=1 nuScenes-shape scan, native executor) from CUPTI activity records
(torch.profiler): for every kernel its stream, start and duration - so the idle time BETWEEN kernels of the main stream
(launch boundaries, host waits at data-dependent sizes) is measured, not inferred.  Prints the span, the busy / idle split
per stream, the largest gaps with the kernels on either side, and a per-kernel table.  The profiler adds ~1-2 us per launch
on the host; the device-side durations and the gaps between back-to-back kernels are what this is read for.
Usage: python tools/timeline.py [precision] [--json out.json]"""
import json
import os
import sys
from collections import defaultdict

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    from torch.profiler import ProfilerActivity, profile
    from curvecloudnet_b200.configs import build_model
    from curvecloudnet_b200.model import ModelBase
    from curvecloudnet_b200.synthetic import driving_scan_batch, randomize_model_
    precision = sys.argv[1] if len(sys.argv) > 1 and not sys.argv[1].startswith("--") else "bf16x3"
    out_json = sys.argv[sys.argv.index("--json") + 1] if "--json" in sys.argv else None
    model = randomize_model_(build_model(ModelBase, "nuscenes", precision=precision), 1234).cuda().eval()
    scans = [driving_scan_batch([s]).to("cuda") for s in range(3)]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    with torch.no_grad():
        for i in range(6):
            torch.manual_seed(i)
            model(scans[i % 3])
        torch.cuda.synchronize()
        with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
            for i in range(3):
                flush.zero_()
                torch.cuda.synchronize()
                torch.manual_seed(i)
                model(scans[i % 3])
                torch.cuda.synchronize()
    ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
    ks = sorted(((e.time_range.start, e.time_range.end, e.name, getattr(e, "device_resource_id", 0)) for e in ev), key=lambda t: t[0])
    # split into the three forwards: the flush kernel (a 256 MiB fill) separates them
    fills = [i for i, k in enumerate(ks) if "FillFunctor<unsigned char>" in k[2] and k[1] - k[0] > 20]
    if len(fills) < 3:
        print("could not find the flush kernels (%d)" % len(fills))
        return
    last = ks[fills[-1] + 1:]
    last = [k for k in last if "Memset" not in k[2] or True]
    t0, t1 = last[0][0], max(k[1] for k in last)
    print("forward: %d device activities, span %.1f us" % (len(last), t1 - t0))
    by_stream = defaultdict(list)
    for k in last:
        by_stream[k[3]].append(k)
    main_stream = max(by_stream, key=lambda s: len(by_stream[s]))
    report = {"span_us": t1 - t0, "streams": {}}
    for s, lst in sorted(by_stream.items(), key=lambda kv: -len(kv[1])):
        busy = sum(k[1] - k[0] for k in lst)
        print("stream %s: %d activities, busy %.1f us, first %.1f last %.1f" % (s, len(lst), busy, lst[0][0] - t0, lst[-1][1] - t0))
        report["streams"][str(s)] = {"n": len(lst), "busy_us": busy}
    lst = by_stream[main_stream]
    gaps = []
    for a, b in zip(lst[:-1], lst[1:]):
        gaps.append((b[0] - a[1], a, b))
    tot_gap = sum(g[0] for g in gaps if g[0] > 0)
    busy = sum(k[1] - k[0] for k in lst)
    print("main stream: busy %.1f us + idle %.1f us between %d consecutive activities (mean gap %.2f us)" %
          (busy, tot_gap, len(gaps), tot_gap / max(len(gaps), 1)))
    hist = defaultdict(lambda: [0, 0.0])
    for g, _, _ in gaps:
        b = "<1" if g < 1 else "1-2" if g < 2 else "2-4" if g < 4 else "4-8" if g < 8 else "8-20" if g < 20 else ">=20"
        hist[b][0] += 1
        hist[b][1] += max(g, 0)
    for b in ("<1", "1-2", "2-4", "4-8", "8-20", ">=20"):
        print("  gap %-5s us: %4d gaps, %8.1f us" % (b, hist[b][0], hist[b][1]))
    print("largest gaps (us): after -> before")
    for g, a, b in sorted(gaps, key=lambda t: -t[0])[:25]:
        print("  %7.1f  at %8.1f   %-50s -> %-50s" % (g, a[1] - t0, a[2][:50], b[2][:50]))
    per = defaultdict(lambda: [0, 0.0])
    for k in last:
        per[k[2][:70]][0] += 1
        per[k[2][:70]][1] += k[1] - k[0]
    print("per kernel (all streams):")
    for name, (n, us) in sorted(per.items(), key=lambda kv: -kv[1][1])[:40]:
        print("  %-70s %4d  %8.1f us  avg %6.2f" % (name, n, us, us / n))
    report.update({"main_busy_us": busy, "main_idle_us": tot_gap, "n_main": len(lst)})
    # every Linear launch with its shape: the python executor issues the same C-ABI calls in the same order, so its
    # ccn_linear_planes / ccn_conv_planes argument lists line up with the linear_planes_kernel records of the native run
    from curvecloudnet_b200 import ops
    shapes, real_call = [], ops.call

    def spy(name, *a):
        if name == "ccn_linear_planes":
            shapes.append(("lin", a[24], a[25], a[3] + a[7] + a[11] + a[15], a[26], 1 + (a[7] > 0) + (a[11] > 0) + (a[15] > 0)))
        elif name == "ccn_conv_planes":
            shapes.append(("conv", 0))
        return real_call(name, *a)
    pym = randomize_model_(build_model(ModelBase, "nuscenes", precision=precision, native_executor=False), 1234).cuda().eval()
    ops.call = spy
    try:
        with torch.no_grad():
            torch.manual_seed(2)
            pym(scans[2])
    finally:
        ops.call = real_call
    lin = [k for k in last if "linear_planes_kernel" in k[2]]
    print("linear launches in order (%d kernels, %d python-side calls):" % (len(lin), len(shapes)))
    if len(lin) == len(shapes):
        tot = 0.0
        for k, sh in zip(lin, shapes):
            d = k[1] - k[0]
            tot += d
            if sh[0] == "lin":
                _, M, N, K, box, nseg = sh
                fl = 2.0 * M * N * K
                print("  t=%7.1f  M=%6d N=%5d K=%5d box=%3d segs=%d  %6.1f us  %6.1f TF/s alg  tiles=%d" %
                      (k[0] - t0, M, N, K, box, nseg, d, fl / d * 1e-6, ((M + 127) // 128) * ((N + box - 1) // box)))
            else:
                print("  t=%7.1f  conv layer %s  %6.1f us" % (k[0] - t0, sh[1:2], d))
        print("  total %.1f us" % tot)
    if out_json:
        report["kernels"] = [(k[0] - t0, k[1] - k[0], k[2][:60], str(k[3])) for k in last]
        json.dump(report, open(out_json, "w"))


if __name__ == "__main__":
    main()
