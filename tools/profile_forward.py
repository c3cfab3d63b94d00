"""Dev tool: per-entry-point device time of one nuScenes-shape forward (CUDA events around every C-ABI call).
usage: python tools/profile_forward.py [precision] [scans per forward]"""
import sys, os, collections
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from curvecloudnet_b200 import native
from curvecloudnet_b200.configs import build_model
from curvecloudnet_b200.model import ModelBase
from curvecloudnet_b200.synthetic import driving_scan_batch, randomize_model_

prec = sys.argv[1] if len(sys.argv) > 1 else "bf16"
model = build_model(ModelBase, "nuscenes", precision=prec)
randomize_model_(model, 1234)
model = model.cuda().eval()
nb = int(sys.argv[2]) if len(sys.argv) > 2 else 1
data = driving_scan_batch(list(range(nb))).to("cuda")
with torch.no_grad():
    for _ in range(3):
        torch.manual_seed(0); model(data)
torch.cuda.synchronize()
import time
ws = []
for _ in range(5):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    with torch.no_grad():
        torch.manual_seed(0); model(data)
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    ws.append(((t1 - t0) * 1e3, (time.perf_counter() - t0) * 1e3))
print("untimed forwards: host return / device done (ms):", ", ".join("%.2f/%.2f" % w for w in ws))
native.profile_begin(list(native.SIGNATURES), pad_cycles=20000)   # spin pad: the event pair brackets the kernel, not the launch gap
native._prof_footprint = []
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
import time
t0 = time.perf_counter()
e0.record()
with torch.no_grad():
    torch.manual_seed(0); model(data)
e1.record()
torch.cuda.synchronize()
wall = time.perf_counter() - t0
ev = list(native._prof_events)
foot = list(native._prof_footprint)
native._prof_footprint = None
prof = native.profile_end()
print("forward: %.3f ms device span, %.3f ms wall (with event overhead)" % (e0.elapsed_time(e1), wall * 1e3))
tot = sum(d["ms"] for d in prof.values())
print("sum of per-call device times: %.3f ms over %d calls" % (tot, sum(d["calls"] for d in prof.values())))
for k, d in sorted(prof.items(), key=lambda kv: -kv[1]["ms"]):
    print("  %-28s calls %4d  %8.3f ms  %5.1f%%" % (k, d["calls"], d["ms"], 100 * d["ms"] / tot))
print("linear calls (M, N, K) us TFLOP/s, algorithmic MB and GB/s:")
rows = []
for name, a, b, work, nbytes in ev:
    if name.startswith("ccn_linear"):
        rows.append((a.elapsed_time(b) * 1e3, work, nbytes))
shapes = native._prof_shapes
for (us, work, nbytes), shp in zip(rows, shapes):
    print("   M=%7d N=%5d K=%5d  %8.1f us  %7.1f TF/s  %8.2f MB  %7.0f GB/s"
          % (shp[0], shp[1], shp[2], us, work / us / 1e6, nbytes / 1e6, nbytes / us / 1e3))

# per-kernel achieved bandwidth: argument footprint (bytes of every tensor argument: inputs read + outputs written;
# gathers count their whole source) / device time, against the measured HBM copy peak
import json
try:
    peak = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    peak = 6650.0
agg = collections.OrderedDict()
for (name, a, b, work, _nb), fb in zip(ev, foot):
    d = agg.setdefault(name, [0.0, 0.0, 0])
    d[0] += fb; d[1] += a.elapsed_time(b) * 1e-3; d[2] += 1
print("per-entry-point footprint bandwidth (peak %.0f GB/s measured copy bandwidth):" % peak)
print("  %-28s %6s %12s %10s %10s %8s" % ("entry point", "calls", "MB / call", "us / call", "GB/s", "of peak"))
for name, (fb, sec, n) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    gbs = fb / sec / 1e9 if sec > 0 else 0.0
    print("  %-28s %6d %12.3f %10.1f %10.1f %7.1f%%" % (name, n, fb / n / 1e6, sec / n * 1e6, gbs, 100 * gbs / peak))
