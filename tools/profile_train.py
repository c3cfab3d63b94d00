"""Dev tool: per-entry-point device time of one training step (fwd + loss + bwd) - CUDA events around every C-ABI call.
usage: python tools/profile_train.py [precision] [config] [batch]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from curvecloudnet_b200 import autograd as A  # noqa: E402
from curvecloudnet_b200 import native  # noqa: E402
from curvecloudnet_b200.configs import build_model  # noqa: E402
from curvecloudnet_b200.model import ModelBase  # noqa: E402
from curvecloudnet_b200.synthetic import driving_scan_batch, object_scan_batch, randomize_model_  # noqa: E402

prec = sys.argv[1] if len(sys.argv) > 1 else "bf16x3"
name = sys.argv[2] if len(sys.argv) > 2 else "shapenet-seg"
B = int(sys.argv[3]) if len(sys.argv) > 3 else 16
model = build_model(ModelBase, name, precision=prec)
randomize_model_(model, 1234)
model = model.cuda().train()
data = (driving_scan_batch(list(range(B))) if name in ("nuscenes", "kitti", "a2d2") else object_scan_batch(list(range(B)))).to("cuda")
n_cls = model.n_out
target = torch.randint(0, n_cls, (data.pos.shape[0],), device="cuda")


def step():
    for p in model.parameters():
        p.grad = None
    torch.manual_seed(0)
    loss = A.nll_loss(model(data), target)
    loss.backward()
    return loss


for _ in range(3):
    step()
torch.cuda.synchronize()
for _ in range(3):
    t0 = time.perf_counter()
    step()
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    print("step: host return %.2f ms, device done %.2f ms" % ((t1 - t0) * 1e3, (time.perf_counter() - t0) * 1e3))
print("peak memory %.2f GB" % (torch.cuda.max_memory_allocated() / 2 ** 30))
native.profile_begin(list(native.SIGNATURES), pad_cycles=20000)
native._prof_footprint = []
step()
torch.cuda.synchronize()
ev = list(native._prof_events)
foot = list(native._prof_footprint)
native._prof_footprint = None
shapes = list(native._prof_shapes)
prof = native.profile_end()
tot = sum(d["ms"] for d in prof.values())
print("%s %s B=%d (%d points): sum of per-call device times %.3f ms over %d calls"
      % (name, prec, B, data.pos.shape[0], tot, sum(d["calls"] for d in prof.values())))
for k, d in sorted(prof.items(), key=lambda kv: -kv[1]["ms"]):
    print("  %-28s calls %4d  %8.3f ms  %5.1f%%" % (k, d["calls"], d["ms"], 100 * d["ms"] / tot))
rows = [(n, a.elapsed_time(b) * 1e3, w) for n, a, b, w, _nb in ev if n in native._WORK]
print("GEMM-shaped calls (M, N, K) us TFLOP/s:")
for (n, us, w), shp in zip(rows, shapes):
    print("   %-18s M=%7d N=%5d K=%5d  %8.1f us  %7.1f TF/s" % (n, shp[0], shp[1], shp[2], us, w / us / 1e6))

# per-entry-point footprint bandwidth (bytes of every tensor argument / device time) against the measured HBM copy peak
import collections
import json
import os
try:
    peak = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    peak = 6650.0
agg = collections.OrderedDict()
for (nm, a, b, w, _nb), fb in zip(ev, foot):
    d = agg.setdefault(nm, [0.0, 0.0, 0])
    d[0] += fb; d[1] += a.elapsed_time(b) * 1e-3; d[2] += 1
print("per-entry-point footprint bandwidth (peak %.0f GB/s measured copy bandwidth):" % peak)
print("  %-28s %6s %12s %10s %10s %8s" % ("entry point", "calls", "MB / call", "us / call", "GB/s", "of peak"))
for nm, (fb, sec, n) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    gbs = fb / sec / 1e9 if sec > 0 else 0.0
    print("  %-28s %6d %12.3f %10.1f %10.1f %7.1f%%" % (nm, n, fb / n / 1e6, sec / n * 1e6, gbs, 100 * gbs / peak))
