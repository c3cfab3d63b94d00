"""Dev tool: where the host waits for the device inside one nuScenes-shape forward (every tolist / item on a CUDA tensor,
by call site)."""
import sys, os, time, collections, traceback
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from curvecloudnet_b200.configs import build_model
from curvecloudnet_b200.model import ModelBase
from curvecloudnet_b200.synthetic import driving_scan_batch, randomize_model_
model = build_model(ModelBase, "nuscenes", precision="bf16x3"); randomize_model_(model, 1234); model = model.cuda().eval()
data = driving_scan_batch([0]).to("cuda")
with torch.no_grad():
    for _ in range(3): model(data)
rec = collections.OrderedDict()
def wrap(name):
    orig = getattr(torch.Tensor, name)
    def f(self, *a, **k):
        if not self.is_cuda: return orig(self, *a, **k)
        t0 = time.perf_counter(); r = orig(self, *a, **k); dt = time.perf_counter() - t0
        st = traceback.extract_stack(limit=4)
        key = " <- ".join("%s:%d" % (os.path.basename(s.filename), s.lineno) for s in reversed(st[:-1]))
        rec.setdefault(key, []).append(dt * 1e6); return r
    setattr(torch.Tensor, name, f)
for n in ("tolist", "item"): wrap(n)
for it in range(4):
    rec.clear(); torch.cuda.synchronize(); t0 = time.perf_counter()
    with torch.no_grad(): torch.manual_seed(0); model(data)
    host = (time.perf_counter() - t0) * 1e3
print("host %.2f ms" % host)
for k, v in rec.items(): print("%7.1f us x%d  %s" % (sum(v), len(v), k))
print("total %.1f us" % sum(sum(v) for v in rec.values()))
