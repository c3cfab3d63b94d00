"""Dev tool: where does the HOST time of one nuScenes-shape forward go (cProfile + time inside host syncs)."""
import cProfile
import os
import pstats
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from curvecloudnet_b200.configs import build_model
from curvecloudnet_b200.model import ModelBase
from curvecloudnet_b200.synthetic import driving_scan_batch, randomize_model_

prec = sys.argv[1] if len(sys.argv) > 1 else "bf16x3"
# the native step executor (csrc/executor.cu): host wall time of ccn_forward_run, split into waiting and working
nmodel = randomize_model_(build_model(ModelBase, "nuscenes", precision=prec), 1234).cuda().eval()
ndata = driving_scan_batch([0]).to("cuda")
with torch.no_grad():
    for i in range(6):
        torch.manual_seed(0)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        nmodel(ndata)
        t1 = time.perf_counter()
        torch.cuda.synchronize()
        t2 = time.perf_counter()
        st = nmodel.__dict__["_native"][1].host_stats()
        if i >= 3:
            print("native executor: forward call %.2f ms (device done %.2f ms); inside ccn_forward_run %.2f ms = %.2f ms host work + "
                  "%.2f ms blocked in %d waits for data-dependent sizes" % ((t1 - t0) * 1e3, (t2 - t0) * 1e3, st["total_ms"],
                                                                          st["total_ms"] - st["blocked_ms"], st["blocked_ms"], st["syncs"]))
del nmodel
print("python executor (native_executor=False):")
model = build_model(ModelBase, "nuscenes", precision=prec, native_executor=False)
randomize_model_(model, 1234)
model = model.cuda().eval()
data = driving_scan_batch([0]).to("cuda")
with torch.no_grad():
    for _ in range(3):
        torch.manual_seed(0); model(data)
torch.cuda.synchronize()

sync_t = [0.0, 0]
_tolist, _item = torch.Tensor.tolist, torch.Tensor.item


def tolist(self):
    t = time.perf_counter(); r = _tolist(self); sync_t[0] += time.perf_counter() - t; sync_t[1] += 1; return r


def item(self):
    t = time.perf_counter(); r = _item(self); sync_t[0] += time.perf_counter() - t; sync_t[1] += 1; return r


torch.Tensor.tolist, torch.Tensor.item = tolist, item
with torch.no_grad():
    for _ in range(3):
        sync_t[:] = [0.0, 0]
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        torch.manual_seed(0); model(data)
        t1 = time.perf_counter()
        torch.cuda.synchronize()
        print("forward host %.2f ms, of which %.2f ms inside %d tolist/item calls" % ((t1 - t0) * 1e3, sync_t[0] * 1e3, sync_t[1]))
torch.Tensor.tolist, torch.Tensor.item = _tolist, _item

# time inside the foreign calls (kernel launches) and inside torch.empty
from curvecloudnet_b200 import native
ff = {}
for name, (fn, ptrs) in list(native._fns.items()):
    def mk(fn, name):
        def w(*a):
            t = time.perf_counter(); r = fn(*a); d = time.perf_counter() - t
            e = ff.setdefault(name, [0.0, 0]); e[0] += d; e[1] += 1
            return r
        return w
    native._fns[name] = (mk(fn, name), ptrs)
_empty = torch.empty
et = [0.0, 0]
def empty(*a, **k):
    t = time.perf_counter(); r = _empty(*a, **k); et[0] += time.perf_counter() - t; et[1] += 1; return r
torch.empty = empty
with torch.no_grad():
    for rep in range(2):
        ff.clear(); et[:] = [0.0, 0]
        torch.cuda.synchronize(); t0 = time.perf_counter()
        torch.manual_seed(0); model(data)
        t1 = time.perf_counter(); torch.cuda.synchronize()
    tot = sum(v[0] for v in ff.values())
    print("forward host %.2f ms (with wrappers); inside foreign calls %.2f ms over %d calls; torch.empty %.2f ms over %d" % ((t1 - t0) * 1e3, tot * 1e3, sum(v[1] for v in ff.values()), et[0] * 1e3, et[1]))
    for k, v in sorted(ff.items(), key=lambda kv: -kv[1][0])[:12]:
        print("   %-28s %3d calls %7.1f us each" % (k, v[1], v[0] / v[1] * 1e6))
torch.empty = _empty
for name in list(native._fns):
    native._fns.pop(name)
pr = cProfile.Profile()
with torch.no_grad():
    torch.manual_seed(0)
    pr.enable(); model(data); pr.disable()
torch.cuda.synchronize()
st = pstats.Stats(pr)
st.sort_stats("tottime").print_stats(45)
st.sort_stats("cumulative").print_stats(45)
