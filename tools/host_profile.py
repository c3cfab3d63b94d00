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
model = build_model(ModelBase, "nuscenes", precision=prec)
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
pr = cProfile.Profile()
with torch.no_grad():
    torch.manual_seed(0)
    pr.enable(); model(data); pr.disable()
torch.cuda.synchronize()
st = pstats.Stats(pr)
st.sort_stats("tottime").print_stats(28)
