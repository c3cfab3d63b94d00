"""Three seeded headline forwards (native executor) and nothing else: the process ncu attaches to when a particular launch of
the forward is captured (`-k regex:linear_planes_kernel -s <73 * 2 + index> -c <n>`; tools/timeline.py prints the index).
Usage: python tools/ncu_forward.py [precision] [forwards]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from curvecloudnet_b200.configs import build_model                      # noqa: E402
from curvecloudnet_b200.model import ModelBase                         # noqa: E402
from curvecloudnet_b200.synthetic import driving_scan_batch, randomize_model_   # noqa: E402

precision = sys.argv[1] if len(sys.argv) > 1 else "bf16x3"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 3
model = randomize_model_(build_model(ModelBase, "nuscenes", precision=precision), 1234).cuda().eval()
scan = driving_scan_batch([2]).to("cuda")
with torch.no_grad():
    for i in range(n):
        torch.manual_seed(2)
        y = model(scan)
torch.cuda.synchronize()
print("ok", tuple(y.shape))
