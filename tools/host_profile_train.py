"""Dev tool: cProfile of the host side of one training step (ShapeNet-seg B=16)."""
import cProfile
import os
import pstats
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from curvecloudnet_b200 import autograd as A  # noqa: E402
from curvecloudnet_b200.configs import build_model  # noqa: E402
from curvecloudnet_b200.model import ModelBase  # noqa: E402
from curvecloudnet_b200.synthetic import object_scan_batch, randomize_model_  # noqa: E402

model = build_model(ModelBase, "shapenet-seg", precision=sys.argv[1] if len(sys.argv) > 1 else "bf16x3")
randomize_model_(model, 1234)
model = model.cuda().train()
data = object_scan_batch(list(range(16))).to("cuda")
target = torch.randint(0, 50, (data.pos.shape[0],), device="cuda")


def step():
    for p in model.parameters():
        p.grad = None
    torch.manual_seed(0)
    loss = A.nll_loss(model(data), target)
    loss.backward()


for _ in range(3):
    step()
torch.cuda.synchronize()
pr = cProfile.Profile()
pr.enable()
step()
pr.disable()
torch.cuda.synchronize()
st = pstats.Stats(pr)
st.sort_stats("tottime").print_stats(35)
st.sort_stats("cumtime").print_stats(45)
