"""Dev tool: per-step feature comparison (train mode) between the CUDA model and the CPU oracle on the train.pt case."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from golden_utils import load, rel_err  # noqa: E402
from curvecloudnet_b200.configs import build_model  # noqa: E402
from curvecloudnet_b200.model import Batch, ModelBase  # noqa: E402
from curvecloudnet_b200.synthetic import randomize_model_  # noqa: E402
from oracle.ccn_model import ModelBase as OracleModel  # noqa: E402
from oracle.tp_nn import Batch as OBatch  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "nuscenes"
g = load("train.pt")[name]
kw = dict(x=g["x"], pos=g["pos"], batch=g["batch"], curve_idxs=g["curve"])
if g["labels"] is not None:
    kw["labels"] = g["labels"]
outs = {}


def hook(tag):
    def f(mod, args, out):
        o = out[0] if isinstance(out, tuple) else out
        outs.setdefault(tag, []).append(o.detach().cpu() if torch.is_tensor(o) else None)
    return f


ref = build_model(OracleModel, name); randomize_model_(ref, g["weight_seed"]); ref.train()
ours = build_model(ModelBase, name, precision="fp32"); randomize_model_(ours, g["weight_seed"]); ours = ours.cuda().train()
for i, (a, b) in enumerate(zip(ref.steps, ours.steps)):
    a.register_forward_hook(hook("ref")); b.register_forward_hook(hook("ours"))
torch.manual_seed(g["rng_seed"]); lr = ref(OBatch(**kw))
torch.manual_seed(g["rng_seed"]); lo = ours(Batch(**kw).to("cuda"))
for i, (a, b) in enumerate(zip(outs["ref"], outs["ours"])):
    if a is None or b is None:
        continue
    same = a.shape == b.shape
    print("step %2d %-14s %-18s %s" % (i, ours.step_names[i], tuple(a.shape), "%.3e" % rel_err(b, a) if same else "SHAPE %s" % (tuple(b.shape),)))
print("logits", rel_err(lo.detach().cpu(), lr.detach()), "golden", rel_err(lo.detach().cpu(), g["batch_stats"]["logits"]))
