"""Per-parameter gradient comparison of one training step: ours (GPU) vs the fp32 CPU oracle vs the oracle in fp64
(the last tells how much of a difference is conditioning).  usage: python tests/dev/debug_train.py <config> <precision>"""
import copy
import os
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from curvecloudnet_b200 import autograd as A                     # noqa: E402
from curvecloudnet_b200.configs import build_model               # noqa: E402
from curvecloudnet_b200.model import ModelBase                   # noqa: E402
from curvecloudnet_b200.synthetic import driving_scan_batch, object_scan_batch, randomize_model_   # noqa: E402
from oracle.ccn_model import ModelBase as OracleModel            # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "shapenet-seg"
precision = sys.argv[2] if len(sys.argv) > 2 else "fp32"
sub = int(sys.argv[3]) if len(sys.argv) > 3 else 1500
freeze = len(sys.argv) > 4 and sys.argv[4] == "freeze"


def freeze_bn(m):
    if freeze:
        for mod in m.modules():
            if isinstance(mod, torch.nn.BatchNorm1d):
                mod.eval()
    return m

data = driving_scan_batch([3], subsample=sub) if name in ("nuscenes", "kitti", "a2d2") else object_scan_batch([1, 2])

ref = build_model(OracleModel, name)
randomize_model_(ref, 1234)
ref.train()
freeze_bn(ref)
state = copy.deepcopy(ref.state_dict())


def run_ref(model, d):
    torch.manual_seed(11)
    out = model(d)
    gen = torch.Generator().manual_seed(5)
    target = torch.randint(0, out.shape[1], (out.shape[0],), generator=gen)
    loss = F.nll_loss(F.log_softmax(out, 1), target)
    loss.backward()
    return out.detach(), float(loss), target


want, loss_ref, target = run_ref(ref, data)
g32 = {k: (p.grad.clone() if p.grad is not None else None) for k, p in ref.named_parameters()}

g64, want64 = None, None

model = build_model(ModelBase, name, precision=precision)
model.load_state_dict(state, strict=True)
model = freeze_bn(model.cuda().train())
torch.manual_seed(11)
got = model(data.to("cuda"))
loss = A.nll_loss(got, target.cuda())
loss.backward()


def rel(a, b):
    return float((a.double() - b.double()).abs().max() / b.double().abs().max().clamp(min=1e-30))


print("logits ours-vs-ref32 %.3e" % rel(got.detach().cpu(), want), "loss", float(loss), loss_ref)
if want64 is not None:
    print("logits ref32-vs-ref64 %.3e  ours-vs-ref64 %.3e" % (rel(want, want64), rel(got.detach().cpu(), want64)))
rows = []
for k, p in model.named_parameters():
    r = g32[k]
    if r is None:
        continue
    e = rel(p.grad.cpu(), r)
    e64 = rel(r, g64[k]) if g64 is not None and g64[k] is not None else float("nan")
    o64 = rel(p.grad.cpu(), g64[k]) if g64 is not None and g64[k] is not None else float("nan")
    rows.append((e, e64, o64, float(r.abs().max()), k, tuple(p.shape)))
gmax = max(r[3] for r in rows)
print("params with max|g| < 1e-4 of the largest gradient entry (%.3e) are left out of the table" % gmax)
rows = [r for r in rows if r[3] >= 1e-4 * gmax]
if os.environ.get("ORDER") != "model":
    rows.sort(reverse=True)
print("%10s %10s %10s %10s  %s" % ("ours/ref32", "ref32/ref64", "ours/ref64", "max|g|", "param"))
for e, e64, o64, m, k, shp in (rows if os.environ.get('ORDER') == 'model' else rows[:40]):
    print("%10.3e %10.3e %10.3e %10.3e  %s %s" % (e, e64, o64, m, k, shp))
print("median ours/ref32 %.3e" % sorted(r[0] for r in rows)[len(rows) // 2])
