"""Loader for the committed fixtures made by oracle/make_golden.py (int32 -> int64 widening)."""
import os

import torch

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _widen(obj):
    if torch.is_tensor(obj):
        return obj.to(torch.int64) if obj.dtype == torch.int32 else obj
    if isinstance(obj, dict):
        return {k: _widen(v) for k, v in obj.items()}
    return obj


def load(name):
    return _widen(torch.load(os.path.join(GOLDEN, name), weights_only=False))


class FixedDraws:
    """Draws object that replays recorded values (same interface as oracle.ccn_point_ops.Draws)."""

    def __init__(self, u=None, noise=None, starts=None):
        self.u, self.noise, self.starts = u, noise, list(starts) if starts is not None else None

    def curve_phase(self):
        return self.u

    def voxel_noise(self, n):
        assert self.noise.shape[0] == n
        return self.noise

    def fps_start(self, length):
        return int(self.starts.pop(0))


def rel_err(a, b):
    """max |a-b| / max |b|  (the 'relative' of BASELINE.json's 1e-4 / 2e-2 tolerances)."""
    return float((a.double() - b.double()).abs().max() / b.double().abs().max().clamp(min=1e-30))


def grad_projection(name, numel):
    """Same +-1 vector per parameter name as oracle/make_golden.py grad_projection."""
    import zlib
    gen = torch.Generator().manual_seed(zlib.crc32(name.encode()))
    return torch.randint(0, 2, (numel,), generator=gen).double() * 2 - 1


def check_train_golden(model, logits, loss, rec, logit_tol, grad_tol, label=""):
    """Compare one training step with a record of tests/golden/train.pt (made from the REAL reference sources):
    logits, loss, and for every parameter the gradient's L2 norm, its projection on a fixed +-1 vector and (for
    vectors) every entry.  Errors are relative to the tensor's norm, floored by 1e-3 of the largest norm."""
    e_logits = rel_err(logits.detach().cpu(), rec["logits"])
    e_loss = abs(float(loss) - rec["loss"]) / abs(rec["loss"])
    params = dict(model.named_parameters())
    nmax = max(g["norm"] for g in rec["grads"].values())
    worst = ("", 0.0)
    for k, g in rec["grads"].items():
        assert params[k].grad is not None, "no gradient for %s" % k
        gd = params[k].grad.detach().cpu().double().flatten()
        scale = max(g["norm"], 1e-3 * nmax)
        errs = [abs(float(gd.norm()) - g["norm"]) / scale,
                abs(float((gd * grad_projection(k, gd.numel())).sum()) - g["proj"]) / (scale * gd.numel() ** 0.5) * 3]
        if "full" in g:
            errs.append(float((gd - g["full"].double().flatten()).norm()) / scale)
        if max(errs) > worst[1]:
            worst = (k, max(errs))
    print("%s train golden: logits %.3e loss %.3e worst gradient %s %.3e" % (label, e_logits, e_loss, worst[0], worst[1]))
    assert e_logits <= logit_tol and e_loss <= logit_tol, (e_logits, e_loss)
    assert worst[1] <= grad_tol, "gradient of %s off by %.3e" % worst


def loss_case(name, n, C, seed):
    """Inputs of a tests/golden/losses.pt record (logits, labels with ~30 % ignored zeros, class weights)."""
    gen = torch.Generator().manual_seed(seed)
    logits = torch.randn(n, C, generator=gen) * 2.5
    gt = torch.randint(0, C, (n,), generator=gen)
    gt[torch.rand(n, generator=gen) < 0.3] = 0
    if name == "small":
        gt[gt == 5] = 6                                                # an absent class
    w = torch.rand(C, generator=gen) + 0.5
    return logits, gt, w
