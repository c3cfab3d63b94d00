"""Loss / metrics epilogue of the segmentation runs on the device:
`seg_loss` (src/run/nuscenes_seg.py:229-243: weighted NLL with ignore_index + Lovasz-softmax), `lovasz_softmax_flat`
(src/models/utils/lovasz_losses.py:175-204), `fast_hist` / `fast_hist_crop` / `per_class_iu` (src/utils/utils.py:4-12,
nuscenes_seg.py:245-249) and the test-time-augmentation loop of `val()` (nuscenes_seg.py:86-116).  All arithmetic on
the per-point tensors is in csrc/loss_metrics.cu / train_ops.cu."""
from __future__ import annotations

from typing import Optional

import torch
from torch.autograd import Function

from . import autograd as A
from . import native, ops
from .native import call


class _LovaszSoftmax(Function):
    @staticmethod
    def forward(ctx, logits, target, ignore_index):
        logits, ld = ops._rows(logits)
        target = ops._i64(target)
        n, C = logits.shape
        npad = int(native.load().ccn_lovasz_pad(n))
        dev = logits.device
        keys = torch.empty((C * npad,), dtype=torch.int64, device=dev)
        fg_count = torch.empty((C,), dtype=torch.int32, device=dev)
        lse = torch.empty((n,), dtype=torch.float32, device=dev)
        grad = torch.empty((C, n), dtype=torch.float32, device=dev)
        acc = torch.empty((2,), dtype=torch.float64, device=dev)
        loss = torch.empty((1,), dtype=torch.float32, device=dev)
        call("ccn_lovasz_softmax_fwd", logits, ld, target, n, C, int(ignore_index), keys, fg_count, lse, grad, acc, loss)
        ctx.save_for_backward(logits, target, lse, grad, fg_count, acc)
        ctx.ignore = int(ignore_index)
        return loss.view(())

    @staticmethod
    def backward(ctx, g):
        logits, target, lse, grad, fg_count, acc = ctx.saved_tensors
        logits, ld = ops._rows(logits)
        n, C = logits.shape
        d = torch.empty((n, C), dtype=torch.float32, device=logits.device)
        call("ccn_lovasz_softmax_bwd", logits, ld, target, n, C, ctx.ignore, lse, grad, fg_count, acc,
             g.reshape(1).contiguous().float(), d, C, 0)
        return d, None, None


def lovasz_softmax_flat(logits: torch.Tensor, labels: torch.Tensor, ignore_index: int = -1) -> torch.Tensor:
    """lovasz_softmax_flat(F.softmax(logits)[labels != ignore], labels[labels != ignore], classes='present')
    (lovasz_losses.py:175-204); the softmax and the ignore filter of seg_loss are fused in.  Elements with EXACTLY equal
    errors are ordered by (foreground, point index) - the loss does not depend on that order, the sub-gradient of the
    tied elements does (torch.sort leaves it unspecified)."""
    if logits.shape[0] == 0:
        return logits.sum() * 0.0
    return _LovaszSoftmax.apply(logits, labels, ignore_index)


def seg_loss(pred: torch.Tensor, gt: torch.Tensor, ignore: int = 0, use_lovasz: bool = False,
             ce_weights: Optional[torch.Tensor] = None):
    """seg_loss of src/run/nuscenes_seg.py:229-243 on raw logits `pred`: -> (loss, per-point NLL)."""
    perpnt = torch.empty((pred.shape[0],), dtype=torch.float32, device=pred.device)
    loss = A.nll_loss(pred, gt, ce_weights, ignore, perpnt)
    if use_lovasz:
        loss = loss + lovasz_softmax_flat(pred, gt, ignore)
    return loss, perpnt


# ---------------------------------------------------------------------------------------------------- metrics
def fast_hist(pred: torch.Tensor, label: torch.Tensor, n: int, hist: Optional[torch.Tensor] = None) -> torch.Tensor:
    """fast_hist (utils.py:4-8): (n, n) int64 counts [label][pred] of the points with 0 <= label < n; `hist` accumulates."""
    if hist is None:
        hist = torch.zeros((n, n), dtype=torch.int64, device=pred.device)
    call("ccn_confusion_hist", 0, 0, 0, ops._i64(pred), ops._i64(label), pred.shape[0], n, 0, 0, hist)
    return hist


def fast_hist_crop(pred: torch.Tensor, label: torch.Tensor, unique_label) -> torch.Tensor:
    """fast_hist_crop (nuscenes_seg.py:245-249)."""
    ul = torch.as_tensor(unique_label, dtype=torch.int64, device=pred.device)
    hist = fast_hist(pred, label, int(ul.max()) + 2)
    return hist[ul + 1][:, ul + 1]


def per_class_iu(hist: torch.Tensor) -> torch.Tensor:
    """per_class_iu (utils.py:11-12) on the (classes x classes) histogram (a 16 x 16 host-size array)."""
    h = hist.double()
    d = torch.diagonal(h)
    return d / (h.sum(1) + h.sum(0) - d)


# ---------------------------------------------------------------------------------------------------- val() TTA
TTA_FLIPS = ((1.0, 1.0), (-1.0, 1.0), (1.0, -1.0), (-1.0, -1.0))       # nuscenes_seg.py:93-98
TTA_SCALES = (1.0, 0.95, 1.05)                                         # :100-103


def tta_stack(data, factors):
    """The augmented copies of `data` as ONE batch: copy g (xy factors `factors[g]`) becomes samples g*B .. g*B + B - 1.
    Per-point and per-sample tensors are repeated; positions are scaled by ccn_scale_xy."""
    from .model import Batch
    pos0 = ops._f32(data.pos)
    n = pos0.shape[0]
    G = len(factors)
    nb = int(data.batch[-1].item()) + 1 if n else 0                     # host sync (once per stack)
    pos = torch.empty((G * n, 3), dtype=torch.float32, device=pos0.device)
    for g, (sx, sy) in enumerate(factors):
        call("ccn_scale_xy", pos0, n, float(sx), float(sy), pos[g * n:(g + 1) * n])
    out = Batch()
    for k, v in data.__dict__.items():
        if k == "pos":
            out.pos = pos
        elif k == "batch":
            out.batch = (v.unsqueeze(0) + torch.arange(G, device=v.device, dtype=v.dtype).unsqueeze(1) * nb).reshape(-1)
        elif torch.is_tensor(v) and v.dim() >= 1 and v.shape[0] in (n, nb):
            out.__dict__[k] = v.repeat(G, *([1] * (v.dim() - 1)))
        else:
            out.__dict__[k] = v
    return out


@torch.no_grad()
def tta_forward(model, data, flips=TTA_FLIPS, scales=TTA_SCALES, group: int = 1):
    """The augmentation loop of val() (nuscenes_seg.py:86-116): 12 forwards over xy-flipped / xy-scaled positions,
    out = log_softmax(sum(outs) / 12), predict = argmax.  -> (out (N, C) log-probabilities, predict_labels (N) int64).

    group = 1: one forward per augmented copy, every forward draws its own sampling randomness exactly like the
    reference loop (draw-for-draw reproducible against it under a seed).
    group = G > 1: G augmented copies are stacked into one batch of G*B samples per forward (12 / G forwards).  Every op
    of the path is per sample and eval-mode BatchNorm is frozen, so each copy is processed exactly as if alone; only the
    ORDER in which the random sampling draws are consumed differs from the loop (the reference's draws are unseeded).
    The stacked forward runs the device at its batched rate instead of the launch-bound B=1 rate."""
    pos0 = ops._f32(data.pos)
    n = pos0.shape[0]
    factors = [(fx * sc, fy * sc) for fx, fy in flips for sc in scales]
    acc, first = None, True
    if group <= 1:
        orig = data.pos
        try:
            for sx, sy in factors:
                pos = torch.empty_like(pos0)
                call("ccn_scale_xy", pos0, n, float(sx), float(sy), pos)
                data.pos = pos
                out = model(data, labels=getattr(data, "y", None))
                out, ld = ops._rows(out)
                if acc is None:
                    acc = torch.empty((n, out.shape[1]), dtype=torch.float32, device=out.device)
                call("ccn_accumulate", out, ld, n, out.shape[1], acc, acc.shape[1], int(first))
                first = False
        finally:
            data.pos = orig               # the caller's batch gets its positions back even if a forward raises
    else:
        for g0 in range(0, len(factors), group):
            stacked = tta_stack(data, factors[g0:g0 + group])
            out = model(stacked, labels=getattr(stacked, "y", None))
            out, ld = ops._rows(out)
            if acc is None:
                acc = torch.empty((n, out.shape[1]), dtype=torch.float32, device=out.device)
            for g in range(len(factors[g0:g0 + group])):
                call("ccn_accumulate", out[g * n:(g + 1) * n], ld, n, out.shape[1], acc, acc.shape[1], int(first))
                first = False
    C = acc.shape[1]
    logp = torch.empty_like(acc)
    pred = torch.empty((n,), dtype=torch.int64, device=acc.device)
    call("ccn_mean_logsoftmax", acc, C, n, C, 1.0 / len(factors), logp, C, pred)
    return logp, pred
