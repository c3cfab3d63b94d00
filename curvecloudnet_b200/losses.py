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


# ---------------------------------------------------------------------------------------------------- object runs
SHAPENET_CATEGORY_NAMES = ['Airplane', 'Bag', 'Cap', 'Car', 'Chair', 'Earphone', 'Guitar', 'Knife', 'Lamp', 'Laptop',
                           'Motorbike', 'Mug', 'Pistol', 'Rocket', 'Skateboard', 'Table']        # src/run/globals.py:88-89
# torch_geometric.datasets.ShapeNet.seg_classes (the part labels of every category, used at shapenet_seg.py:93)
SHAPENET_SEG_CLASSES = {'Airplane': [0, 1, 2, 3], 'Bag': [4, 5], 'Cap': [6, 7], 'Car': [8, 9, 10, 11], 'Chair': [12, 13, 14, 15],
                        'Earphone': [16, 17, 18], 'Guitar': [19, 20, 21], 'Knife': [22, 23], 'Lamp': [24, 25, 26, 27],
                        'Laptop': [28, 29], 'Motorbike': [30, 31, 32, 33, 34, 35], 'Mug': [36, 37], 'Pistol': [38, 39, 40],
                        'Rocket': [41, 42, 43], 'Skateboard': [44, 45, 46], 'Table': [47, 48, 49]}
_part_tables = {}


def object_seg_loss(pred: torch.Tensor, gt: torch.Tensor):
    """seg_loss of src/run/shapenet_seg.py:182-186 on raw logits: log_softmax, per-point NLL, its mean -> (loss, per-point)."""
    return seg_loss(pred, gt, ignore=-100)


def part_iou(scores: torch.Tensor, y: torch.Tensor, ptr: torch.Tensor, categories, seg_classes=None, names=None,
             absent_score: float = 1.0) -> torch.Tensor:
    """The per-object loop of shapenet_seg.py:88-99 as one launch: for object b (points ptr[b]..ptr[b+1], category
    categories[b]) jaccard_index(scores[:, part].argmax(-1), y_map[y], num_classes=len(part), absent_score) -> (B,) float32
    on the device (the reference appends `iou.item()` per object: one host sync each)."""
    seg_classes = SHAPENET_SEG_CLASSES if seg_classes is None else seg_classes
    names = SHAPENET_CATEGORY_NAMES if names is None else names
    dev = scores.device
    key = (tuple((nme, tuple(seg_classes[nme])) for nme in names), str(dev))
    if key not in _part_tables:
        flat, pp = [], [0]
        for nme in names:
            flat += list(seg_classes[nme])
            pp.append(len(flat))
        _part_tables[key] = (torch.tensor(flat, dtype=torch.int32, device=dev), torch.tensor(pp, dtype=torch.int32, device=dev),
                             max(pp[i + 1] - pp[i] for i in range(len(names))))
    parts, parts_ptr, pmax = _part_tables[key]
    sc, ld = ops._rows(scores)
    cat = torch.as_tensor(categories, dtype=torch.int64, device=dev)
    B = cat.shape[0]
    out = torch.empty((B,), dtype=torch.float32, device=dev)
    flags = torch.empty((1,), dtype=torch.int32, device=dev)
    call("ccn_part_iou", sc, ld, ops._i64(y), ptr.to(torch.int32), cat, B, parts, parts_ptr, len(names), pmax, float(absent_score),
         out, 0, flags)
    return out


def category_iou_summary(ious: torch.Tensor, categories: torch.Tensor):
    """shapenet_seg.py:103-111 on the collected per-object IoUs (host-size vectors): per-category mean (scatter mean sized
    max category + 1), instance mean, class mean, median."""
    cat = categories.to(torch.int64)
    n = int(cat.max()) + 1 if cat.numel() else 0
    tot = torch.zeros(n, dtype=ious.dtype, device=ious.device).scatter_add_(0, cat, ious)
    cnt = torch.zeros(n, dtype=ious.dtype, device=ious.device).scatter_add_(0, cat, torch.ones_like(ious)).clamp(min=1)
    per_cat = tot / cnt
    return {"per_category": per_cat, "inst_iou": float(ious.mean()), "mean_iou": float(per_cat.mean()),
            "median_iou": float(per_cat.median())}


def classification_metrics(log_probs: torch.Tensor, labels: torch.Tensor, n_classes: int):
    """The epilogue of src/run/shapenet_classification.py:77-93 for a batch / a whole split: NLL loss on log-probabilities,
    the confusion histogram of the arg-max (fused kernel), accuracy (= sklearn accuracy_score) and class-mean accuracy
    (= sklearn balanced_accuracy_score: mean recall over the classes that occur).  -> dict with the histogram on the device."""
    hist = torch.zeros((n_classes, n_classes), dtype=torch.int64, device=log_probs.device)
    sc, ld = ops._rows(log_probs)
    call("ccn_confusion_hist", sc, ld, n_classes, 0, ops._i64(labels), sc.shape[0], n_classes, 0, 0, hist)
    h = hist.double()
    support = h.sum(1)
    diag = torch.diagonal(h)
    seen = support > 0
    return {"hist": hist, "accuracy": float(diag.sum() / h.sum().clamp(min=1)),
            "class_accuracy": float((diag[seen] / support[seen]).mean()) if bool(seen.any()) else 0.0}


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
