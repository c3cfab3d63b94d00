"""CPU restatement of the loss / metrics epilogue around the path.
TEST INFRASTRUCTURE - see oracle/__init__.py.

  lovasz_grad, lovasz_softmax_flat   src/models/utils/lovasz_losses.py:21-33, 175-204
  seg_loss                           src/run/nuscenes_seg.py:229-243
  fast_hist, per_class_iu            src/utils/utils.py:4-12;  fast_hist_crop nuscenes_seg.py:245-249
  tta_forward                        the augmentation loop of val(), nuscenes_seg.py:86-116
"""
from __future__ import annotations

import torch
import torch.nn.functional as F


def lovasz_grad(gt_sorted):
    """lovasz_losses.py:21-33."""
    p = len(gt_sorted)
    gts = gt_sorted.sum()
    intersection = gts - gt_sorted.float().cumsum(0)
    union = gts + (1 - gt_sorted).float().cumsum(0)
    jaccard = 1.0 - intersection / union
    if p > 1:
        jaccard[1:p] = jaccard[1:p] - jaccard[0:-1]
    return jaccard


def lovasz_softmax_flat(probas, labels):
    """lovasz_losses.py:175-204 with classes='present' (the only mode seg_loss uses)."""
    if probas.numel() == 0:
        return probas * 0.0
    losses = []
    for c in range(probas.size(1)):
        fg = (labels == c).float()
        if fg.sum() == 0:
            continue
        errors = (fg - probas[:, c]).abs()
        errors_sorted, perm = torch.sort(errors, 0, descending=True)
        losses.append(torch.dot(errors_sorted, lovasz_grad(fg[perm]).detach()))
    return sum(losses) / len(losses) if losses else probas.sum() * 0.0


def seg_loss(pred, gt, ignore=0, use_lovasz=False, ce_weights=None):
    """nuscenes_seg.py:229-243."""
    pred_log_soft = F.log_softmax(pred, dim=-1)
    loss = F.nll_loss(pred_log_soft, gt, reduction="mean", ignore_index=ignore, weight=ce_weights)
    perpnt_loss = F.nll_loss(pred_log_soft, gt, reduction="none")
    if use_lovasz:
        pred_soft = F.softmax(pred, dim=-1)
        keep = gt != ignore
        loss = loss + torch.mean(lovasz_softmax_flat(pred_soft[keep], gt[keep]))
    return loss, perpnt_loss


def fast_hist(pred, label, n):
    """utils.py:4-8."""
    k = (label >= 0) & (label < n)
    return torch.bincount(n * label[k] + pred[k], minlength=n ** 2)[:n ** 2].reshape(n, n)


def fast_hist_crop(output, target, unique_label):
    """nuscenes_seg.py:245-249."""
    ul = torch.as_tensor(unique_label, dtype=torch.int64)
    hist = fast_hist(output.flatten(), target.flatten(), int(ul.max()) + 2)
    return hist[ul + 1, :][:, ul + 1]


def per_class_iu(hist):
    """utils.py:11-12."""
    h = hist.double()
    d = torch.diagonal(h)
    return d / (h.sum(1) + h.sum(0) - d)


@torch.no_grad()
def tta_forward(model, data):
    """nuscenes_seg.py:86-116: 4 xy-flips x 3 xy-scales, out = log_softmax(sum(outs) / 12), predict = argmax."""
    orig = data.pos.clone()
    outs = []
    for axis_flips in range(4):
        for scales in range(3):
            pos = orig.clone()
            if axis_flips == 1:
                pos[:, 0] *= -1
            elif axis_flips == 2:
                pos[:, 1] *= -1
            elif axis_flips == 3:
                pos[:, :2] *= -1
            if scales == 1:
                pos[:, :2] *= 0.95
            elif scales == 2:
                pos[:, :2] *= 1.05
            data.pos = pos
            outs.append(model(data, labels=getattr(data, "y", None)))
    data.pos = orig
    out = F.log_softmax(sum(outs) / 12, dim=-1)
    return out, torch.argmax(out, dim=-1)


def tta_forward_stacked(model, data, group):
    """The same 12 augmented copies, `group` of them stacked into one batch per forward (copy g -> samples g*B ..):
    the CPU statement of curvecloudnet_b200.losses.tta_forward(group > 1).  Every op of the path is per sample, so this
    differs from the loop above only in the order the sampling draws are consumed."""
    import copy
    factors = [(fx * sc, fy * sc) for fx, fy in ((1.0, 1.0), (-1.0, 1.0), (1.0, -1.0), (-1.0, -1.0))
               for sc in (1.0, 0.95, 1.05)]
    n = data.pos.shape[0]
    nb = int(data.batch[-1]) + 1
    total = None
    for g0 in range(0, len(factors), group):
        fs = factors[g0:g0 + group]
        G = len(fs)
        st = copy.copy(data)
        parts = []
        for sx, sy in fs:
            p = data.pos.clone()
            p[:, 0] *= sx
            p[:, 1] *= sy
            parts.append(p)
        st.pos = torch.cat(parts)
        st.batch = torch.cat([data.batch + g * nb for g in range(G)])
        for k, v in data.__dict__.items():
            if k not in ("pos", "batch") and torch.is_tensor(v) and v.dim() >= 1 and v.shape[0] in (n, nb):
                setattr(st, k, v.repeat(G, *([1] * (v.dim() - 1))))
        out = model(st, labels=getattr(st, "y", None))
        for g in range(G):
            total = out[g * n:(g + 1) * n] if total is None else total + out[g * n:(g + 1) * n]
    out = F.log_softmax(total / len(factors), dim=-1)
    return out, torch.argmax(out, dim=-1)


# ----------------------------------------------------------------------------------------------------------------------
# object runs: src/run/shapenet_seg.py:88-111 (part IoU per object) - torchmetrics is absent here, its jaccard_index is
# restated from its published definition (confusion matrix; intersection / union per class, absent_score where the union is
# empty, mean over the classes): parity unpinned for that one function
# ----------------------------------------------------------------------------------------------------------------------
def jaccard_index(pred, target, num_classes, absent_score=0.0):
    conf = torch.zeros((num_classes, num_classes), dtype=torch.int64)
    conf.index_put_((target.long(), pred.long()), torch.ones_like(target, dtype=torch.int64), accumulate=True)
    inter = torch.diagonal(conf)
    union = conf.sum(0) + conf.sum(1) - inter
    scores = inter.float() / union.float()
    scores[union == 0] = absent_score
    return scores.mean()


def part_iou_loop(out, y, ptr, categories, seg_classes, names, n_seg_classes=50):
    """The object loop of shapenet_seg.py:88-99 -> list of per-object IoUs."""
    y_map = torch.zeros(n_seg_classes, dtype=torch.int64)
    sizes = (ptr[1:] - ptr[:-1]).tolist()
    ious = []
    for o, t, c in zip(out.split(sizes), y.split(sizes), categories):
        part = torch.tensor(seg_classes[names[int(c)]])
        y_map[part] = torch.arange(part.shape[0])
        ious.append(float(jaccard_index(o[:, part].argmax(dim=-1), y_map[t], num_classes=part.shape[0], absent_score=1.0)))
    return ious
