"""CPU restatement of the reference's scan ingest (TEST INFRASTRUCTURE: only tests/, smoke() and bench.py's
CPU legs may import this).

Follows SemNuScenes.get / _load_frame / _get_curves / _training_augmentation
(src/data/nuscenes_dataset.py:42-149) and SemKITTI.get / _load_frame / _get_curves / _training_augmentation
(src/data/kitti_dataset.py:40-152).  Pinned by tests/golden/ingest.pt, generated from those methods' own source
text by oracle/make_golden.py (mode `ingest`).

PolarMix: `polarmix` / `swap` / `rotate_copy` (data_utils.py:5-85) are restated below and pinned by golden vectors of
the real functions; the dataset methods that drive them are broken as shipped (the nuScenes `_apply_polarmix` calls
`self._load_nusc_data`, `SemKITTI.get` calls `self._polarmix`; neither exists), so the restatement follows the body of
`_apply_polarmix` with the partner sweep handed in.
"""
from __future__ import annotations

import numpy as np
import torch

from .ccn_point_ops import get_curves

POS_NORMALIZE = 20.0          # nuscenes_dataset.py:18, kitti_dataset.py:17
CURVE_THRESH = 0.08           # nuscenes_dataset.py:17, kitti_dataset.py:16


def draw_augmentation(dataset: str, rng=np.random):
    """The random draws of `_training_augmentation`, in the reference's call order
    (nuscenes_dataset.py:120-149: rotation over 360 deg; kitti_dataset.py:123-152: over 180 deg)."""
    if dataset == "nuscenes":
        rad = np.deg2rad(rng.random() * 360) - np.pi
    else:
        rad = np.deg2rad(rng.random() * 180) - np.pi / 2
    flip = int(rng.choice(4, 1)[0])
    scale = float(rng.uniform(0.95, 1.05))
    t = [float(rng.normal(0, 0.1, 1)[0]) for _ in range(3)]
    return {"rot_c": float(np.cos(rad)), "rot_s": float(np.sin(rad)), "flip": flip, "scale": scale, "translate": t}


def training_augmentation(xyz: np.ndarray, aug) -> np.ndarray:
    """nuscenes_dataset.py:120-149 with the draws given: f64 rotation rounded into the f32 array, sign flips, f32
    scale of x and y, f64 translation rounded once."""
    xyz = xyz.copy()
    c, s = aug["rot_c"], aug["rot_s"]
    x, y = xyz[:, 0].astype(np.float64), xyz[:, 1].astype(np.float64)
    xyz[:, 0] = x * c + y * (-s)
    xyz[:, 1] = x * s + y * c
    if aug["flip"] & 1:
        xyz[:, 0] = -xyz[:, 0]
    if aug["flip"] & 2:
        xyz[:, 1] = -xyz[:, 1]
    sc = np.float32(aug["scale"])
    xyz[:, 0] = sc * xyz[:, 0]
    xyz[:, 1] = sc * xyz[:, 1]
    xyz[:, :3] = xyz[:, :3].astype(np.float64) + np.asarray(aug["translate"], dtype=np.float64)[None]
    return xyz


def ingest_scan(raw: np.ndarray, labels, learning_map, dataset: str = "nuscenes", aug=None, pos_div=POS_NORMALIZE):
    """raw: (n,5) f32 nuScenes records or (n,4) f32 KITTI records; labels: u8 (nuScenes) / i32 (KITTI) / None (test
    split); learning_map: dict raw label -> class.  -> dict(pos, x, y, curve_idxs, reorder, inv_reorder)."""
    n = raw.shape[0]
    if dataset == "nuscenes":
        assert raw.shape[1] == 5
        beam = torch.from_numpy(raw[:, 4].copy())
        lab = np.ones(n, dtype=np.uint8) if labels is None else labels.reshape(-1)
    else:
        assert raw.shape[1] == 4
        beam = torch.zeros(n)
        lab = np.zeros(n, dtype=np.int64) if labels is None else (labels.reshape(-1).astype(np.int64) & 0xFFFF)
    if labels is not None:
        lab = np.array([learning_map[int(v)] for v in lab], dtype=np.int64)
    lab = lab.astype(np.uint8)
    pts = torch.from_numpy(raw[:, :3].copy())
    reorder, inv, cid = get_curves(pts, beam, CURVE_THRESH)
    xyz = pts[reorder].numpy()
    if aug is not None:
        xyz = training_augmentation(xyz, aug)
    pos = torch.from_numpy(xyz) / pos_div
    return {"pos": pos, "x": torch.from_numpy(raw[:, 3].copy())[reorder][:, None],
            "y": torch.from_numpy(lab)[reorder].long(), "curve_idxs": cid.long(), "reorder": reorder, "inv_reorder": inv}


# ------------------------------------------------------------------------------------------------------ PolarMix
def draw_polarmix(n_sweeps: int, rng=np.random):
    """Draw order of `_apply_polarmix` + `polarmix` (nuscenes_dataset.py:151-166, data_utils.py:69-85)."""
    idx2 = int(rng.randint(n_sweeps))
    alpha = (rng.random() - 1) * np.pi
    omega = [rng.random() * np.pi * 2 / 3, (rng.random() + 1) * np.pi * 2 / 3]
    return {"idx2": idx2, "alpha": alpha, "beta": alpha + np.pi, "omega": omega,
            "do_swap": bool(rng.random() < 0.5), "do_paste": bool(rng.random() < 0.5)}


def _sector(rows: np.ndarray, lo, hi) -> np.ndarray:
    yaw = -np.arctan2(rows[:, 1], rows[:, 0])                      # f32 in, f32 out (data_utils.py:7-9)
    return (yaw > np.float32(lo)) & (yaw < np.float32(hi))


def polarmix(d1, d2, mix, instance_classes):
    """data_utils.py:5-85 with the draws given, on two ingested sweeps (pos in sensor units, dense curve ids), plus
    the id bookkeeping of `_apply_polarmix` (nuscenes_dataset.py:151-178): sweep-2 ids += max(sweep-1 ids) before, dense
    ranks (`torch.unique(..., return_inverse=True)`) after.  Rows are [x y z reflectance curve_id] in f32."""
    def rows_of(d, off):
        cid = (d["curve_idxs"].numpy() + off).astype(np.float32)
        return np.concatenate([d["pos"].numpy(), d["x"].numpy().reshape(-1, 1), cid[:, None]], axis=1).astype(np.float32)

    r1 = rows_of(d1, 0)
    r2 = rows_of(d2, int(d1["curve_idxs"].max()))
    l1, l2 = d1["y"].numpy().astype(np.uint8), d2["y"].numpy().astype(np.uint8)
    rows, labs = r1, l1
    if mix["do_swap"]:                                              # swap(): sweep 1 outside the sector + sweep 2 inside
        m1, m2 = _sector(r1, mix["alpha"], mix["beta"]), _sector(r2, mix["alpha"], mix["beta"])
        rows = np.concatenate([r1[~m1], r2[m2]])
        labs = np.concatenate([l1[~m1], l2[m2]])
    if mix["do_paste"]:                                             # rotate_copy() of sweep 2's instance points
        sel = np.concatenate([np.nonzero(l2 == c)[0] for c in instance_classes])
        sel = sel[np.argsort(r2[sel, 4], kind="stable")]
        inst, ilab = r2[sel], l2[sel]
        copies, prev = [inst], inst
        for om in mix["omega"]:
            c, s = np.cos(om), np.sin(om)
            new = np.zeros_like(inst)
            x, y = inst[:, 0].astype(np.float64), inst[:, 1].astype(np.float64)
            new[:, 0] = x * c + y * (-s)
            new[:, 1] = x * s + y * c
            new[:, 2] = inst[:, 2]
            new[:, 3] = inst[:, 3]
            if len(inst):
                new[:, 4] = inst[:, 4] + np.max(prev[:, 4])
            copies.append(new)
            prev = new
        pasted = np.concatenate(copies)
        pasted[:, 4] += np.max(rows[:, 4])
        rows = np.concatenate([rows, pasted])
        labs = np.concatenate([labs, ilab, ilab, ilab])
    dense = torch.unique(torch.from_numpy(rows[:, 4].copy()), return_inverse=True)[1]
    return {"pos": torch.from_numpy(rows[:, :3].copy()), "x": torch.from_numpy(rows[:, 3:4].copy()),
            "y": torch.from_numpy(labs).long(), "curve_idxs": dense.long()}


# ----------------------------------------------------------------------------------------------------------------------
# A2D2 (src/data/audi_dataset.py): no beam ids - curves from a kNN graph per lidar.  Pinned by tests/golden/a2d2.pt,
# generated from SemA2D2.get / _get_curves / _partition_curves_at_discontinuities / _training_augmentation's own source text
# (oracle/make_golden.py, mode `a2d2`; torch_cluster.knn is absent there and stands as exact (d2, index) kNN: unpinned).
# ----------------------------------------------------------------------------------------------------------------------
A2D2_CURVE_THRESHES = (0.1, 0.17, 0.1, 0.12, 0.1)    # audi_dataset.py:19
A2D2_KNN = (4, 4, 4, 3, 4)                            # audi_dataset.py:20
A2D2_POS_NORMALIZE = 30.0                            # audi_dataset.py:21


def _components_by_lowest_point(n: int, src, dst) -> np.ndarray:
    """Weakly connected components, numbered in the order of their lowest point (what scipy's connected_components returns:
    it opens a new label at the first unlabelled node while walking 0..n-1).  Plain union-find."""
    parent = list(range(n))

    def find(a):
        while parent[a] != a:
            parent[a] = parent[parent[a]]
            a = parent[a]
        return a

    for a, b in zip(src, dst):
        ra, rb = find(a), find(b)
        if ra != rb:
            parent[max(ra, rb)] = min(ra, rb)
    roots = [find(a) for a in range(n)]
    label, out = {}, np.empty(n, dtype=np.int64)
    for a in range(n):                      # a root is the lowest point of its component, so labels appear in that order
        if roots[a] not in label:
            label[roots[a]] = len(label)
        out[a] = label[roots[a]]
    return out


def a2d2_partition(points: torch.Tensor, comp: torch.Tensor, thresh: float) -> torch.Tensor:
    """audi_dataset.py:214-237: a cut where the component changes, or where the f64 edges to the previous AND to the
    second-previous point (10 where there is none) both exceed thresh * sqrt(|p.xy|) (f32); ids = running count - 1."""
    n = points.shape[0]
    pd = points.double()
    e1 = torch.full((n,), 10.0, dtype=torch.float64)
    e2 = torch.full((n,), 10.0, dtype=torch.float64)
    e1[1:] = torch.linalg.norm(pd[1:] - pd[:-1], dim=-1)
    e2[2:] = torch.linalg.norm(pd[2:] - pd[:-2], dim=-1)
    lim = thresh * torch.sqrt(torch.linalg.norm(points[:, :2], dim=-1))
    cut = (e1 > lim) & (e2 > lim)
    cut[1:] |= comp[1:] != comp[:-1]
    return torch.cumsum(cut, dim=0) - 1


def a2d2_get_curves(points: np.ndarray, lidar_id: np.ndarray, timestamp: np.ndarray, knn=A2D2_KNN,
                    curve_threshes=A2D2_CURVE_THRESHES):
    """audi_dataset.py:156-211 -> (reorder (n) into the input, curve_idxs (n) int64, sorted points (n,3)); see the header of
    csrc/a2d2_curves.cu for the steps.  Quirks kept: a sensor's ids are offset by the sum of the LARGEST ids (not the
    counts) of the sensors before it (:197,205), so its first curve takes the id of the previous sensor's last; the final
    np.unique compaction."""
    from . import tp_ops as tp
    order_all, ids_all, total = [], [], 0
    for sid in np.unique(lidar_id.astype(np.int64)):
        sel = np.flatnonzero(lidar_id == sid)
        p = torch.from_numpy(points[sel])
        m = p.shape[0]
        k = min(int(knn[sid]), m)
        nbr = tp.knn_points(p[None], p[None], torch.tensor([m]), torch.tensor([m]), k)[1][0]            # (m, k), self included
        q = torch.arange(m)[:, None].expand(-1, k)
        elen = torch.norm(p[q.reshape(-1)] - p[nbr.reshape(-1)], dim=1)
        lim = curve_threshes[sid] * torch.sqrt(torch.linalg.norm(p[:, :2], dim=-1))
        keep = elen < lim[q.reshape(-1)]
        comp = _components_by_lowest_point(m, q.reshape(-1)[keep].tolist(), nbr.reshape(-1)[keep].tolist())
        order = np.lexsort((timestamp[sel], comp))                     # by component, then time, ties in file order
        ids = torch.from_numpy(comp[order])
        if m > 1:
            ids = a2d2_partition(p[order], ids, 0.9 * curve_threshes[sid])
        order_all.append(sel[order])
        ids_all.append(ids.numpy() + total)
        total += int(ids.max())
    reorder = np.concatenate(order_all)
    cid = np.unique(np.concatenate(ids_all), return_inverse=True)[1]
    return reorder, torch.from_numpy(cid.astype(np.int64)), points[reorder]


def a2d2_draw_augmentation(rng=np.random):
    """The draws of SemA2D2._training_augmentation in its call order (audi_dataset.py:81-90)."""
    return {"flip_y": int(rng.choice(2, 1)[0]), "scale": float(rng.uniform(0.95, 1.05))}


def a2d2_ingest_frame(frame: dict, aug=None):
    """SemA2D2.get (audi_dataset.py:120-141) after _load_frame: curves, positions / 30, reflectance / 255, the optional
    y flip and x, y scale -> dict(pos, x, reorder, curve_idxs)."""
    reorder, cid, pts = a2d2_get_curves(frame["points"], frame["lidar_id"], frame["timestamp"])
    pos = torch.from_numpy(pts.copy())
    pos /= A2D2_POS_NORMALIZE
    x = torch.from_numpy(frame["reflectance"][reorder]).view(-1, 1) / 255
    if aug is not None:
        xyz = pos.numpy()
        if aug["flip_y"] == 1:
            xyz[:, 1] = -xyz[:, 1]
        sc = np.float32(aug["scale"])
        xyz[:, 0] = sc * xyz[:, 0]
        xyz[:, 1] = sc * xyz[:, 1]
        pos = torch.from_numpy(xyz).float()
    return {"pos": pos.float(), "x": x, "reorder": torch.from_numpy(reorder), "curve_idxs": cid}
