"""CPU restatement of the reference's scan ingest (TEST INFRASTRUCTURE: only tests/, smoke() and bench.py's
CPU legs may import this).

Follows SemNuScenes.get / _load_frame / _get_curves / _training_augmentation
(src/data/nuscenes_dataset.py:42-149) and SemKITTI.get / _load_frame / _get_curves / _training_augmentation
(src/data/kitti_dataset.py:40-152).  Pinned by tests/golden/ingest.pt, generated from those methods' own source
text by oracle/make_golden.py (mode `ingest`).

PolarMix (`_apply_polarmix`, data_utils.py:5-85) is unreachable in the reference as shipped: the nuScenes method
calls `self._load_nusc_data` and SemKITTI.get calls `self._polarmix`, neither of which exists - it is not restated.
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


def ingest_scan(raw: np.ndarray, labels, learning_map, dataset: str = "nuscenes", aug=None):
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
    pos = torch.from_numpy(xyz) / POS_NORMALIZE
    return {"pos": pos, "x": torch.from_numpy(raw[:, 3].copy())[reorder][:, None],
            "y": torch.from_numpy(lab)[reorder].long(), "curve_idxs": cid.long(), "reorder": reorder, "inv_reorder": inv}
