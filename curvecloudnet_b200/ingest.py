"""Scan ingest on the device: raw sweep records -> the `Data` the model consumes.

Host-side mirror of the reference's dataset `get()` bodies - SemNuScenes (src/data/nuscenes_dataset.py:42-149) and
SemKITTI (src/data/kitti_dataset.py:40-152): the stable sort by ring id, the curve split, the label map, the training
augmentation and the position normalisation run in ONE C-ABI call (`ccn_ingest_scan`, csrc/scan_ingest.cu) on the raw
file bytes; the host only draws the augmentation's random numbers (same numpy calls, same order as the reference, so a
seeded run reproduces the reference's augmented sweep) and moves the bytes.

PolarMix (`_apply_polarmix`, data_utils.py:5-85) is unreachable in the reference as shipped (it calls methods that do not
exist: `_load_nusc_data`, `_polarmix`), and A2D2's five-sensor curve builder (audi_dataset.py:140-210) is a different
algorithm; neither is part of this entry point.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence

import numpy as np
import torch

from . import native
from .model import Batch

POS_NORMALIZE = 20.0          # nuscenes_dataset.py:18, kitti_dataset.py:17
CURVE_THRESH = 0.08           # nuscenes_dataset.py:17, kitti_dataset.py:16

ERR_RING = 1
ERR_LABEL = 2


@dataclass
class ScanAugmentation:
    """The drawn values of `_training_augmentation` (nuscenes_dataset.py:120-149, kitti_dataset.py:123-152)."""
    rot_c: float = 1.0
    rot_s: float = 0.0
    flip: int = 0                                   # np.random.choice(4): 0 none, 1 x, 2 y, 3 x and y
    scale: float = 1.0
    translate: Sequence[float] = field(default_factory=lambda: (0.0, 0.0, 0.0))

    @classmethod
    def draw(cls, dataset: str = "nuscenes", rng=np.random) -> "ScanAugmentation":
        """Same numpy calls in the same order as the reference: with the same `np.random.seed` the draws are equal."""
        if dataset == "nuscenes":
            rad = np.deg2rad(rng.random() * 360) - np.pi
        elif dataset == "kitti":
            rad = np.deg2rad(rng.random() * 180) - np.pi / 2
        else:
            raise ValueError("dataset must be 'nuscenes' or 'kitti'")
        flip = int(rng.choice(4, 1)[0])
        scale = float(rng.uniform(0.95, 1.05))
        t = tuple(float(rng.normal(0, 0.1, 1)[0]) for _ in range(3))
        return cls(float(np.cos(rad)), float(np.sin(rad)), flip, scale, t)


class LearningMap:
    """`learning_map` of the dataset YAML (raw label -> class) as a device lookup table; absent keys are -1 and make
    the ingest raise, like the reference's `learning_map.__getitem__` KeyError."""

    def __init__(self, mapping: Dict[int, int], device="cuda"):
        size = max(mapping) + 1 if mapping else 1
        lut = np.full(size, -1, dtype=np.int32)
        for k, v in mapping.items():
            lut[int(k)] = int(v)
        self.size = size
        self.lut = torch.from_numpy(lut).to(device)


class _Staging:
    """Ring of pinned host buffers for the file bytes: the H2D copy of sweep i overlaps the host work of sweep i+1; a
    buffer is reused only after the copy that read it has completed (its event)."""

    def __init__(self, slots: int = 4):
        self.bufs = [None] * slots
        self.events = [None] * slots
        self.next = 0

    def upload(self, src: np.ndarray, device) -> torch.Tensor:
        i = self.next
        self.next = (i + 1) % len(self.bufs)
        if self.events[i] is not None:
            self.events[i].synchronize()
        nbytes = src.nbytes
        if self.bufs[i] is None or self.bufs[i].numel() < nbytes:
            self.bufs[i] = torch.empty((max(nbytes, 1 << 20),), dtype=torch.uint8, pin_memory=True)
        host = self.bufs[i][:nbytes]
        host.numpy()[:] = src.reshape(-1).view(np.uint8)
        out = host.to(device, non_blocking=True)
        ev = torch.cuda.Event()
        ev.record()
        self.events[i] = ev
        return out


_staging = _Staging()


def _as_device(buf, dtype, device):
    """bytes / numpy array (staged through pinned memory) or torch tensor (host or device) -> flat device tensor."""
    if isinstance(buf, (bytes, bytearray, memoryview)):
        buf = np.frombuffer(buf, dtype=np.uint8)
    if isinstance(buf, np.ndarray):
        return _staging.upload(np.ascontiguousarray(buf), device).view(dtype)
    if not torch.is_tensor(buf):
        raise TypeError("expected bytes, a numpy array or a torch tensor")
    buf = buf.contiguous().reshape(-1)
    if buf.dtype != dtype:
        buf = buf.view(torch.uint8).view(dtype)             # reinterpret the file bytes
    return buf.to(device, non_blocking=True)


def ingest_scan(raw, labels=None, learning_map: Optional[LearningMap] = None, dataset: str = "nuscenes",
                augmentation: Optional[ScanAugmentation] = None, n_beams: int = 32, device="cuda",
                check: bool = True) -> Batch:
    """One sweep -> `Data` fields (a one-sample attribute bag; `ingest_batch` concatenates).

    raw     the `.bin` file content: (n,5) f32 x y z intensity ring for nuScenes, (n,4) f32 x y z remission for KITTI
    labels  the lidarseg / .label file content: u8 per point (nuScenes), i32 per point (KITTI, low 16 bits used), or None
            (test split: ones for nuScenes, zeros for KITTI, like the reference)
    augmentation  None (validation / test) or the drawn training augmentation
    check   read the error flags back (one host sync) and raise on a ring id outside [0, n_beams) / an unmapped label
    """
    if dataset not in ("nuscenes", "kitti"):
        raise ValueError("dataset must be 'nuscenes' or 'kitti'")
    native.load()
    rec = 5 if dataset == "nuscenes" else 4
    nbytes = raw.nbytes if isinstance(raw, np.ndarray) else (raw.numel() * raw.element_size() if torch.is_tensor(raw) else len(raw))
    if nbytes % (4 * rec):
        raise ValueError("raw sweep: %d bytes is not a multiple of the %d-float record" % (nbytes, rec))
    rawd = _as_device(raw, torch.float32, device)
    n = rawd.numel() // rec
    nb = int(n_beams) if dataset == "nuscenes" else 1
    label_bytes, labd, lut, lut_size = 0, 0, 0, (-1 if dataset == "nuscenes" else 0)
    if labels is not None:
        if learning_map is None:
            raise ValueError("labels need a LearningMap")
        label_bytes = 1 if dataset == "nuscenes" else 4
        labd = _as_device(labels, torch.uint8 if label_bytes == 1 else torch.int32, device)
        if labd.numel() != n:
            raise ValueError("labels: %d entries for %d points" % (labd.numel(), n))
        lut, lut_size = learning_map.lut, learning_map.size
    kw = dict(device=rawd.device)
    pos = torch.empty((n, 3), dtype=torch.float32, **kw)
    x = torch.empty((n, 1), dtype=torch.float32, **kw)
    y = torch.empty((n,), dtype=torch.int64, **kw)
    cid = torch.empty((n,), dtype=torch.int64, **kw)
    reorder = torch.empty((n,), dtype=torch.int64, **kw)
    inv = torch.empty((n,), dtype=torch.int64, **kw)
    counts = torch.zeros((2,), dtype=torch.int32, **kw)
    if n > 0:
        scratch = torch.empty((nb + n,), dtype=torch.int32, **kw)
        beam_tmp = torch.empty((n,), dtype=torch.int64, **kw)
        sorted_tmp = torch.empty((n, 3), dtype=torch.float32, **kw)
        a = augmentation
        t = a.translate if a is not None else (0.0, 0.0, 0.0)
        native.call("ccn_ingest_scan", rawd, n, rec, labd, label_bytes, 0xFFFF if label_bytes == 4 else 0xFF, lut, lut_size,
                    nb, CURVE_THRESH, 1 if a is not None else 0, float(a.rot_c) if a else 1.0, float(a.rot_s) if a else 0.0,
                    int(a.flip) if a else 0, float(a.scale) if a else 1.0, float(t[0]), float(t[1]), float(t[2]),
                    POS_NORMALIZE, scratch, beam_tmp, sorted_tmp, reorder, inv, pos, x, y, cid, counts)
        if check:
            flags = int(counts[1].item())                    # host sync
            if flags & ERR_RING:
                raise ValueError("ingest_scan: ring id outside [0, %d)" % nb)
            if flags & ERR_LABEL:
                raise KeyError("ingest_scan: label without an entry in the learning map")
    return Batch(x=x, pos=pos, y=y, curve_idxs=cid, reorder=inv, counts=counts)


def ingest_batch(sweeps: List[dict], learning_map: Optional[LearningMap] = None, dataset: str = "nuscenes",
                 augment: bool = False, rng=np.random, n_beams: int = 32, device="cuda") -> Batch:
    """`DataLoader` collate of the reference (torch_geometric Batch.from_data_list) over device-ingested sweeps.
    sweeps: [{"raw": ..., "labels": ... or None}]; augment=True draws one training augmentation per sweep."""
    out = []
    for s in sweeps:
        aug = ScanAugmentation.draw(dataset, rng) if augment else None
        out.append(ingest_scan(s["raw"], s.get("labels"), learning_map, dataset, aug, n_beams, device, check=False))
    flags = torch.stack([d.counts[1] for d in out]).max() if out else None
    if flags is not None and int(flags.item()):              # ONE host sync per batch
        raise ValueError("ingest_batch: ring id out of range or unmapped label (flags %d)" % int(flags.item()))
    return Batch.from_data_list(out)
