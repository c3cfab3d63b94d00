"""Scan ingest on the device: raw sweep records -> the `Data` the model consumes.

Host-side mirror of the reference's dataset `get()` bodies - SemNuScenes (src/data/nuscenes_dataset.py:42-149) and
SemKITTI (src/data/kitti_dataset.py:40-152): the stable sort by ring id, the curve split, the label map, the training
augmentation and the position normalisation run in ONE C-ABI call (`ccn_ingest_scan`, csrc/scan_ingest.cu) on the raw
file bytes; the host only draws the augmentation's random numbers (same numpy calls, same order as the reference, so a
seeded run reproduces the reference's augmented sweep) and moves the bytes.

PolarMix (`polarmix` / `swap` / `rotate_copy`, data_utils.py:5-85, driven like `_apply_polarmix` drives them) runs on the
device as well (`ccn_polarmix`, csrc/polarmix.cu) between the grouping and the augmentation: `ingest_scan_polarmix`.  (The
dataset methods that call it are broken as shipped - `_load_nusc_data` / `_polarmix` do not exist - so the partner sweep is
handed in by the caller.)  A2D2's five-sensor kNN curve builder (audi_dataset.py:140-210) is a different algorithm and not
part of this module.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence

import numpy as np
import torch

from . import native
from . import ops
from .model import Batch

POS_NORMALIZE = 20.0          # nuscenes_dataset.py:18, kitti_dataset.py:17
CURVE_THRESH = 0.08           # nuscenes_dataset.py:17, kitti_dataset.py:16

# polarmix instance classes (nuscenes_dataset.py:12-13, kitti_dataset.py:10-12)
INSTANCE_CLASSES = {"nuscenes": (2, 3, 4, 5, 6, 7, 9, 10), "kitti": (1, 2, 3, 4, 5, 6, 7, 8)}

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
                check: bool = True, pos_div: float = POS_NORMALIZE) -> Batch:
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
                    float(pos_div), scratch, beam_tmp, sorted_tmp, reorder, inv, pos, x, y, cid, counts)
        if check:
            flags = int(counts[1].item())                    # host sync
            if flags & ERR_RING:
                raise ValueError("ingest_scan: ring id outside [0, %d)" % nb)
            if flags & ERR_LABEL:
                raise KeyError("ingest_scan: label without an entry in the learning map")
    return Batch(x=x, pos=pos, y=y, curve_idxs=cid, reorder=inv, counts=counts)


@dataclass
class PolarMixDraw:
    """The drawn values of `_apply_polarmix` + `polarmix` (nuscenes_dataset.py:151-166, data_utils.py:69-85)."""
    idx2: int
    alpha: float
    beta: float
    omega0: float
    omega1: float
    do_swap: bool
    do_paste: bool

    @classmethod
    def draw(cls, n_sweeps: int, rng=np.random) -> "PolarMixDraw":
        """Same numpy calls in the same order as the reference (partner sweep, sector, rotation angles, two gates)."""
        idx2 = int(rng.randint(n_sweeps))
        alpha = (rng.random() - 1) * np.pi
        beta = alpha + np.pi
        omega0 = rng.random() * np.pi * 2 / 3
        omega1 = (rng.random() + 1) * np.pi * 2 / 3
        do_swap = bool(rng.random() < 0.5)
        do_paste = bool(rng.random() < 0.5)
        return cls(idx2, float(alpha), float(beta), float(omega0), float(omega1), do_swap, do_paste)


def polarmix(d1: Batch, d2: Batch, mix: PolarMixDraw, instance_classes: Sequence[int]) -> Batch:
    """`polarmix` (data_utils.py:69-85) + the curve-id bookkeeping of `_apply_polarmix` on two ingested sweeps whose
    positions are still in sensor units (`ingest_scan(..., pos_div=1.0)`, no augmentation).  One host sync (the size of
    the result)."""
    import ctypes
    n1, n2 = d1.pos.shape[0], d2.pos.shape[0]
    if n1 == 0 or n2 == 0:
        raise ValueError("polarmix needs two non-empty sweeps")       # np.max of an empty array raises upstream too
    dev = d1.pos.device
    n_max = n1 + 4 * n2
    pos = torch.empty((n_max, 3), dtype=torch.float32, device=dev)
    x = torch.empty((n_max, 1), dtype=torch.float32, device=dev)
    y = torch.empty((n_max,), dtype=torch.int64, device=dev)
    cid = torch.empty((n_max,), dtype=torch.int64, device=dev)
    counts = torch.zeros((2,), dtype=torch.int32, device=dev)
    scratch = torch.empty((2 * n1 + 7 * n2 + 8,), dtype=torch.int32, device=dev)
    cls = (ctypes.c_int32 * max(len(instance_classes), 1))(*[int(c) for c in instance_classes])
    native.call("ccn_polarmix", d1.pos, d1.x, d1.y, d1.curve_idxs, n1, d2.pos, d2.x, d2.y, d2.curve_idxs, n2, d2.counts,
                int(mix.do_swap), float(np.float32(mix.alpha)), float(np.float32(mix.beta)), int(mix.do_paste),
                ctypes.addressof(cls), len(instance_classes), float(mix.omega0), float(mix.omega1), scratch, pos, x, y, cid,
                counts)
    n_out, n_curves = counts.tolist()                                   # host sync
    return Batch(x=x[:n_out], pos=pos[:n_out], y=y[:n_out], curve_idxs=cid[:n_out], counts=counts)


def augment_points(pos: torch.Tensor, augmentation: Optional[ScanAugmentation], pos_div: float = POS_NORMALIZE):
    """`_training_augmentation` + `points /= POS_NORMALIZE` on grouped points (the tail of `get()` after PolarMix)."""
    out = torch.empty_like(pos)
    a = augmentation
    t = a.translate if a is not None else (0.0, 0.0, 0.0)
    native.call("ccn_augment_points", pos, pos.shape[0], 1 if a is not None else 0, float(a.rot_c) if a else 1.0,
                float(a.rot_s) if a else 0.0, int(a.flip) if a else 0, float(a.scale) if a else 1.0, float(t[0]), float(t[1]),
                float(t[2]), float(pos_div), out)
    return out


def ingest_scan_polarmix(sweep1: dict, sweep2: dict, mix: PolarMixDraw, learning_map: LearningMap,
                         dataset: str = "nuscenes", augmentation: Optional[ScanAugmentation] = None, n_beams: int = 32,
                         device="cuda") -> Batch:
    """`get()` of the training split with `polarmix_aug` (nuscenes_dataset.py:42-65): ingest both sweeps, mix them, then
    augment and normalise.  sweep = {"raw": ..., "labels": ...}; `mix.idx2` selected sweep2 (the caller owns the files)."""
    d1 = ingest_scan(sweep1["raw"], sweep1.get("labels"), learning_map, dataset, None, n_beams, device, False, 1.0)
    d2 = ingest_scan(sweep2["raw"], sweep2.get("labels"), learning_map, dataset, None, n_beams, device, False, 1.0)
    out = polarmix(d1, d2, mix, INSTANCE_CLASSES[dataset])
    flags = int(torch.maximum(d1.counts[1], d2.counts[1]).item())
    if flags:
        raise ValueError("ingest_scan_polarmix: ring id out of range or unmapped label (flags %d)" % flags)
    out.pos = augment_points(out.pos, augmentation)
    return out


@dataclass
class A2D2Augmentation:
    """The draws of SemA2D2._training_augmentation (audi_dataset.py:81-90): flip of the y axis, one scale for x and y."""
    flip_y: int
    scale: float

    @classmethod
    def draw(cls, rng=np.random) -> "A2D2Augmentation":
        flip = int(rng.choice(2, 1)[0])
        return cls(flip, float(rng.uniform(0.95, 1.05)))


A2D2_POS_NORMALIZE = 30.0        # audi_dataset.py:21


def ingest_a2d2_frame(frame: dict, labels=None, augmentation: Optional[A2D2Augmentation] = None, device="cuda",
                      check: bool = True) -> Batch:
    """SemA2D2.get (audi_dataset.py:120-141) after `_load_frame`: frame = the npz fields {"points" (n,3) f32, "reflectance"
    (n,) f32, "timestamp" (n,) i64, "lidar_id" (n,) i64} as numpy arrays or tensors; labels = per-point class ids (the
    reference reads them off the undistorted label image on the host, :92-117 - file and image decoding stay with the
    caller).  Curves come from ccn_a2d2_curves (kNN graph per lidar -> components -> time order -> cuts), positions are
    divided by 30 and, for training, y-flipped and x, y-scaled - all on the device.  -> Batch(x, pos, y, curve_idxs, reorder
    = index of every output point in the input, counts = {curves, error flags})."""
    native.load()
    pts = np.ascontiguousarray(frame["points"], dtype=np.float32) if isinstance(frame["points"], np.ndarray) else frame["points"]
    n = int(pts.shape[0])
    ptd = _as_device(pts, torch.float32, device).view(n, 3)
    lid = _as_device(_np_i64(frame["lidar_id"]), torch.int64, device)
    tsd = _as_device(_np_i64(frame["timestamp"]), torch.int64, device)
    refd = _as_device(frame["reflectance"] if torch.is_tensor(frame["reflectance"])
                      else np.ascontiguousarray(frame["reflectance"], dtype=np.float32), torch.float32, device)
    labd = _as_device(_np_i64(labels), torch.int64, device) if labels is not None else None
    reorder, cid, sorted_pts, counts = ops.a2d2_curves(ptd, lid, tsd)
    x = torch.empty((n, 1), dtype=torch.float32, device=ptd.device)
    y = torch.empty((n,), dtype=torch.int64, device=ptd.device)
    native.call("ccn_a2d2_finish", refd, labd if labd is not None else 0, reorder, n, 255.0, x, y)
    if n == 0:
        return Batch(x=x, pos=sorted_pts, y=y, curve_idxs=cid, reorder=reorder, counts=counts)
    pos = augment_points(sorted_pts, None, A2D2_POS_NORMALIZE)                      # points /= 30 (:137)
    if augmentation is not None:                                                    # then flip / scale (:140-141)
        a = ScanAugmentation(1.0, 0.0, 2 if augmentation.flip_y == 1 else 0, augmentation.scale, (0.0, 0.0, 0.0))
        pos = augment_points(pos, a, 1.0)
    if check:
        flags = int(counts[1].item())                        # host sync
        if flags & 1:
            raise ValueError("ingest_a2d2_frame: lidar id outside [0, %d)" % len(ops.A2D2_KNN))
        if flags & 2:
            raise ValueError("ingest_a2d2_frame: timestamps span more than 2^30 ticks")
    return Batch(x=x, pos=pos, y=y, curve_idxs=cid, reorder=reorder, counts=counts)


@dataclass
class ObjectAugmentation:
    """The draws of the shapenet-seg training transform (object_datasets.py:31-36), one object: RandomScale's
    `random.uniform(0.95, 1.05)` (python's generator) and `(torch.rand((1, 3)) - 0.5) * 0.05` (torch's)."""
    scale: float
    shift: Sequence[float]

    @classmethod
    def draw(cls) -> "ObjectAugmentation":
        import random
        scale = random.uniform(0.95, 1.05)
        shift = ((torch.rand((1, 3)) - 0.5) * 0.05)[0]
        return cls(float(scale), tuple(float(v) for v in shift))


def ingest_object_scans(pos: torch.Tensor, batch: torch.Tensor, augmentations: Optional[List[ObjectAugmentation]] = None):
    """CurvesInMemoryDataset.get (object_datasets.py:21-38) for a batch of stored objects: positions (N, 3) with their object
    ids -> normalised (and, for the shapenet-seg training split, transformed) positions.  Loading the `.pth` file is the
    caller's; curve ids, normals and labels pass through unchanged."""
    native.load()
    pos = pos.to(torch.float32).contiguous()
    sptr = ops.sample_ptr(batch)
    B = sptr.n_samples
    out = torch.empty_like(pos)
    aug = 0
    if augmentations is not None:
        if len(augmentations) != B:
            raise ValueError("one ObjectAugmentation per object")
        aug = torch.tensor([[a.scale, a.shift[0], a.shift[1], a.shift[2]] for a in augmentations], dtype=torch.float32).to(pos.device)
    native.call("ccn_object_normalize", pos, sptr.dev, B, aug, out)
    return out


def ingest_kortx_scan(points, raw_curve_idxs, segmentation=None, npoints: int = 2048, samples_per_scan: int = 5,
                      capture_setting: str = "table", device="cuda"):
    """The per-file body of scanning_simulator/kortx/generate_kortx.py:38-75 on the device: one long scan (the CSV's columns
    0-2 xyz in f64, 4 raw curve counter, 5 segmentation) -> `fix_curve_idxs`, `cut_curves` (threshold 0.2 for 'table'
    captures, else 1.0), `samples_per_scan` windows of `npoints` points scaled into the unit cube's half, curve ids renumbered
    per window.  -> dict(pos (S, npoints, 3) f64, curve_idxs (S, npoints) i64, y (S, npoints) i64 or None, all_curve_idxs (n,)).
    Reading the CSV / info.json stays with the caller."""
    native.load()
    pts = _as_device(np.ascontiguousarray(points, dtype=np.float64) if isinstance(points, np.ndarray) else points.to(torch.float64),
                     torch.float64, device).view(-1, 3)
    n = pts.shape[0]
    if n < npoints:
        raise ValueError("ingest_kortx_scan: %d points, fewer than one window of %d" % (n, npoints))
    raw = _as_device(_np_i64(raw_curve_idxs), torch.int64, device)
    seg = _as_device(_np_i64(segmentation), torch.int64, device) if segmentation is not None else None
    dev = pts.device
    scratch = torch.empty((2 * n,), dtype=torch.int32, device=dev)
    curve = torch.empty((n,), dtype=torch.int64, device=dev)
    native.call("ccn_kortx_curves", pts, raw, n, 0.2 if capture_setting == "table" else 1.0, scratch, curve)
    S = int(samples_per_scan)
    pos = torch.empty((S, npoints, 3), dtype=torch.float64, device=dev)
    cid = torch.empty((S, npoints), dtype=torch.int64, device=dev)
    y = torch.empty((S, npoints), dtype=torch.int64, device=dev) if seg is not None else None
    native.call("ccn_kortx_windows", pts, curve, seg if seg is not None else 0, n, int(npoints), S, pos, cid, y if y is not None else 0)
    return {"pos": pos, "curve_idxs": cid, "y": y, "all_curve_idxs": curve}


def _np_i64(a):
    return np.ascontiguousarray(a, dtype=np.int64) if isinstance(a, np.ndarray) else a


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
