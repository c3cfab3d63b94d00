"""Synthetic laser-scan inputs - datasets and checkpoints are offline.

This is the INPUT GENERATOR for tests and bench.py (host numpy, seeded), not part of
the hot path.  It emits exactly the `data` contract of SURVEY.md Appendix B:
pos f32 (N,3), x f32 (N,1) | None, batch i64 (N), curve_idxs i64 (N) (dense from 0 per
sample, non-decreasing), labels i64 (B) for object scans.

Driving scans mimic the reference's nuScenes/KITTI loaders
(src/data/nuscenes_dataset.py:42-66): a spinning multi-beam LiDAR ray-cast against a
ground plane, random boxes and an enclosing far wall; points are emitted in acquisition
(azimuth-major) order with a beam id per point, exactly what `_get_curves`
(nuscenes_dataset.py:91-118) consumes, and positions are divided by POS_NORMALIZE=20.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import List, Optional

import numpy as np
import torch

NUSCENES = dict(beams=32, azimuth_steps=1085, elev_deg=(-30.67, 10.67), pos_normalize=20.0,
                curve_thresh=0.08, refl_scale=1.0)
KITTI = dict(beams=64, azimuth_steps=1875, elev_deg=(-24.8, 2.0), pos_normalize=20.0,
             curve_thresh=0.08, refl_scale=1.0 / 255.0)


@dataclass
class RawScan:
    """One sweep in acquisition order, metres, before curve grouping."""
    points: np.ndarray       # (N,3) f32
    reflectance: np.ndarray  # (N,)  f32
    beam_ids: np.ndarray     # (N,)  f32 (the reference reads the beam id as the 5th float column)


def raycast_driving_scan(seed: int, beams=32, azimuth_steps=1085, elev_deg=(-30.67, 10.67),
                         n_boxes=40, drop=0.05, range_noise=0.02, max_range=70.0, refl_scale=1.0,
                         **_unused) -> RawScan:
    rng = np.random.default_rng(seed)
    elev = np.deg2rad(np.linspace(elev_deg[0], elev_deg[1], beams))
    az = np.linspace(-np.pi, np.pi, azimuth_steps, endpoint=False)
    # acquisition order: azimuth-major, beams interleaved inside one firing
    A, E = np.meshgrid(az, elev, indexing="ij")
    beam = np.broadcast_to(np.arange(beams)[None, :], A.shape)
    d = np.stack([np.cos(E) * np.cos(A), np.cos(E) * np.sin(A), np.sin(E)], axis=-1).reshape(-1, 3)
    beam = beam.reshape(-1)

    ground_z = -1.84
    t = np.full(d.shape[0], np.inf)
    dz = d[:, 2]
    hit = dz < -1e-6
    t[hit] = ground_z / dz[hit]

    # random axis-aligned boxes standing on the ground + one enclosing wall box seen from inside
    centers = rng.uniform(-50.0, 50.0, size=(n_boxes, 2))
    near = np.linalg.norm(centers, axis=1) < 4.0
    centers[near] += 8.0
    sizes = rng.uniform(1.0, 10.0, size=(n_boxes, 2))
    heights = rng.uniform(1.0, 10.0, size=(n_boxes,))
    lo = np.concatenate([centers - sizes / 2, np.full((n_boxes, 1), ground_z)], axis=1)
    hi = np.concatenate([centers + sizes / 2, (ground_z + heights)[:, None]], axis=1)
    with np.errstate(divide="ignore", invalid="ignore"):
        inv = 1.0 / d
        for b in range(n_boxes):
            t0 = lo[b][None, :] * inv
            t1 = hi[b][None, :] * inv
            tmin = np.nanmax(np.minimum(t0, t1), axis=1)
            tmax = np.nanmin(np.maximum(t0, t1), axis=1)
            ok = (tmax >= np.maximum(tmin, 0.0)) & (tmin > 0.0)
            t = np.where(ok & (tmin < t), tmin, t)
        # far wall: exit distance of the ray from a 110 m x 110 m x 40 m room
        wall_lo = np.array([-55.0, -55.0, ground_z])
        wall_hi = np.array([55.0, 55.0, 40.0])
        texit = np.nanmin(np.maximum(wall_lo[None, :] * inv, wall_hi[None, :] * inv), axis=1)
    t = np.minimum(t, texit)

    t = t + rng.normal(0.0, range_noise, size=t.shape)
    keep = np.isfinite(t) & (t > 1.0) & (t < max_range) & (rng.random(t.shape) > drop)
    pts = (d * t[:, None])[keep].astype(np.float32)
    refl = (rng.integers(0, 256, size=keep.sum()).astype(np.float32) * np.float32(refl_scale)).astype(np.float32)
    return RawScan(points=pts, reflectance=refl, beam_ids=beam[keep].astype(np.float32))


def a2d2_frame(seed: int, n_sensors: int = 5, beams: int = 16, azimuth_steps: int = 900, fov_deg: float = 60.0,
               shuffle: bool = True) -> dict:
    """An A2D2-shaped frame (src/data/audi_dataset.py:92-117 `_load_frame`: the npz keys points / reflectance / timestamp /
    lidar_id): `n_sensors` 16-beam lidars at different mounts, each cut to the front camera's field of view, NO beam ids,
    a timestamp per firing (all beams of a firing share it) and the points in shuffled file order.  Returns numpy arrays:
    points f32 (n,3), reflectance f32 (n,), timestamp i64 (n,), lidar_id i64 (n,), label i64 (n,)."""
    rng = np.random.default_rng(1000 + seed)
    pts, refl, ts, ids = [], [], [], []
    for sid in range(n_sensors):
        lo = float(rng.uniform(-22.0, -12.0))
        raw = raycast_driving_scan(100 * seed + sid, beams=beams, azimuth_steps=azimuth_steps, elev_deg=(lo, lo + 26.0),
                                   n_boxes=30, drop=0.08)
        p = raw.points
        az = np.arctan2(p[:, 1], p[:, 0])
        front = np.abs(az) < np.deg2rad(fov_deg)
        p, az = p[front], az[front]
        firing = np.round((az + np.pi) / (2 * np.pi) * azimuth_steps).astype(np.int64)
        mount = np.array([rng.uniform(-0.5, 1.5), rng.uniform(-0.8, 0.8), rng.uniform(-0.2, 0.3)], dtype=np.float32)
        pts.append(p + mount[None, :])
        refl.append(raw.reflectance[front])
        ts.append(1_554_000_000_000_000 + 37 * sid + 55 * firing)
        ids.append(np.full(p.shape[0], sid, dtype=np.int64))
    points = np.concatenate(pts).astype(np.float32)
    out = {"points": points, "reflectance": np.concatenate(refl).astype(np.float32),
           "timestamp": np.concatenate(ts).astype(np.int64), "lidar_id": np.concatenate(ids)}
    out["label"] = rng.integers(0, 38, size=points.shape[0]).astype(np.int64)
    if shuffle:
        perm = rng.permutation(points.shape[0])
        out = {k: np.ascontiguousarray(v[perm]) for k, v in out.items()}
    return out


def group_curves_host(points: torch.Tensor, beam_ids: torch.Tensor, curve_thresh: float = 0.08):
    """Host-side statement of the curve split used only by this generator: stable sort by
    beam, split where the beam changes or |p[i+1]-p[i]| (f64) > thresh*sqrt(|p[i+1].xy|).
    Returns (reorder, curve_idxs).  The device implementation of the same operator is
    ccn_get_curves; its parity test compares against the oracle, not against this."""
    reorder = torch.sort(beam_ids, stable=True)[1]
    p = points[reorder]
    b = beam_ids[reorder]
    split_beam = (b[1:] - b[:-1]) != 0
    en = torch.linalg.norm(p[1:].double() - p[:-1].double(), dim=-1)
    rad = torch.linalg.norm(p[1:, :2], dim=-1)
    split = (en > (curve_thresh * torch.sqrt(rad))) | split_beam
    cid = torch.cat([torch.zeros(1, dtype=torch.int64), torch.cumsum(split, dim=0)])
    return reorder, cid


class ScanBatch:
    """The `data` object handed to ModelBase.forward (stand-in for a PyG Batch)."""

    def __init__(self, x, pos, batch, curve_idxs, y=None, labels=None):
        self.x, self.pos, self.batch, self.curve_idxs, self.y = x, pos, batch, curve_idxs, y
        if labels is not None:
            self.labels = labels

    def to(self, device, non_blocking=False):
        kw = {}
        for k, v in self.__dict__.items():
            kw[k] = v.to(device, non_blocking=non_blocking) if torch.is_tensor(v) else v
        out = ScanBatch.__new__(ScanBatch)
        out.__dict__.update(kw)
        return out

    def pin_memory(self):
        out = ScanBatch.__new__(ScanBatch)
        out.__dict__.update({k: (v.pin_memory() if torch.is_tensor(v) else v) for k, v in self.__dict__.items()})
        return out

    @property
    def num_points(self):
        return int(self.pos.shape[0])

    def input_bytes(self):
        return sum(v.numel() * v.element_size() for v in self.__dict__.values() if torch.is_tensor(v))


def _collate(samples: List[dict], with_x: bool, labels: Optional[List[int]] = None) -> ScanBatch:
    pos = torch.cat([s["pos"] for s in samples])
    x = torch.cat([s["x"] for s in samples]) if with_x else None
    batch = torch.cat([torch.full((s["pos"].shape[0],), i, dtype=torch.int64) for i, s in enumerate(samples)])
    curve = torch.cat([s["curve_idxs"] for s in samples])
    y = torch.cat([s["y"] for s in samples])
    lab = torch.tensor(labels, dtype=torch.int64) if labels is not None else None
    return ScanBatch(x, pos, batch, curve, y, lab)


def driving_scan_sample(seed: int, kind: str = "nuscenes", n_classes: int = 17, subsample: Optional[int] = None,
                        **overrides) -> dict:
    cfg = dict(NUSCENES if kind == "nuscenes" else KITTI)
    cfg.update(overrides)
    raw = raycast_driving_scan(seed, **cfg)
    pts = torch.from_numpy(raw.points)
    beam = torch.from_numpy(raw.beam_ids)
    refl = torch.from_numpy(raw.reflectance)
    if subsample is not None and subsample < pts.shape[0]:
        # keep a contiguous azimuth sector so curves stay curves
        pts, beam, refl = pts[:subsample], beam[:subsample], refl[:subsample]
    reorder, cid = group_curves_host(pts, beam, cfg["curve_thresh"])
    pos = pts[reorder] / cfg["pos_normalize"]
    g = torch.Generator().manual_seed(seed + 7919)
    y = torch.randint(0, n_classes, (pos.shape[0],), generator=g)
    return dict(pos=pos.contiguous(), x=refl[reorder][:, None].contiguous(), curve_idxs=cid, y=y,
                raw_points=pts, raw_beam=beam, raw_reflectance=refl)


def driving_scan_batch(seeds, kind="nuscenes", n_classes=17, subsample=None, **overrides) -> ScanBatch:
    return _collate([driving_scan_sample(s, kind, n_classes, subsample, **overrides) for s in seeds], with_x=True)


def object_scan_sample(seed: int, n_points: int = 2048, pattern: str = "random") -> dict:
    """Heightfield object scanned along 24-40 laser lines (cf. scanning_simulator/utils/scanlines.py:31-38,
    curve split at gap>0.01 after unit-sphere normalisation, curve_cloud.py:40-47, object_datasets.py:61-66)."""
    rng = np.random.default_rng(seed)
    n_lines = max(2, int(rng.integers(24, 41)) * n_points // 2048)   # ~50-85 points per laser line
    mu = rng.uniform(-0.6, 0.6, size=(5, 2))
    sig = rng.uniform(0.15, 0.5, size=(5,))
    amp = rng.uniform(-0.6, 0.8, size=(5,))

    def height(xy):
        z = np.zeros(xy.shape[0])
        for k in range(5):
            z += amp[k] * np.exp(-((xy - mu[k]) ** 2).sum(1) / (2 * sig[k] ** 2))
        return z

    per = [n_points // n_lines + (1 if i < n_points % n_lines else 0) for i in range(n_lines)]
    pts, line_id = [], []
    for i, n in enumerate(per):
        if pattern == "parallel":
            th = 0.3
        elif pattern == "grid":
            th = 0.0 if i % 2 == 0 else math.pi / 2
        else:
            th = rng.uniform(0, math.pi)
        off = rng.uniform(-0.9, 0.9)
        dirv = np.array([math.cos(th), math.sin(th)])
        nrm = np.array([-dirv[1], dirv[0]])
        s = np.linspace(-1.0, 1.0, n)
        xy = off * nrm[None, :] + s[:, None] * dirv[None, :]
        # knock out a random stretch so lines break into several curves
        a = rng.integers(0, max(n - 8, 1))
        keep = np.ones(n, dtype=bool)
        keep[a:a + int(rng.integers(0, 6))] = False
        xy = xy[keep]
        pts.append(np.concatenate([xy, height(xy)[:, None]], axis=1))
        line_id.append(np.full(xy.shape[0], i))
    p = np.concatenate(pts).astype(np.float32)
    lid = np.concatenate(line_id)
    p = p - p.mean(0, keepdims=True)
    p = p / np.linalg.norm(p, axis=1).max()
    gap = np.linalg.norm(p[1:] - p[:-1], axis=1) > 0.04
    split = gap | (lid[1:] != lid[:-1])
    cid = np.concatenate([[0], np.cumsum(split)]).astype(np.int64)
    pos = torch.from_numpy(p.astype(np.float32))
    g = torch.Generator().manual_seed(seed + 104729)
    y = torch.randint(0, 50, (pos.shape[0],), generator=g)
    return dict(pos=pos, x=None, curve_idxs=torch.from_numpy(cid), y=y)


def object_scan_batch(seeds, n_points=2048, pattern="random") -> ScanBatch:
    samples = [object_scan_sample(s, n_points, pattern) for s in seeds]
    labels = [int(s % 16) for s in seeds]
    return _collate(samples, with_x=False, labels=labels)


def randomize_model_(model: torch.nn.Module, seed: int = 1234) -> torch.nn.Module:
    """Random-init weights of SURVEY §8(d): Linear/Conv weights ~ U(+-sqrt(3/fan_in)) (so signals
    neither vanish nor explode through the ~100 layers of the nuScenes step list), BatchNorm gamma~U(0.5,1.5), beta~N(0,0.1), running_mean~N(0,0.1), running_var~U(0.5,1.5)
    so that BN is not the identity.  Works on the oracle model and on the product model alike (same
    state_dict layout), iterating in state_dict order."""
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for name, t in model.state_dict().items():
            if name.endswith("num_batches_tracked"):
                continue
            if name.endswith("running_mean"):
                t.copy_(torch.randn(t.shape, generator=g) * 0.1)
            elif name.endswith("running_var"):
                t.copy_(torch.rand(t.shape, generator=g) + 0.5)
            elif ".norms." in name or "norm_modules" in name:
                if name.endswith("weight"):
                    t.copy_(torch.rand(t.shape, generator=g) + 0.5)
                else:
                    t.copy_(torch.randn(t.shape, generator=g) * 0.1)
            elif t.dim() >= 2:
                fan_in = t[0].numel()
                bound = math.sqrt(3.0 / fan_in)  # var 1/fan_in: measured to keep activations O(0.3..100) over all 33 steps
                t.copy_((torch.rand(t.shape, generator=g) * 2 - 1) * bound)
            else:  # bias
                t.copy_((torch.rand(t.shape, generator=g) * 2 - 1) * 0.1)
    return model


# ----------------------------------------------------------------------------------------------
# Synthetic triangle meshes for the scanning simulator (curvecloudnet_b200/simulator.py): the ShapeNet / Kortx meshes the
# reference's generator reads (scanning_simulator/shapenet_seg/generate_shapenet.py) are not available offline.  Objects fit
# a sphere of radius ~0.3 around the origin (the camera sits at distance 1, sampling.py:16,37), every triangle is present in
# both windings (the rasteriser culls one winding, rasterizer_kernel.cu:194-198), so every view sees a closed surface.
# ----------------------------------------------------------------------------------------------
def _grid_faces(nu: int, nv: int, wrap_u: bool, wrap_v: bool):
    import numpy as np
    iu = np.arange(nu if wrap_u else nu - 1)
    iv = np.arange(nv if wrap_v else nv - 1)
    a, b = np.meshgrid(iu, iv, indexing="ij")
    a1, b1 = (a + 1) % nu, (b + 1) % nv
    v00, v10, v01, v11 = a * nv + b, a1 * nv + b, a * nv + b1, a1 * nv + b1
    return np.concatenate([np.stack([v00, v10, v11], -1).reshape(-1, 3), np.stack([v00, v11, v01], -1).reshape(-1, 3)], 0)


def _box(lo, hi):
    import numpy as np
    lo, hi = np.asarray(lo, float), np.asarray(hi, float)
    c = np.array([[x, y, z] for x in (lo[0], hi[0]) for y in (lo[1], hi[1]) for z in (lo[2], hi[2])])
    quads = [(0, 1, 3, 2), (4, 6, 7, 5), (0, 4, 5, 1), (2, 3, 7, 6), (0, 2, 6, 4), (1, 5, 7, 3)]
    f = np.array([[q[0], q[1], q[2]] for q in quads] + [[q[0], q[2], q[3]] for q in quads])
    return c, f


def synthetic_mesh(kind: str = "blob", seed: int = 0, detail: int = 64):
    """-> vertices (nv, 3) float32, faces (nf, 3) int32 (CPU tensors).  kind: "sphere" | "blob" (sphere with seeded radial
    harmonics: self-occlusion, silhouettes) | "torus" | "table" (few LARGE triangles: a top and four legs)."""
    import numpy as np
    rng = np.random.RandomState(seed)
    if kind in ("sphere", "blob"):
        nu, nv = detail, 2 * detail
        th = np.linspace(0.02, np.pi - 0.02, nu)[:, None]
        ph = (np.arange(nv) * (2 * np.pi / nv))[None, :]
        r = 0.28 * np.ones((nu, nv))
        if kind == "blob":
            for _ in range(4):
                a, b, p1, p2 = rng.randint(1, 5), rng.randint(1, 5), rng.rand() * 6.28, rng.rand() * 6.28
                r = r + 0.03 * np.sin(a * th + p1) * np.cos(b * ph + p2)
        v = np.stack([r * np.sin(th) * np.cos(ph), r * np.cos(th) + 0 * ph, r * np.sin(th) * np.sin(ph)], -1).reshape(-1, 3)
        f = _grid_faces(nu, nv, False, True)
    elif kind == "torus":
        nu, nv = 2 * detail, detail
        u = (np.arange(nu) * (2 * np.pi / nu))[:, None]
        w = (np.arange(nv) * (2 * np.pi / nv))[None, :]
        R0, r0 = 0.2, 0.08
        v = np.stack([(R0 + r0 * np.cos(w)) * np.cos(u), r0 * np.sin(w) + 0 * u, (R0 + r0 * np.cos(w)) * np.sin(u)], -1).reshape(-1, 3)
        f = _grid_faces(nu, nv, True, True)
    elif kind == "table":
        parts = [_box([-0.28, 0.10, -0.2], [0.28, 0.13, 0.2])]
        for sx in (-1, 1):
            for sz in (-1, 1):
                parts.append(_box([sx * 0.25 - 0.02, -0.2, sz * 0.17 - 0.02], [sx * 0.25 + 0.02, 0.10, sz * 0.17 + 0.02]))
        vs, fs, off = [], [], 0
        for c, ff in parts:
            vs.append(c)
            fs.append(ff + off)
            off += c.shape[0]
        v, f = np.concatenate(vs, 0), np.concatenate(fs, 0)
    else:
        raise KeyError(kind)
    # a seeded rigid tilt, so that no face is axis-aligned with the image grid
    ax = rng.randn(3)
    ax /= np.linalg.norm(ax)
    ang = 0.5 * rng.rand() + 0.2
    K = np.array([[0, -ax[2], ax[1]], [ax[2], 0, -ax[0]], [-ax[1], ax[0], 0]])
    Rm = np.eye(3) + np.sin(ang) * K + (1 - np.cos(ang)) * K @ K
    v = v @ Rm.T
    f = np.concatenate([f, f[:, ::-1]], 0)
    return torch.tensor(v, dtype=torch.float32), torch.tensor(np.ascontiguousarray(f), dtype=torch.int32)
