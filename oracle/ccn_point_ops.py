"""CPU restatement of /root/reference/src/models/utils/point_ops.py and
src/models/modules/fps_ops.py (the op-level boundary, SURVEY.md §8(a) a2,a6,a7,a10-a15,a19,a20).

TEST INFRASTRUCTURE - see oracle/__init__.py.  Every function cites the reference lines it
follows.  Random draws the reference takes from torch's global RNG are explicit arguments
(see `Draws`).  Float arithmetic that decides an INDEX is written as explicit single
operations that reproduce what the reference's torch calls compute on CPU:

  torch.linalg.norm(v, dim=-1), v (N,3) f32  ==  sqrt(fma(z,z,fma(y,y,x*x)))   (norm3)
  torch.cumsum(f32)                           ==  sequential sum in f64, rounded per element
  scalar / tensor (Tensor.__rtruediv__)       ==  reciprocal(tensor) * scalar
  tensor % scalar                             ==  fmod with the sign fix (exact)

(all four are pinned against torch itself in tests/test_oracle_numerics.py).
"""
from __future__ import annotations

import math

import torch

from . import tp_ops as tp


# ----------------------------------------------------------------------------------
# random draws
# ----------------------------------------------------------------------------------
class Draws:
    """The reference's random draws, in call order, from torch's GLOBAL CPU generator:
    CurveFPS `torch.rand(1)` (fps_ops.py:30), VoxelFPS `torch.rand(N)` (fps_ops.py:56), the
    pytorch3d random FPS start `torch.randint(high=len, size=(1,))` per cloud
    (point_ops.py:64) and SAModule's `torch.randperm` (pointnet2.py:51).  Seeding the global
    generator identically therefore gives the reference, the oracle and the CUDA path the
    same draws."""

    def curve_phase(self) -> torch.Tensor:
        return torch.rand(1)

    def voxel_noise(self, n: int) -> torch.Tensor:
        return torch.rand(n)

    def fps_start(self, length: int) -> int:
        return int(torch.randint(high=int(length), size=(1,)).item())

    def randperm(self, n: int) -> torch.Tensor:
        return torch.randperm(n)


# ----------------------------------------------------------------------------------
# exact emulation of the torch-CPU float kernels that feed index decisions
# ----------------------------------------------------------------------------------
def fma32(a, b, c):
    """fl32(a*b + c): product exact in f64, one f64 add, then round to f32."""
    return (a.double() * b.double() + c.double()).float()


def sqrt_cr(v: torch.Tensor) -> torch.Tensor:
    """Correctly rounded f32 sqrt (sqrt in f64, then round: innocuous double rounding).
    NOTE torch.sqrt on a CPU f32 *vector* is NOT correctly rounded (about 0.6 % of inputs are
    1 ulp off on AVX512 builds), whereas the scalar sqrt inside torch.linalg.norm is - and so is
    CUDA's sqrtf without fast-math."""
    return v.double().sqrt().float()


def norm3(v: torch.Tensor) -> torch.Tensor:
    """torch.linalg.norm(v, dim=-1) for (N,3) f32 as the AVX2/AVX512 CPU kernel computes it."""
    x, y, z = v[..., 0], v[..., 1], v[..., 2]
    return sqrt_cr(fma32(z, z, fma32(y, y, x * x)))


def cumsum_f32(v: torch.Tensor) -> torch.Tensor:
    """torch.cumsum of an f32 vector on CPU: sequential f64 accumulation, rounded per element."""
    return torch.cumsum(v.double(), dim=0).float()


# ----------------------------------------------------------------------------------
# a2: CSR helpers
# ----------------------------------------------------------------------------------
def batch_to_ptr(batch: torch.Tensor, with_ends: bool = False) -> torch.Tensor:
    """point_ops.py:47-54: positions where a sorted id vector changes (+[0, N] with_ends)."""
    assert bool(torch.all(batch[1:] >= batch[:-1])), "ids must be sorted"
    inner = torch.nonzero(batch[1:] != batch[:-1]).flatten() + 1
    if with_ends:
        n = torch.tensor([batch.shape[0]], dtype=inner.dtype)
        inner = torch.cat([torch.zeros(1, dtype=inner.dtype), inner, n])
    return inner


def curve_ids_global(curve: torch.Tensor, batch: torch.Tensor) -> torch.Tensor:
    """point_ops.py:20-44: offset per-sample curve ids by the cumulative curve count of the
    preceding samples (count = last id of the sample + 1); identity for one sample (:27-28)."""
    ends = batch_to_ptr(batch, with_ends=True)
    if ends.shape[0] - 1 == 1:
        return curve
    assert bool(torch.all(curve[1:] - curve[:-1]) >= 0)  # (sic) point_ops.py:31,464-465
    last = curve[ends[1:-1] - 1] + 1
    offs = torch.cat([torch.zeros(1, dtype=last.dtype), torch.cumsum(last, dim=0)])
    return curve + offs[batch]


def pad_by_batch(t: torch.Tensor, batch: torch.Tensor):
    """point_ops.py:358-381 to_batch_padded: ragged (N,C) -> (B,Nmax,C) zero padded + mask."""
    ends = batch_to_ptr(batch, with_ends=True)
    counts = ends[1:] - ends[:-1]
    B = counts.shape[0]
    if B == 1:
        return t.unsqueeze(0), torch.ones((1, t.shape[0]), dtype=torch.bool)
    nmax = int(counts.max())
    local = torch.arange(batch.shape[0]) - ends[batch]
    out = torch.zeros((B, nmax) + tuple(t.shape[1:]), dtype=t.dtype)
    out[batch, local] = t
    mask = torch.zeros((B, nmax), dtype=torch.bool)
    mask[batch, local] = True
    return out, mask


def _padded(x, batch):
    """point_ops.py:264-284 dense2padded_pyg."""
    xp, mask = pad_by_batch(x, batch)
    lengths = mask.sum(dim=-1)
    offsets = torch.cumsum(lengths, dim=0)[:-1]
    return xp, mask, lengths, offsets


def _flatten_idx(idx, mask, offsets):
    """point_ops.py:287-290 pytorch3didx2pyg."""
    idx = idx.clone()
    idx[1:] += offsets.view(-1, 1, 1)
    return idx[mask]


# ----------------------------------------------------------------------------------
# a11: FPS
# ----------------------------------------------------------------------------------
def fps_count(lengths: torch.Tensor, ratio: float) -> torch.Tensor:
    """point_ops.py:63: K = ceil(lengths * ratio), evaluated in FLOAT32 (int64 tensor * python
    float promotes to f32) - e.g. 3000*0.3 -> 900.00006 -> 901."""
    return torch.ceil(lengths * ratio).to(torch.int64)


def fps_sorted(pos, batch, ratio, draws: Draws) -> torch.Tensor:
    """point_ops.py:57-70 fps_pytorch3d: per-cloud FPS from a random start, flattened to
    global indices and SORTED ascending (so only the selected set is observable)."""
    pp, mask = pad_by_batch(pos, batch)
    lengths = mask.sum(dim=-1)
    K = fps_count(lengths, ratio)
    starts = [draws.fps_start(int(l)) for l in lengths]
    idx = tp.sample_farthest_points(pp, lengths, K, starts)
    keep = idx != -1
    idx = idx.clone()
    idx[1:] += torch.cumsum(lengths, dim=0)[:-1].unsqueeze(-1)
    return idx[keep].sort()[0]


# ----------------------------------------------------------------------------------
# a13-a15: neighbour grouping
# ----------------------------------------------------------------------------------
def knn_ball_group(p1, p2, batch1, batch2, operation="ball-group", radius=None, knn=None,
                   accel_knn=True, return_dense=False):
    """point_ops.py:73-111 knn_ball_group_pytorch3d.
    "ball-group": pytorch3d.ball_query K=128 (:81); "knn" + accel_knn: FRNN fixed-radius kNN
    (:87, default radius 0.25); "knn" otherwise: exact knn_points (:91).  Dense return =
    (idx (B,P1,K), lengths2, mask1); else flat (row=query ordinal, col=point index) over the
    entries != -1 of valid queries, row-major."""
    a, m1, l1, o1 = _padded(p1, batch1)
    b, m2, l2, o2 = _padded(p2, batch2)
    if operation == "ball-group":
        assert radius is not None
        idx = tp.ball_query(a, b, l1, l2, 128, radius)
    elif operation == "knn":
        if accel_knn:
            if radius is None:
                radius = 0.25
            _, idx = tp.frnn_grid_points(a, b, l1, l2, knn, radius)
        else:
            _, idx = tp.knn_points(a, b, l1, l2, knn)
        if return_dense:
            return idx, l2, m1
    else:
        raise RuntimeError()
    B, P1, K = idx.shape
    keep = (idx != -1) & m1.unsqueeze(-1)
    col = _flatten_idx(idx, keep, o2)
    q = torch.arange(P1).view(1, P1, 1).repeat(B, 1, K)
    row = _flatten_idx(q, keep, o1)
    return row, col


# ----------------------------------------------------------------------------------
# a6: CurveFPS, a10: VoxelFPS
# ----------------------------------------------------------------------------------
def _edge_norms(pos, glob):
    """|p[i+1]-p[i]| along the point order, 0 across curve boundaries (fps_ops.py:19-21,
    point_ops.py:149-151)."""
    en = norm3(pos[1:] - pos[:-1])
    return torch.where(glob[1:] == glob[:-1], en, torch.zeros_like(en))


def curve_fps(pos, batch, curve, arclen: float, u: torch.Tensor) -> torch.Tensor:
    """fps_ops.py:16-39 CurveFPS.forward.  u = the torch.rand(1) draw of :30.
    s_i   = cumsum(edge norms)[i] - cumsum[start of i's curve]
    s_i  += ((start_i * 117) * u) % arclen
    keep i if round(s_i/arclen) != round(s_{i-1}/arclen), and every curve start."""
    glob = curve_ids_global(curve, batch)
    en = _edge_norms(pos, glob)
    starts = torch.cat([torch.zeros(1, dtype=torch.int64), batch_to_ptr(glob)])
    start_of = starts[glob]
    run = torch.cat([torch.zeros(1), cumsum_f32(en)])
    s = run - run[start_of]
    s = s + torch.remainder((start_of * 117) * u.to(torch.float32), arclen)
    q = torch.round(s / arclen)
    keep = torch.cat([torch.ones(1, dtype=torch.bool), q[1:] != q[:-1]])
    keep[start_of] = True
    return torch.nonzero(keep).flatten()


def voxel_fps(pos, batch, voxel: float, noise: torch.Tensor) -> torch.Tensor:
    """fps_ops.py:51-60 VoxelFPS.forward.  noise = the torch.rand(N) draw of :56.
    One point per occupied voxel (batch, floor(p/v)): argmin of |floor(p/v) - p/v| + noise*v/4,
    ties -> lowest index; output in lexicographic voxel-key order (torch.unique(dim=0))."""
    pv = pos / voxel
    cell = torch.floor(pv).long()
    key = torch.cat([batch.view(-1, 1), cell], dim=-1)
    _, inv = torch.unique(key, dim=0, return_inverse=True)
    d = norm3(cell - pv)
    d = d + noise.to(d) * voxel / 4
    size = int(inv.max()) + 1
    _, arg = tp.scatter_min(d.view(-1, 1), inv, 0, None, size)
    return arg.flatten()


# ----------------------------------------------------------------------------------
# a7: grouping along curves
# ----------------------------------------------------------------------------------
def _interleaved_offsets(k: int) -> torch.Tensor:
    """[0, -1, +1, -2, +2, ..., -k, +k]  (point_ops.py:170-171, 207-208)."""
    o = (torch.tensor([[-1, 1]]) * torch.arange(1, k + 1)[:, None]).flatten()
    return torch.cat([torch.zeros(1, dtype=torch.int64), o])


def radius_1d_group_subset(points, idxs, curve, batch, radius: float):
    """point_ops.py:143-193.  Per curve k_c = ceil(radius / (curve length / #points)) (inf -> 1),
    `radius / t` being reciprocal(t)*radius; K = min(max k_c, max #points); every query idxs[q]
    looks at idx + [0,-1,+1,...,-K,+K], keeps in-range same-curve candidates and of those the
    first k ones in that order, k = k_c looked up with the LOCAL curve id of the query (:185, a
    reference quirk for B>1 that is reproduced).  Returns (row=query ordinal, col=point index)."""
    N, Q = points.shape[0], idxs.shape[0]
    glob = curve_ids_global(curve, batch)
    en = _edge_norms(points, glob)
    ncurves = int(glob.max()) + 1
    clen = torch.zeros(ncurves).scatter_add_(0, glob[1:], en)          # global_add_pool :156
    cnt = torch.zeros(ncurves).scatter_add_(0, glob, torch.ones(N))      # :157
    avg = clen / cnt
    kc = torch.ceil(torch.reciprocal(avg) * radius)                      # :159 (__rtruediv__)
    kc[torch.isinf(kc)] = 1
    K = int(min(float(kc.max()), float(cnt.max())))                       # :165-167
    off = _interleaved_offsets(K)
    cand = idxs[:, None] + off[None, :]
    ok = (cand < N) & (cand >= 0)
    cand = torch.where(ok, cand, torch.zeros_like(cand))
    ok &= glob[idxs].view(Q, 1) == glob[cand]
    kq = kc[curve[idxs].long()]                                           # :185 LOCAL id
    ok &= torch.cumsum(ok, dim=1) <= kq[:, None]
    row = torch.arange(Q)[:, None].expand(Q, 2 * K + 1)[ok]
    return row, cand[ok]


def knn_1d_group_subset(points, idxs, curve, batch, knn: int):
    """point_ops.py:114-140 (reached by no shipped config), stated as a plain loop - small cases only: for query q the
    candidates idxs[q] + [0, -1, +1, ..., -knn, +knn] that are inside the array and on the query's (global) curve, the first
    knn of them in that order (the query itself counts, :133).  Returns (row = query ordinal, col = point index)."""
    glob = curve_ids_global(curve, batch).tolist()
    n = len(glob)
    rows, cols = [], []
    for q, c in enumerate(idxs.tolist()):
        kept = 0
        for s in range(2 * knn + 1):
            j = c + (0 if s == 0 else (-((s + 1) // 2) if s % 2 else s // 2))
            if 0 <= j < n and glob[j] == glob[c] and kept < knn:
                rows.append(q)
                cols.append(j)
                kept += 1
    return torch.tensor(rows, dtype=torch.int64), torch.tensor(cols, dtype=torch.int64)


def group_first_k_values(values, batch, k):
    """point_ops.py:384-428 (caller: scanning_simulator/utils/curve_cloud.py:223), stated as a plain loop - small cases only:
    for every batch id that occurs (ascending) its first min(count, k) values in order of appearance, zero padded to
    k_max = k or k.max(); k a per-occurring-batch tensor or an int.  Returns (grouped, mask)."""
    ids = batch.tolist()
    present = sorted(set(ids))
    per_batch = torch.is_tensor(k)
    k_max = int(k.max()) if per_batch else int(k)
    v2 = values.reshape(values.shape[0], -1)
    out = torch.zeros((len(present), k_max, v2.shape[1]), dtype=values.dtype)
    mask = torch.zeros((len(present), k_max), dtype=torch.bool)
    for p, g in enumerate(present):
        members = [i for i, b in enumerate(ids) if b == g]
        lim = min(len(members), int(k[p]) if per_batch else k_max)
        for t in range(lim):
            out[p, t] = v2[members[t]]
            mask[p, t] = True
    return (out.squeeze(-1) if values.dim() == 1 else out), mask


def knn_1d_group_superset(points, idxs, curve, batch, knn: int):
    """point_ops.py:196-260.  For every point i: a_i = #sampled points at positions <= i; the
    candidates are the sampled ordinals a_i + [0,-1,+1,...,-(knn+1),+(knn+1)] that exist and
    lie on i's curve; they are ordered by Euclidean distance to i (ties: candidate order) and
    the first knn are kept.  Returns (row=point index, col=sampled ORDINAL)."""
    N, Q = points.shape[0], idxs.shape[0]
    glob = curve_ids_global(curve, batch)
    hit = torch.zeros(N, dtype=torch.bool)
    hit[idxs] = True
    a = torch.cumsum(hit, dim=0)
    off = _interleaved_offsets(knn + 1)
    W = off.shape[0]
    cand = a[:, None] + off[None, :]
    ok = (cand < Q) & (cand >= 0)
    cand = torch.where(ok, cand, torch.zeros_like(cand))
    cpt = idxs[cand]                                                     # (N, W) point index
    ok &= glob.view(N, 1) == glob[cpt]
    d = norm3(points[cpt] - points.view(N, 1, 3))
    d = torch.where(ok, d, torch.full_like(d, 100.0))
    order = torch.sort(d, dim=1, stable=True)[1]
    cand = torch.gather(cand, 1, order)
    ok = torch.gather(ok, 1, order)
    ok &= torch.cumsum(ok, dim=1) <= knn
    assert bool(torch.all(ok.sum(dim=1) > 0))
    row = torch.arange(N)[:, None].expand(N, W)[ok]
    return row, cand[ok]


# ----------------------------------------------------------------------------------
# a19/a20: inverse-distance interpolation
# ----------------------------------------------------------------------------------
def _blend(x, x_idx, y_idx, w, n_out):
    num = tp.scatter_add(x[x_idx] * w, y_idx, 0, None, n_out)
    den = tp.scatter_add(w, y_idx, 0, None, n_out)
    return num / den


def knn_interpolate(x, pos_x, pos_y, batch_x, batch_y, k=3):
    """point_ops.py:293-341: exact k-NN of every pos_y in pos_x (knn_points, no radius),
    w = 1/clamp(d2, 1e-16), y = sum(w x)/sum(w)."""
    y_idx, x_idx = knn_ball_group(pos_y, pos_x, batch_y, batch_x, operation="knn", knn=k, accel_knn=False)
    diff = pos_x[x_idx] - pos_y[y_idx]
    d2 = (diff * diff).sum(dim=-1, keepdim=True)
    w = 1.0 / torch.clamp(d2, min=1e-16)
    return _blend(x, x_idx, y_idx, w, pos_y.shape[0])


def knn_interpolate_1d(x, idx, pos_y, batch_y, curve_y, k):
    """point_ops.py:344-355: same blend over the along-curve neighbours of knn_1d_group_superset;
    the weight uses norm()**2."""
    pos_x = pos_y[idx]
    y_idx, x_idx = knn_1d_group_superset(pos_y, idx, curve_y, batch_y, k)
    d = (pos_x[x_idx] - pos_y[y_idx]).norm(dim=-1, keepdim=True) ** 2
    w = 1.0 / torch.clamp(d, min=1e-16)
    return _blend(x, x_idx, y_idx, w, pos_y.shape[0])


# ----------------------------------------------------------------------------------
# a3: curve grouping of a raw sweep
# ----------------------------------------------------------------------------------
def get_curves(points, beam_ids, curve_thresh: float = 0.08):
    """src/data/nuscenes_dataset.py:91-118 (== kitti_dataset.py:73-92): stable sort by beam id;
    a new curve starts where the beam changes or the f64 edge length exceeds
    thresh*sqrt(|p[i+1].xy|) (radius in f32); curve id = running count of splits.
    Returns (reorder, inverse reorder, curve_idxs)."""
    reorder = torch.sort(beam_ids, stable=True)[1]
    inv = torch.empty_like(reorder)
    inv[reorder] = torch.arange(points.shape[0])
    p = points[reorder]
    b = beam_ids[reorder]
    beam_split = (b[1:] - b[:-1]) != 0
    e = p[1:].double() - p[:-1].double()
    en = torch.linalg.norm(e, dim=-1)
    rad = torch.linalg.norm(p[1:, :2], dim=-1)
    # the reference's torch.sqrt(f32 vector) is up to 1 ulp off on CPU (see sqrt_cr); the oracle
    # fixes the correctly rounded value, which is also what the CUDA kernel computes
    split = (en > (curve_thresh * sqrt_cr(rad))) | beam_split
    cid = torch.cat([torch.zeros(1, dtype=torch.int64), torch.cumsum(split, dim=0)])
    return reorder, inv, cid
