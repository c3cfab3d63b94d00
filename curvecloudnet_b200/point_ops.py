"""The reference's op-level API (src/models/utils/point_ops.py, src/models/modules/fps_ops.py) on the
sm_100a kernels: same function names, argument meaning, return layouts (int64 indices, (row, col)
edge lists, padded (B, P, K) tables) and error behaviour.  The model path (modules.py) uses the
int32/CSR forms from `ops` directly; the conversions to the reference's layouts below are plain
reshapes of kernel outputs at the API boundary.
"""
from __future__ import annotations

from typing import Optional

import torch

from . import ops
from .modules import fps_sorted
from .nn import DEFAULT_DRAWS, Draws


def _need_cuda(t: torch.Tensor, what: str):
    if not t.is_cuda:
        # the reference's accelerated kNN raises on CPU tensors as well (point_ops.py:439-440)
        raise TypeError("%s: CUDA tensors required (curvecloudnet_b200 has no CPU path)" % what)


def curveidx_local2global(point2curveidx: torch.Tensor, batch: torch.Tensor) -> torch.Tensor:
    """point_ops.py:20-44."""
    _need_cuda(batch, "curveidx_local2global")
    return ops.curve_csr(batch, point2curveidx).curve_glob


def batch2ptr(batch: torch.Tensor, with_ends: bool = False) -> torch.Tensor:
    """point_ops.py:47-54: positions where the sorted id vector changes (+ [0, N] with_ends)."""
    _need_cuda(batch, "batch2ptr")
    ptr = ops.sample_ptr_any(batch)
    return ptr if with_ends else ptr[1:-1]


def to_batch_padded(t: torch.Tensor, batch: torch.Tensor):
    """point_ops.py:358-381: ragged (N,C) -> (B,Nmax,C) zero padded + mask."""
    _need_cuda(t, "to_batch_padded")
    sptr = ops.sample_ptr(batch)
    B, lens = sptr.n_samples, sptr.lengths
    if B == 1:
        return t.unsqueeze(0), torch.ones((1, t.shape[0]), dtype=torch.bool, device=t.device)
    nmax = max(lens)
    out = torch.zeros((B, nmax) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
    mask = torch.zeros((B, nmax), dtype=torch.bool, device=t.device)
    for b in range(B):
        out[b, :lens[b]] = t[sptr.host[b]:sptr.host[b + 1]]
        mask[b, :lens[b]] = True
    return out, mask


def fps_pytorch3d(pos: torch.Tensor, batch: torch.Tensor, ratio: float, draws: Draws = DEFAULT_DRAWS) -> torch.Tensor:
    """point_ops.py:57-70: per-cloud farthest point sampling, result sorted ascending."""
    _need_cuda(pos, "fps_pytorch3d")
    return fps_sorted(pos, ops.sample_ptr(batch), ratio, draws)[0]


def _dense_local(idx: torch.Tensor, batch1: torch.Tensor, sptr1: ops.SamplePtr, sptr2: ops.SamplePtr, fill: int):
    """(n1, K) global int32 table -> the reference's padded (B, P1max, K) int64 table of LOCAL indices."""
    B, K = sptr1.n_samples, idx.shape[1]
    p1max = max(sptr1.lengths) if B else 0
    out = torch.full((B, p1max, K), fill, dtype=torch.int64, device=idx.device)
    mask1 = torch.zeros((B, p1max), dtype=torch.bool, device=idx.device)
    for b in range(B):
        lo, hi = sptr1.host[b], sptr1.host[b + 1]
        blk = idx[lo:hi].to(torch.int64)
        out[b, :hi - lo] = torch.where(blk >= 0, blk - sptr2.host[b], blk)
        mask1[b, :hi - lo] = True
    return out, mask1


def fast_knn(points1, points2, lengths1, lengths2, K, r, return_nn=False):
    """point_ops.py:431-461 on padded inputs: idx (B,P1,K) int64, -1 padded."""
    if not points1.is_cuda:
        raise TypeError("fast_knn only supports CUDA tensors")       # point_ops.py:439-440
    B = points1.shape[0]
    l1, l2 = [int(v) for v in lengths1.tolist()], [int(v) for v in lengths2.tolist()]
    p1 = torch.cat([points1[b, :l1[b]] for b in range(B)]).contiguous()
    p2 = torch.cat([points2[b, :l2[b]] for b in range(B)]).contiguous()
    b1 = torch.cat([torch.full((l1[b],), b, dtype=torch.int64, device=p1.device) for b in range(B)])
    b2 = torch.cat([torch.full((l2[b],), b, dtype=torch.int64, device=p1.device) for b in range(B)])
    s1, s2 = ops.sample_ptr(b1), ops.sample_ptr(b2)
    idx, _, _ = ops.frnn_knn(p1, b1, p2, b2, s2, K, r)
    dense, _ = _dense_local(idx, b1, s1, s2, -1)
    out = torch.full((B, points1.shape[1], K), -1, dtype=torch.int64, device=p1.device)
    out[:, :dense.shape[1]] = dense
    return out


def knn_ball_group_pytorch3d(p1, p2, batch1, batch2, operation="ball-group", radius=None, knn=None, accel_knn=True,
                             return_dense=False):
    """point_ops.py:73-111.  "ball-group": first 128 in radius (index order); "knn" + accel_knn: fixed-radius
    kNN (default radius 0.25); "knn" otherwise: exact kNN.  Dense return = (idx (B,P1,K) local, lengths2,
    mask1); else flat (row = query ordinal, col = global point index) over the valid entries, row-major."""
    _need_cuda(p1, "knn_ball_group_pytorch3d")
    s1, s2 = ops.sample_ptr(batch1), ops.sample_ptr(batch2)
    if operation == "ball-group":
        assert radius is not None
        idx, _ = ops.ball_query(p1, batch1, p2, s2, radius, 128)
    elif operation == "knn":
        if accel_knn:
            idx, _, _ = ops.frnn_knn(p1, batch1, p2, batch2, s2, knn, 0.25 if radius is None else radius)
        else:
            idx, _ = ops.knn_exact(p1, batch1, p2, s2, knn, zero_fill=True, want_d2=False)
        if return_dense:
            dense, mask1 = _dense_local(idx, batch1, s1, s2, -1 if accel_knn else 0)
            return dense, torch.tensor(s2.lengths, dtype=torch.int64, device=p1.device), mask1
    else:
        raise RuntimeError()
    keep = idx >= 0
    row = torch.arange(idx.shape[0], device=idx.device).view(-1, 1).expand_as(idx)[keep]
    return row, idx[keep].to(torch.int64)


def radius_1d_group_subset(points, idxs, point2curveidx, batch, radius):
    """point_ops.py:143-193 -> (row = query ordinal, col = point index)."""
    _need_cuda(points, "radius_1d_group_subset")
    topo = ops.curve_csr(batch, point2curveidx)
    e = ops.radius_1d_group(points, topo, idxs, radius)
    return e.erow.to(torch.int64), e.ecol.to(torch.int64)


def knn_1d_group_subset(points, idxs, point2curveidx, batch, knn):
    """point_ops.py:114-140 -> (row = query ordinal, col = point index): up to knn points around each query along the point
    order and on its curve, the query itself first.  (The reference forms the candidate indices in float32; identical below
    2^24 points.)"""
    _need_cuda(points, "knn_1d_group_subset")
    e = ops.knn_1d_group(curveidx_local2global(point2curveidx, batch), idxs, int(knn))
    return e.erow.to(torch.int64), e.ecol.to(torch.int64)


def knn_1d_group_superset(points, idxs, point2curveidx, batch, knn):
    """point_ops.py:196-260 -> (row = point index, col = sampled ORDINAL)."""
    _need_cuda(points, "knn_1d_group_superset")
    topo = ops.curve_csr(batch, point2curveidx)
    nbr, _ = ops.knn_1d_superset(points, topo.curve_ord, idxs, knn)
    keep = nbr >= 0
    assert bool(torch.all(keep.sum(dim=1) > 0))                          # point_ops.py:252
    row = torch.arange(nbr.shape[0], device=nbr.device).view(-1, 1).expand_as(nbr)[keep]
    return row, nbr[keep].to(torch.int64)


def knn_interpolate_pytorch3d(x, pos_x, pos_y, batch_x=None, batch_y=None, k=3):
    """point_ops.py:293-341."""
    _need_cuda(x, "knn_interpolate_pytorch3d")
    nbr, d2 = ops.knn_exact(pos_y, batch_y, pos_x, ops.sample_ptr(batch_x), k, zero_fill=True, want_d2=True)
    return ops.interp_blend(x, nbr, d2, from_d2=True)


def knn_interpolate_1D_pytorch3d(x, idx, pos_y, batch_y, point2curveidx_y, k):
    """point_ops.py:344-355."""
    _need_cuda(x, "knn_interpolate_1D_pytorch3d")
    topo = ops.curve_csr(batch_y, point2curveidx_y)
    nbr, wgt = ops.knn_1d_superset(pos_y, topo.curve_ord, idx, k)
    return ops.interp_blend(x, nbr, wgt, from_d2=False)


class CurveFPS(torch.nn.Module):
    """fps_ops.py:7-39."""

    def __init__(self, arclen_spacing=0.3):
        super().__init__()
        self.arclen_spacing = arclen_spacing

    def forward(self, pos, batch, point2curveidx, draws: Draws = DEFAULT_DRAWS):
        _need_cuda(pos, "CurveFPS")
        topo = ops.curve_csr(batch, point2curveidx)
        return ops.curve_fps(pos, topo, self.arclen_spacing, draws.curve_phase())[0]


class VoxelFPS(torch.nn.Module):
    """fps_ops.py:42-60."""

    def __init__(self, voxel_size=0.05):
        super().__init__()
        self.voxel_size = voxel_size

    def forward(self, pos, batch, draws: Draws = DEFAULT_DRAWS):
        _need_cuda(pos, "VoxelFPS")
        sptr = ops.sample_ptr(batch)
        return ops.voxel_fps(pos, batch, self.voxel_size, draws.voxel_noise(pos.shape[0]), sptr.n_samples)[0]


def compute_feature_diffs(x, point2curveidx, batch):
    """fast_conv1d.py:190-205."""
    _need_cuda(x, "compute_feature_diffs")
    topo = ops.curve_csr(batch, point2curveidx)
    c = x.shape[1]
    return ops.curve_feat_prep(x, x, False, True, topo.curve_ord)[:, c:]


def get_curves(points, beam_ids, labels=None, reflectance=None, curve_thresh: float = 0.08, n_beams=None):
    """SemNuScenes._get_curves / SemKITTI._get_curves (src/data/nuscenes_dataset.py:91-118, kitti_dataset.py:73-92) on the
    device: -> (points sorted by beam, curve_idxs, labels sorted, reflectance sorted, inv_reorder) like the reference
    method; points may carry extra columns after xyz (they are reordered with them)."""
    reorder, inv, cid, sorted_xyz, counts = ops.get_curves(points, beam_ids, curve_thresh, n_beams)
    if int(counts[1].item()):                                    # host sync (error flag)
        raise ValueError("get_curves: beam id outside [0, n_beams)")
    pts = sorted_xyz if points.shape[1] == 3 else ops.gather_rows(points, reorder)
    lab = None if labels is None else (ops.gather_i64(labels.reshape(-1), reorder).reshape(-1, *labels.shape[1:])
                                       if labels.dtype == torch.int64 else labels[reorder])
    ref = None if reflectance is None else ops.gather_rows(reflectance.reshape(reflectance.shape[0], -1).float(), reorder)
    return pts, cid, lab, ref, inv


def dense2padded_pyg(x: torch.Tensor, batch: torch.Tensor):
    """point_ops.py:264-284 -> (x_padded (B, Nmax, C), mask (B, Nmax), lengths (B,), offsets (B-1,))."""
    x_padded, mask = to_batch_padded(x, batch)
    lengths = mask.sum(dim=-1)
    return x_padded, mask, lengths, torch.cumsum(lengths, dim=0)[:-1]


def pytorch3didx2pyg(idx: torch.Tensor, mask: torch.Tensor, offsets: torch.Tensor) -> torch.Tensor:
    """point_ops.py:287-290: padded (B, P, K) local indices -> flat global indices of the entries selected by mask
    (shifts idx in place, as the reference does)."""
    idx[1:, :] += offsets.view(-1, 1, 1)
    return idx[mask].flatten()


def is_sorted(t: torch.Tensor) -> torch.Tensor:
    """point_ops.py:464-465, quirk kept: `all(diff) >= 0` compares the BOOLEAN all-differences-nonzero with 0, so the result
    is always True (the assert of curveidx_local2global, :31, never fires upstream either)."""
    return torch.all(t[1:] - t[:-1]) >= 0


def group_first_k_values(values: torch.Tensor, batch: torch.Tensor, k):
    """point_ops.py:384-428 -> (grouped (num_batches, k_max[, D]) zero padded, mask (num_batches, k_max)): the first k values of
    every batch id that occurs, in order of appearance; k an int or a per-batch tensor (the limit of batch p is k[p], the
    width k.max()).  Host syncs: the largest id and the per-id counts (the reference synchronises in torch.unique)."""
    _need_cuda(values, "group_first_k_values")
    if values.dtype != torch.float32:
        raise TypeError("group_first_k_values: float32 values (got %s)" % values.dtype)
    one_d = values.dim() == 1
    v2 = values.view(values.shape[0], -1)
    k_max = int(torch.max(k).item()) if torch.is_tensor(k) else int(k)
    n_ids = int(batch.max().item()) + 1 if batch.numel() else 0                # host sync
    out, cnt = ops.group_first_k(v2, batch, n_ids, k_max)
    present = cnt > 0
    out, cnt = out[present], cnt[present].to(torch.int64)                          # ids that do not occur vanish (torch.unique)
    lim = torch.clamp(cnt, max=k.to(cnt.device) if torch.is_tensor(k) else k_max)
    mask = torch.arange(k_max, device=values.device).unsqueeze(0) < lim.unsqueeze(1)
    out = torch.where(mask.unsqueeze(-1), out, torch.zeros_like(out))
    return (out.squeeze(-1) if one_d else out), mask
