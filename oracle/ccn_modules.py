"""CPU restatement of the step modules in /root/reference/src/models/modules/*.py
(SURVEY.md §8(a) a4,a5,a8,a9,a17-a22).  TEST INFRASTRUCTURE - see oracle/__init__.py.

Same class names, constructor arguments, forward signatures and parameter names as the
reference so a state_dict moves freely between reference, oracle and the CUDA modules.
Differences in form only: third-party calls go to oracle.tp_ops, random draws come from a
`Draws` object (`draws=` keyword, default = torch's global RNG like the reference).
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import ccn_point_ops as po
from . import tp_ops as tp
from .ccn_point_ops import Draws

_DEFAULT_DRAWS = Draws()


def _cat_xyz(x, pos, with_xyz):
    """the `x=None / with_xyz` preamble shared by the modules (e.g. fast_conv1d.py:42-45)."""
    if not with_xyz:
        return x
    return pos if x is None else torch.cat([x, pos], dim=1)


# ----------------------------------------------------------------------------------
# fast_conv1d.py
# ----------------------------------------------------------------------------------
def feature_diffs(x, curve, batch):
    """fast_conv1d.py:190-205 compute_feature_diffs: |mean of the backward and forward feature
    differences along the curve| (a one-sided difference at curve ends, 0 for 1-point curves)."""
    glob = po.curve_ids_global(curve, batch)
    same = glob[1:] == glob[:-1]
    e = torch.where(same[:, None], x[1:] - x[:-1], torch.zeros_like(x[1:]))
    z = torch.zeros((1, x.shape[1]), dtype=x.dtype)
    e = torch.cat([z, e, z], dim=0)
    v = torch.cat([torch.zeros(1, dtype=torch.bool), same, torch.zeros(1, dtype=torch.bool)]).to(x.dtype)
    den = torch.clip(v[1:] + v[:-1], min=1)
    return torch.abs((e[1:] + e[:-1]) / den[:, None])


class SymmetricConv1d(nn.Module):
    """fast_conv1d.py:148-187: Conv1d whose learned weight (C_out, C_in, k//2+1) is mirrored
    around tap 0 into a width-k kernel; 'same' zero padding; bias on by default."""

    def __init__(self, in_channels, out_channels, kernel_size, bias=True):
        super().__init__()
        self.weight = nn.Parameter(torch.empty(out_channels, in_channels, kernel_size))
        self.bias = nn.Parameter(torch.empty(out_channels)) if bias else None
        nn.init.kaiming_uniform_(self.weight, a=5 ** 0.5)
        if bias:
            nn.init.uniform_(self.bias, -0.1, 0.1)

    def full_weight(self):
        w = self.weight
        if w.shape[2] > 1:
            w = torch.cat([torch.flip(w[:, :, 1:], dims=[2]), w], dim=2)
        return w

    def forward(self, x):  # x (1?, C_in, L)
        return F.conv1d(x.to(self.weight.dtype), self.full_weight(), self.bias, padding="same")


def _pad_layout(glob, n_points, pad, with_ends):
    """Row layout of the zero-padded signal: `pad` rows are inserted at every pointer of
    batch2ptr(glob, with_ends) (fast_conv1d.py:48-61 / :113-125).  Returns (n_rows, valid mask)."""
    ptr = po.batch_to_ptr(glob, with_ends=with_ends)
    n_ptr = ptr.shape[0]
    n_rows = n_points + n_ptr * pad
    valid = torch.ones(n_rows, dtype=torch.bool)
    if pad > 0 and n_ptr > 0:
        pad_rows = ptr.view(-1, 1) + torch.arange(n_ptr * pad).view(-1, pad)
        valid[pad_rows.flatten()] = False
    return n_rows, valid


class _CurveConvBase(nn.Module):
    def __init__(self, feat_dims, kernel_size, bias, with_xyz, with_diff, diff_every_layer):
        super().__init__()
        self.kernel_size, self.feat_dims = kernel_size, feat_dims
        self.with_xyz, self.with_diff = with_xyz, with_diff
        ku = kernel_size // 2 + 1
        convs, norms = [], []
        for i in range(1, len(feat_dims)):
            doubled = with_diff and (diff_every_layer or i == 1)
            convs.append(SymmetricConv1d(feat_dims[i - 1] * (2 if doubled else 1), feat_dims[i], ku, bias=bias))
            norms.append(nn.BatchNorm1d(feat_dims[i]))
        self.conv_modules = nn.ModuleList(convs)
        self.norm_modules = nn.ModuleList(norms)


class SymmetricCurve1DConvV2(_CurveConvBase):
    """fast_conv1d.py:11-75.  One padded signal for the whole stack: (k//2)*L zero rows before
    every curve and after the last one; conv + BN + LeakyReLU run over pad rows too, so from
    layer 2 on the taps at curve ends read the (non-zero) pad-row activations."""

    def __init__(self, feat_dims=(64, 64, 128), kernel_size=5, bias=True, device=None, dtype=None,
                 with_xyz=False, with_diff=False):
        super().__init__(feat_dims, kernel_size, bias, with_xyz, with_diff, diff_every_layer=False)

    def forward(self, x, pos, batch, point2curveidx, **kwargs):
        glob = po.curve_ids_global(point2curveidx, batch)
        assert bool(torch.all(glob[1:] - glob[:-1] >= 0))
        x = _cat_xyz(x, pos, self.with_xyz)
        pad = (self.kernel_size // 2) * (len(self.feat_dims) - 1)
        n_rows, valid = _pad_layout(glob, x.shape[0], pad if self.kernel_size > 1 else 0, with_ends=True)
        if self.with_diff:
            x = torch.cat([x, feature_diffs(x, point2curveidx, batch)], dim=1)
        sig = torch.zeros((n_rows, x.shape[1]), dtype=x.dtype)
        sig[valid] = x
        for conv, norm in zip(self.conv_modules, self.norm_modules):
            sig = conv(sig.t()).t()
            sig = F.leaky_relu(norm(sig))
        return sig[valid], pos, batch, point2curveidx


class SymmetricCurve1DConvFastV1(_CurveConvBase):
    """fast_conv1d.py:78-145.  Per layer: (re)compute the feature differences (C_in doubles at
    EVERY layer), fresh k//2 zero rows at interior curve boundaries only, conv, drop pad rows,
    BN over valid rows, LeakyReLU."""

    def __init__(self, feat_dims=(64, 64, 128), kernel_size=5, bias=True, device=None, dtype=None,
                 with_xyz=False, with_diff=False):
        super().__init__(feat_dims, kernel_size, bias, with_xyz, with_diff, diff_every_layer=True)

    def forward(self, x, pos, batch, point2curveidx, **kwargs):
        glob = po.curve_ids_global(point2curveidx, batch)
        assert bool(torch.all(glob[1:] - glob[:-1] >= 0))
        x = _cat_xyz(x, pos, self.with_xyz)
        pad = self.kernel_size // 2
        n_rows, valid = _pad_layout(glob, x.shape[0], pad if self.kernel_size > 1 else 0, with_ends=False)
        for conv, norm in zip(self.conv_modules, self.norm_modules):
            if self.with_diff:
                x = torch.cat([x, feature_diffs(x, point2curveidx, batch)], dim=1)
            sig = torch.zeros((n_rows, x.shape[1]), dtype=x.dtype)
            sig[valid] = x
            x = conv(sig.t()).t()[valid]
            x = F.leaky_relu(norm(x))
        return x, pos, batch, point2curveidx


# ----------------------------------------------------------------------------------
# point_conv.py
# ----------------------------------------------------------------------------------
class PointNetConv2(nn.Module):
    """point_conv.py:12-93.  edge_index = [source j ; target i].  message (:60-69) =
    local_nn(cat[x_j, (pos_j - pos_i) / normalize_radius]); aggregate (:71-93) over the target:
    max | mean | sum sigmoid(attend_nn(m)) * m | sum softmax_target(attend_nn(m)) * m; the two
    attention sums are sized max(target)+1 (no dim_size, :87,:92)."""

    def __init__(self, local_nn=None, global_nn=None, attend_nn=None, add_self_loops=True,
                 aggr_type="max", normalize_radius=None, **kwargs):
        super().__init__()
        assert aggr_type in ["max", "attend", "mean", "weighted-sum"]
        assert not add_self_loops, "every reference call site passes add_self_loops=False"
        self.local_nn, self.global_nn, self.attend_nn = local_nn, global_nn, attend_nn
        self.aggr_type, self.normalize_radius = aggr_type, normalize_radius

    def forward(self, x, pos, edge_index):
        x_src = x[0] if isinstance(x, tuple) else x
        pos_src, pos_dst = pos if isinstance(pos, tuple) else (pos, pos)
        j, i = edge_index[0], edge_index[1]
        msg = pos_src[j] - pos_dst[i]
        if self.normalize_radius is not None:
            msg = msg / self.normalize_radius
        if x_src is not None:
            msg = torch.cat([x_src[j], msg], dim=1)
        if self.local_nn is not None:
            msg = self.local_nn(msg)
        n_dst = pos_dst.shape[0]
        if self.aggr_type == "max":
            out = tp.scatter_max(msg, i, 0, None, n_dst)[0]
        elif self.aggr_type == "mean":
            out = tp.scatter_mean(msg, i, 0, None, n_dst)
        elif self.aggr_type == "weighted-sum":
            out = tp.scatter_add(msg * torch.sigmoid(self.attend_nn(msg)), i, 0)
        else:
            out = tp.scatter_add(msg * tp.segment_softmax(self.attend_nn(msg), i), i, 0)
        if self.global_nn is not None:
            out = self.global_nn(out)
        return out


# ----------------------------------------------------------------------------------
# pointnet2.py
# ----------------------------------------------------------------------------------
class SAModule(nn.Module):
    """pointnet2.py:33-78.  down-sample ('random' | 'curve-fps' | 'voxel' | 'fps'), group with
    FRNN kNN (use_fast_knn) or ball_query K=128, PointNetConv2, index pos/batch/curve."""

    def __init__(self, ratio, r, nn, k, curve_fps_arclen=None, voxel_size=None, downsample_type="random",
                 attend_nn=None, aggr_type="max", normalize_radius=False, use_fast_knn=True, **kwargs):
        super().__init__()
        assert downsample_type in ["curve-fps", "random", "fps", "voxel"]
        self.ratio, self.r, self.knn = ratio, r, k
        self.downsample_type, self.use_fast_knn = downsample_type, use_fast_knn
        self.curve_fps_arclen, self.voxel_size = curve_fps_arclen, voxel_size
        self.normalize_radius = r if normalize_radius else None
        self.conv = PointNetConv2(nn, add_self_loops=False, aggr_type=aggr_type, attend_nn=attend_nn,
                                  normalize_radius=self.normalize_radius)

    def forward(self, x, pos, batch, point2curveidx=None, draws: Draws = _DEFAULT_DRAWS, **kwargs):
        if self.downsample_type == "random":
            n = int(pos.shape[0] * self.ratio)
            idx = torch.sort(draws.randperm(pos.shape[0])[:n])[0]
        elif self.downsample_type == "curve-fps":
            idx = po.curve_fps(pos, batch, point2curveidx, self.curve_fps_arclen, draws.curve_phase())
        elif self.downsample_type == "voxel":
            idx = po.voxel_fps(pos, batch, self.voxel_size, draws.voxel_noise(pos.shape[0]))
        else:
            idx = po.fps_sorted(pos, batch, self.ratio, draws)
        op = "knn" if self.use_fast_knn else "ball-group"
        row, col = po.knn_ball_group(pos[idx], pos, batch[idx], batch, operation=op, knn=self.knn, radius=self.r)
        x_dst = None if x is None else x[idx]
        x = self.conv((x, x_dst), (pos, pos[idx]), torch.stack([col, row], dim=0))
        curve = None if point2curveidx is None else point2curveidx[idx]
        return x, pos[idx], batch[idx], curve


class GlobalSAModule(nn.Module):
    """pointnet2.py:81-116: nn(cat[x,pos]) then per-sample max (or mean); state = first point
    of every sample."""

    def __init__(self, nn, **kwargs):
        super().__init__()
        self.nn = nn
        self.pooling = kwargs.get("pooling", "max")

    def forward(self, x, pos, batch, point2curveidx=None, **kwargs):
        x = self.nn(torch.cat([x, pos], dim=1))
        size = int(batch.max()) + 1
        if self.pooling == "max":
            x = tp.scatter_max(x, batch, 0, None, size)[0]
        elif self.pooling == "mean":
            x = tp.scatter_mean(x, batch, 0, None, size)
        else:
            raise NotImplementedError(self.pooling)
        first = torch.cat([torch.zeros(1, dtype=torch.int64), po.batch_to_ptr(batch)])
        curve = None if point2curveidx is None else point2curveidx[first]
        return x, pos[first], batch[first], curve


def _fp_concat(x, x_skip, pos_skip, with_xyz):
    """pointnet2.py:132-139 / :194-201."""
    if x_skip is not None and not with_xyz:
        return torch.cat([x, x_skip], dim=1)
    if x_skip is not None and with_xyz:
        return torch.cat([x, x_skip, pos_skip[:, :3]], dim=1)
    if x_skip is None and with_xyz:
        return torch.cat([x, pos_skip[:, :3]], dim=1)
    return x


class FPModule(nn.Module):
    """pointnet2.py:119-143: 3-NN inverse-d2 interpolation onto the skip level, concat, MLP."""

    def __init__(self, k, nn, with_xyz=False):
        super().__init__()
        self.k, self.nn, self.with_xyz = k, nn, with_xyz

    def forward(self, x, pos, batch, x_skip, pos_skip, batch_skip, point2curveidx=None,
                point2curveidx_skip=None, **kwargs):
        x = po.knn_interpolate(x, pos, pos_skip, batch, batch_skip, k=self.k)
        assert not bool(torch.any(torch.isnan(x)))
        x = self.nn(_fp_concat(x, x_skip, pos_skip, self.with_xyz))
        return x, pos_skip, batch_skip, point2curveidx_skip


class CurveSAModule(nn.Module):
    """pointnet2.py:146-181: CurveFPS (or FPS) -> radius_1d_group_subset -> PointNetConv2;
    returns the down-sampling indices as 6th output for fp-geo."""

    def __init__(self, ratio, r, nn, curve_fps_arclen=None, use_curve_fps=False, global_nn=None, attend_nn=None,
                 with_xyz=False, aggr_type="max", normalize_radius=False, **kwargs):
        super().__init__()
        self.ratio, self.r = ratio, r
        self.curve_fps_arclen, self.use_curve_fps, self.with_xyz = curve_fps_arclen, use_curve_fps, with_xyz
        self.normalize_radius = r if normalize_radius else None
        self.conv = PointNetConv2(nn, add_self_loops=False, global_nn=global_nn, aggr_type=aggr_type,
                                  attend_nn=attend_nn, normalize_radius=self.normalize_radius)

    def forward(self, x, pos, batch, point2curveidx, draws: Draws = _DEFAULT_DRAWS, **kwargs):
        if self.with_xyz:
            x = pos[:, :3] if x is None else torch.cat([x, pos[:, :3]], dim=1)
        if self.use_curve_fps:
            idx = po.curve_fps(pos, batch, point2curveidx, self.curve_fps_arclen, draws.curve_phase())
        else:
            idx = po.fps_sorted(pos, batch, self.ratio, draws)
        row, col = po.radius_1d_group_subset(pos, idx, point2curveidx, batch, self.r)
        x_dst = None if x is None else x[idx]
        x = self.conv((x, x_dst), (pos, pos[idx]), torch.stack([col, row], dim=0))
        return x, pos[idx], batch[idx], point2curveidx[idx], None, idx


class CurveFPModule(FPModule):
    """pointnet2.py:184-205: interpolation along the curve onto the fine points."""

    def forward(self, x, idx, x_skip, pos_skip, batch_skip, point2curveidx_skip=None, **kwargs):
        x = po.knn_interpolate_1d(x, idx, pos_skip, batch_skip, point2curveidx_skip, k=self.k)
        assert not bool(torch.any(torch.isnan(x)))
        x = self.nn(_fp_concat(x, x_skip, pos_skip, self.with_xyz))
        return x, pos_skip, batch_skip, point2curveidx_skip


# ----------------------------------------------------------------------------------
# dgcnn.py (StaticEdgeConv / SGCNNLayer; DynamicEdgeConv / DGCNNLayer / DGCNNLayerRadius - reached by no shipped config)
# ----------------------------------------------------------------------------------
class SGCNNLayer(nn.Module):
    """dgcnn.py:130-266.  kNN graph in xyz.
    forward_fast (:158-207, use_sparse_feat_agg False): FRNN idx (B,N,K) + a self loop in column 0;
      edge feature [x_j, x_i - x_j] with x_j = 0 where idx == -1; the MLP runs over all B*N*(K+1)
      rows; max with invalid rows set to -100 (mean / weighted-sum / attend variants).
    forward_slow (:209-246): edge list from FRNN or knn_points; message nn([x_i, x_j - x_i]);
      scatter_max or softmax attention."""

    def __init__(self, nn, k, aggr="max", r=1.0, num_workers=1, with_xyz=False, attend_nn=None,
                 aggr_type="max", use_sparse_feat_agg=False, use_fast_knn=True, **kwargs):
        super().__init__()
        assert aggr_type in ["max", "attend", "mean", "weighted-sum"]
        self.nn, self.k, self.r, self.with_xyz = nn, k, r, with_xyz
        self.attend_nn, self.aggr_type = attend_nn, aggr_type
        self.use_fast_knn, self.use_sparse_feat_agg = use_fast_knn, use_sparse_feat_agg

    def forward(self, x, pos, batch, point2curveidx=None, **kwargs):
        x = _cat_xyz(x, pos, self.with_xyz)
        out = self._sparse(x, pos, batch) if self.use_sparse_feat_agg else self._dense(x, pos, batch)
        return out, pos, batch, point2curveidx

    def _dense(self, x, pos, batch):
        idxs, len2, mask1 = po.knn_ball_group(pos, pos, batch, batch, operation="knn", knn=self.k,
                                              radius=self.r, return_dense=True)
        B, N = idxs.shape[:2]
        loops = torch.arange(N).view(1, N, 1).repeat(B, 1, 1)
        idxs = torch.cat([loops, idxs], dim=2)
        xp = po.pad_by_batch(x, batch)[0]
        f = tp.frnn_gather(xp, idxs, len2)
        f = torch.cat([f, f[:, :, 0:1, :] - f], dim=-1)
        f = self.nn(f.view(-1, f.shape[-1])).view(B, N, self.k + 1, -1)
        mask = (idxs != -1) & mask1.unsqueeze(-1)
        if self.aggr_type == "mean":
            f = torch.where(mask[..., None], f, torch.zeros_like(f))
            f = f.sum(dim=2) / mask.sum(dim=2).unsqueeze(-1)
        elif self.aggr_type == "max":
            f = torch.where(mask[..., None], f, torch.full_like(f, -1e2)).max(dim=2)[0]
        elif self.aggr_type == "weighted-sum":
            a = self.attend_nn(f.view(B * N * (self.k + 1), -1)).view(B, N, self.k + 1, -1)
            w = torch.where(mask[..., None], torch.sigmoid(a), torch.zeros_like(a))
            w = w / torch.clamp(w.sum(dim=2, keepdim=True), min=1e-3)
            f = (f * w).sum(dim=2)
        else:
            a = self.attend_nn(f.view(B * N * (self.k + 1), -1)).view(B, N, self.k + 1, -1)
            a = torch.where(mask[..., None], a, torch.full_like(a, -5e2))
            f = (f * F.softmax(a, dim=2)).sum(dim=2)
        return f[mask1]

    def _sparse(self, x, pos, batch):
        row, col = po.knn_ball_group(pos, pos, batch, batch, operation="knn", knn=self.k, radius=self.r,
                                     return_dense=False, accel_knn=self.use_fast_knn)
        # edge_index = flip([row, col]) = [col ; row]: source j = neighbour, target i = query
        i, j = row, col
        msg = self.nn(torch.cat([x[i], x[j] - x[i]], dim=-1))
        n = x.shape[0]
        if self.aggr_type == "max":
            return tp.scatter_max(msg, i, 0, None, n)[0]
        return tp.scatter_add(msg * tp.segment_softmax(self.attend_nn(msg), i), i, 0)


class DynamicEdgeConv(nn.Module):
    """dgcnn.py:16-95: the graph is rebuilt from the input rows x (= [x, pos] in the two layers below) with
    knn_ball_group_pytorch3d on D-dimensional points (:80-83: fast kNN with the 0.25 fallback radius, or ball-group);
    edge_index = flip([row, col]) (:84-85): source j = neighbour, target i = query; message nn([x_i, x_j - x_i]) (:91-92);
    MessagePassing's own aggregation over the target (max / mean / add; empty targets give 0)."""

    def __init__(self, nn, k=None, radius=None, aggr="max", num_workers=1, use_knn=True, **kwargs):
        super().__init__()
        self.nn, self.k, self.r, self.aggr, self.use_knn = nn, k, radius, aggr, use_knn

    def edge_conv(self, x, batch):
        if self.use_knn:
            row, col = po.knn_ball_group(x, x, batch, batch, operation="knn", knn=self.k)
        else:
            row, col = po.knn_ball_group(x, x, batch, batch, operation="ball-group", radius=self.r)
        i, j = row, col
        msg = self.nn(torch.cat([x[i], x[j] - x[i]], dim=-1))
        n = x.shape[0]
        if self.aggr == "max":
            return tp.scatter_max(msg, i, 0, None, n)[0]
        if self.aggr == "mean":
            return tp.scatter_mean(msg, i, 0, None, n)
        assert self.aggr in ("add", "sum")
        return tp.scatter_add(msg, i, 0, None, n)


class DGCNNLayer(DynamicEdgeConv):
    """dgcnn.py:98-111."""

    def __init__(self, nn, k, aggr="max", num_workers=1, with_xyz=False, **kwargs):
        super().__init__(nn, k, None, aggr, num_workers)
        self.with_xyz = with_xyz

    def forward(self, x, pos, batch, point2curveidx=None, **kwargs):
        return self.edge_conv(_cat_xyz(x, pos, self.with_xyz), batch), pos, batch, point2curveidx


class DGCNNLayerRadius(DynamicEdgeConv):
    """dgcnn.py:114-127."""

    def __init__(self, nn, r, aggr="max", num_workers=1, with_xyz=False, **kwargs):
        super().__init__(nn, None, r, aggr, num_workers, use_knn=False)
        self.with_xyz = with_xyz

    def forward(self, x, pos, batch, point2curveidx=None, **kwargs):
        return self.edge_conv(_cat_xyz(x, pos, self.with_xyz), batch), pos, batch, point2curveidx


# ----------------------------------------------------------------------------------
# mlp.py / skip_connect.py
# ----------------------------------------------------------------------------------
class SharedMLP(nn.Module):
    """mlp.py:5-22."""

    def __init__(self, dims, use_bias=False, with_xyz=False, act="leaky_relu", **kwargs):
        super().__init__()
        from .tp_nn import MLP
        self.mlp = MLP(dims, dropout=kwargs.get("dropout", 0.0), norm=kwargs.get("norm", "batch_norm"),
                       plain_last=kwargs.get("plain_last", True), act=act, bias=use_bias)
        self.with_xyz = with_xyz

    def forward(self, x, pos, batch, point2curveidx=None, **kwargs):
        return self.mlp(_cat_xyz(x, pos, self.with_xyz)), pos, batch, point2curveidx


class SkipConnect(nn.Module):
    """skip_connect.py:6-15."""

    def __init__(self, nn, num_skips=1):
        super().__init__()
        self.num_skips, self.nn = num_skips, nn

    def forward(self, xs, pos, batch, point2curveidx=None, **kwargs):
        return self.nn(torch.cat(xs, dim=1)), pos, batch, point2curveidx
