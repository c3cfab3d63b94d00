"""Step modules with the reference's class names, constructor arguments, forward signatures and
parameter names (reference src/models/modules/{fast_conv1d,pointnet2,point_conv,dgcnn,mlp,skip_connect}.py),
so `ModelBase` configs and checkpoints drop in.  Bodies call the sm_100a kernels through `ops`;
there is no torch arithmetic on the per-scan path.

`module.eval()`: the inference path - BatchNorm running statistics folded into the weights, activations as bf16 planes.
`module.train()`: every forward dispatches to train_steps.py (batch-statistics BatchNorm, differentiable ops with
hand-written backward kernels, autograd.py).
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn as nn

from . import ops
from . import train_steps as T
from .nn import (DEFAULT_DRAWS, MLP, Draws, PreparedLinear, cat_features, fold_batchnorm, gemm_input, get_cache,
                 mlp_to_features, n_planes)


class _ChainCache:
    """ChainWeights (ops) of an edge MLP's layers after its first + the attention MLP's layers, rebuilt when the folded
    weights change; None when the fused kernels do not cover the shapes (the unfused chain of launches runs then)."""

    def __init__(self):
        self.key, self.chain = None, None

    def get(self, layers, extra=(), dense=False):
        npl = n_planes()
        key = (id(layers), npl, id(extra) if extra else 0, ops.FUSED_EDGE_STEPS, dense)
        if self.key != key:
            rest = list(layers[1:]) + list(extra)
            ok = (npl > 0 and len(rest) >= 1 and
                  ops.ChainWeights.supported(layers[0].out_features, [l.out_features for l in rest], dense) and
                  ops.fused_profitable(layers[0].out_features, npl, bool(extra), dense and len(rest) == 1))
            self.chain = ops.ChainWeights([(l.w, l.b, l.act, l.slope) for l in rest], npl) if ok else None
            self.key, self._keep = key, (layers, extra)
        return self.chain


def _cat_xyz(x, pos, with_xyz, cache=None):
    """cat[x, pos] as the input of a Linear (activation planes on the tensor-core modes)."""
    if not with_xyz:
        return gemm_input(x, cache)
    return gemm_input(pos, cache) if x is None else cat_features([x, pos], cache=cache)


# ----------------------------------------------------------------------------------------------
# fast_conv1d.py
# ----------------------------------------------------------------------------------------------
class SymmetricConv1d(nn.Module):
    """Parameter holder of fast_conv1d.py:148-187: weight (C_out, C_in, k//2+1) mirrored around tap 0."""

    def __init__(self, in_channels, out_channels, kernel_size, bias=True):
        super().__init__()
        self.weight = nn.Parameter(torch.empty(out_channels, in_channels, kernel_size))
        self.bias = nn.Parameter(torch.empty(out_channels)) if bias else None
        nn.init.kaiming_uniform_(self.weight, a=5 ** 0.5)
        if bias:
            bound = 1.0 / (in_channels * kernel_size) ** 0.5
            nn.init.uniform_(self.bias, -bound, bound)

    def full_weight(self):
        w = self.weight
        if w.shape[2] > 1:
            w = torch.cat([torch.flip(w[:, :, 1:], dims=[2]), w], dim=2)   # fast_conv1d.py:175-178
        return w


class _CurveConvBase(nn.Module):
    def __init__(self, feat_dims, kernel_size, bias, with_xyz, with_diff, diff_every_layer):
        super().__init__()
        self.kernel_size, self.feat_dims = kernel_size, list(feat_dims)
        self.with_xyz, self.with_diff = with_xyz, with_diff
        ku = kernel_size // 2 + 1
        convs, norms = [], []
        for i in range(1, len(feat_dims)):
            doubled = with_diff and (diff_every_layer or i == 1)
            convs.append(SymmetricConv1d(feat_dims[i - 1] * (2 if doubled else 1), feat_dims[i], ku, bias=bias))
            norms.append(nn.BatchNorm1d(feat_dims[i]))
        self.conv_modules = nn.ModuleList(convs)
        self.norm_modules = nn.ModuleList(norms)
        self._folded, self._folded_key, self._folded_epoch = None, None, None

    def _layers(self):
        """[(w (k, C_in, C_out), b (C_out))]: mirrored kernel with BatchNorm(eval) folded."""
        if self.training:
            raise RuntimeError("_layers() is the folded inference form; training goes through train_steps.curve_conv_*")
        from . import nn as _nn
        epoch = _nn._EPOCH[0]
        if epoch is not None and self._folded is not None and self._folded_epoch == epoch:
            return self._folded
        self._folded_epoch = epoch
        key = tuple(c.weight._version + n.weight._version + n.bias._version + n.running_mean._version + n.running_var._version +
                    (c.bias._version if c.bias is not None else 0)
                    for c, n in zip(self.conv_modules, self.norm_modules)) + (self.conv_modules[0].weight.data_ptr(),)
        if self._folded is None or key != self._folded_key:
            out = []
            for conv, norm in zip(self.conv_modules, self.norm_modules):
                w, b = fold_batchnorm(conv.full_weight(), conv.bias, norm)
                out.append((w.permute(2, 1, 0).contiguous(), b))
            self._folded, self._folded_key = out, key
        return self._folded


class SymmetricCurve1DConvV2(_CurveConvBase):
    """fast_conv1d.py:11-75.  One zero-padded signal for the whole stack ((k//2)*L pad rows before every
    curve and after the last); conv + BN + LeakyReLU run over pad rows too.  Here: layer 1 reads the
    points through the pad-row map, middle layers work on the padded signal, the last layer writes
    valid rows only."""

    def __init__(self, feat_dims=(64, 64, 128), kernel_size=5, bias=True, device=None, dtype=None,
                 with_xyz=False, with_diff=False):
        super().__init__(feat_dims, kernel_size, bias, with_xyz, with_diff, diff_every_layer=False)

    def forward(self, x, pos, batch, point2curveidx, **kwargs):
        if self.training:
            return T.curve_conv_v2(self, x, pos, batch, point2curveidx, **kwargs)
        topo = get_cache(kwargs).topology(batch, point2curveidx)
        n = pos.shape[0]
        feats = ops.curve_feat_prep(x, pos, self.with_xyz, self.with_diff, topo.curve_ord)
        layers = self._layers()
        k = self.kernel_size
        pad = (k // 2) * len(layers) if k > 1 else 0
        row2pt = ops.curve_row_map(topo, pad)
        rows = row2pt.shape[0]
        # tensor-core layers: from the first layer on whose successors are all GEMM-sized, the k x C_in x C_out contraction
        # of a layer is ONE tcgen05 GEMM over the padded signal (a tap's row shift is a TMA row coordinate), activations
        # stay in bf16 planes between the layers and the last layer stores the curve points only (row map in the
        # epilogue).  Narrow first layers (C_in = 8 at the input of the network) stay on the FP32-pipe tap kernel, which
        # reads the points through the pad-row map.
        tc_from = len(layers)
        if n_planes() > 0 and rows > 0:
            while tc_from > 0 and self._tc_layer(layers[tc_from - 1][0], tc_from - 1 == 0):
                tc_from -= 1
        sig, sigp = feats, None
        for li, (w, b) in enumerate(layers):
            first, last = li == 0, li == len(layers) - 1
            if li >= tc_from:
                if first:
                    sigp = ops.pack_planes([ops.gather_rows_i32(feats, row2pt)], n_planes())   # zero rows at the curve boundaries
                elif sigp is None:
                    sigp = ops.pack_planes([sig], n_planes())
                if last:
                    # the step output feeds Linear layers: fp32 rows and their planes from the same epilogue
                    sig, outp = ops.conv_planes(sigp, w, b, ops.ACT_LEAKY, 0.01, tw=self._tc_weight(li, w), out="both",
                                                out_map=row2pt, out_rows=n)
                    get_cache(kwargs).put_planes(sig, outp)
                else:
                    sigp = ops.conv_planes(sigp, w, b, ops.ACT_LEAKY, 0.01, tw=self._tc_weight(li, w), out="planes")
                continue
            sig = ops.curve_conv_layer(sig, w, b, row2pt, rows, n if last else rows, k, first, last)
        return sig, pos, batch, point2curveidx

    @staticmethod
    def _tc_layer(w, first):
        return w.shape[2] >= 16 and w.shape[1] >= (128 if first else 16)

    def _tc_weight(self, li, w):
        """bf16 planes + TMA descriptors of a layer's folded weight, rebuilt when the folded weights change."""
        cache = self.__dict__.setdefault("_tc_cache", {})
        hit = cache.get(li)
        if hit is None or hit[0] is not w:
            taps, cin, cout = w.shape
            tw = ops.TcWeight(w.permute(2, 0, 1).reshape(cout, taps * cin).contiguous(), [cin] * taps if taps > 1 else None)
            hit = (w, tw)
            cache[li] = hit
        return hit[1]


class SymmetricCurve1DConvFastV1(_CurveConvBase):
    """fast_conv1d.py:78-145.  Every layer: feature differences (C_in doubles each layer), fresh k//2
    zero rows between curves, conv, BN over the valid rows, LeakyReLU."""

    def __init__(self, feat_dims=(64, 64, 128), kernel_size=5, bias=True, device=None, dtype=None,
                 with_xyz=False, with_diff=False):
        super().__init__(feat_dims, kernel_size, bias, with_xyz, with_diff, diff_every_layer=True)

    def forward(self, x, pos, batch, point2curveidx, **kwargs):
        if self.training:
            return T.curve_conv_v1(self, x, pos, batch, point2curveidx, **kwargs)
        topo = get_cache(kwargs).topology(batch, point2curveidx)
        n = pos.shape[0]
        k = self.kernel_size
        pad = k // 2 if k > 1 else 0
        row2pt = ops.curve_row_map(topo, pad)
        rows = row2pt.shape[0]
        for li, (w, b) in enumerate(self._layers()):
            feats = ops.curve_feat_prep(x, pos, self.with_xyz and li == 0, self.with_diff, topo.curve_ord)
            x = ops.curve_conv_layer(feats, w, b, row2pt, rows, n, k, True, True)
        return x, pos, batch, point2curveidx


# ----------------------------------------------------------------------------------------------
# point_conv.py
# ----------------------------------------------------------------------------------------------
class PointNetConv2(nn.Module):
    """point_conv.py:12-93 over a CSR edge list (target-sorted, as every reference call site builds it).
    message  = local_nn(cat[x_j, (pos_j - pos_i) / normalize_radius])            (:60-69)
    aggregate = max | mean | sum sigmoid(attend_nn(m)) m | sum softmax_i(attend_nn(m)) m   (:71-93)
    The first Linear of local_nn is decomposed: its x_j part is one GEMM over the SOURCE points
    (N_src rows instead of E), its 3 positional columns are added per edge together with the folded
    BatchNorm shift and the activation (ccn_edge_layer1); the remaining layers run per edge."""

    def __init__(self, local_nn=None, global_nn=None, attend_nn=None, add_self_loops=True, aggr_type="max",
                 normalize_radius=None, **kwargs):
        super().__init__()
        assert aggr_type in ["max", "attend", "mean", "weighted-sum"]
        if add_self_loops:
            raise NotImplementedError("every reference call site passes add_self_loops=False")
        if local_nn is None:
            raise NotImplementedError("PointNetConv2 without local_nn does not occur on the path")
        self.local_nn, self.global_nn, self.attend_nn = local_nn, global_nn, attend_nn
        self.aggr_type, self.normalize_radius = aggr_type, normalize_radius
        self._dec, self._dec_src = None, None
        self._fused = _ChainCache()

    def decomposed(self, cx):
        """First Linear of local_nn split into its feature part (a PreparedLinear over the cx source channels, None when
        cx = 0) and its positional part wp (3, C1); weight preparation, once per weight version."""
        layers = self.local_nn.layers()
        l0 = layers[0]
        assert l0.in_features == cx + 3, "local_nn input width %d != %d feature + 3 position columns" % (l0.in_features, cx)
        if self._dec is None or self._dec_src is not layers:
            lin_x = PreparedLinear(l0.w[:, :cx].contiguous(), None) if cx > 0 else None
            self._dec, self._dec_src = (lin_x, l0.w[:, cx:].t().contiguous()), layers
        return self._dec

    def forward_edges(self, x_src, pos_src, pos_dst, edges: ops.EdgeCSR, cache=None):
        if self.training:
            return T.pointnet_conv(self, x_src, pos_src, pos_dst, edges)
        x_src = gemm_input(x_src, cache)
        layers = self.local_nn.layers()
        l0 = layers[0]
        cx = 0 if x_src is None else x_src.shape[1]
        assert l0.in_features == cx + 3, "local_nn input width %d != %d feature + 3 position columns" % (l0.in_features, cx)
        lin_x, wp = self.decomposed(cx)                        # x part (C1, cx) and positional part (3, C1)
        H = lin_x(x_src) if lin_x is not None else None
        if edges.max_degree <= ops.FUSED_MAX_DEGREE and (H is None or torch.is_tensor(H)):
            # gather -> layer 1 -> remaining layers [-> attention MLP] -> aggregate in ONE kernel (csrc/edge_fused.cu)
            att = self.aggr_type in ("attend", "weighted-sum")
            extra = self.attend_nn.layers() if att else ()
            chain = self._fused.get(layers, extra) if len(layers) >= 2 else None
            if chain is not None:
                out = ops.pointconv_fused(edges, chain, self.aggr_type, len(layers) - 2 if att else -1, l0.act, l0.slope,
                                          l0.out_features, H=H, wp=wp, bias0=l0.b, pos_src=pos_src, pos_dst=pos_dst,
                                          radius=self.normalize_radius, npl=n_planes(), cache=cache)
                return self.global_nn(out) if self.global_nn is not None else out
        npl = n_planes() if len(layers) > 1 else 0
        msg = ops.edge_layer1(H, wp, l0.b, pos_src, pos_dst, edges, self.normalize_radius, l0.act, l0.slope, npl)
        if self.aggr_type in ("max", "mean"):
            msg = self.local_nn.forward_from(msg, 1)
            out = ops.seg_aggregate(msg, None, edges.row_ptr, edges.nq, self.aggr_type, n_planes(), cache)
        else:
            # the message is needed twice: fp32 for the aggregation, planes as the input of attend_nn
            both = npl > 0
            msg = self.local_nn.forward_from(msg, 1, "both" if both else "f32")
            msg, msg_in = msg if both else (msg, msg)
            att = self.attend_nn(msg_in)
            out = ops.seg_aggregate(msg, att, edges.row_ptr, edges.nq, self.aggr_type, n_planes(), cache)
        if self.global_nn is not None:
            out = self.global_nn(out)
        return out


# ----------------------------------------------------------------------------------------------
# pointnet2.py
# ----------------------------------------------------------------------------------------------
def _subset(x, pos, batch, curve, idx, cache=None):
    posq, batchq, curveq = ops.subset_gather(pos, batch, curve, idx, n_planes(), cache)
    return (None if x is None else ops.gather_rows(x, idx)), posq, batchq, curveq


class SAModule(nn.Module):
    """pointnet2.py:33-78: down-sample ('random' | 'curve-fps' | 'voxel' | 'fps'), group with FRNN kNN
    (use_fast_knn) or ball_query K=128, PointNetConv2."""

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

    def forward(self, x, pos, batch, point2curveidx=None, draws: Draws = DEFAULT_DRAWS, **kwargs):
        if self.training:
            return T.sa_module(self, x, pos, batch, point2curveidx, draws=draws, **kwargs)
        cache = get_cache(kwargs)
        sptr = cache.sample_ptr(batch)
        n = pos.shape[0]
        if self.downsample_type == "random":
            perm = draws.randperm(n)[:int(n * self.ratio)].sort()[0]         # pointnet2.py:49-52
            idx = perm.to(pos.device)
            sptr_q = None
        elif self.downsample_type == "curve-fps":
            topo = cache.topology(batch, point2curveidx)
            idx, sptr_q = ops.curve_fps(pos, topo, self.curve_fps_arclen, draws.curve_phase())
        elif self.downsample_type == "voxel":
            idx, sptr_q = ops.voxel_fps(pos, batch, self.voxel_size, draws.voxel_noise(n), sptr.n_samples)
        else:
            idx, sptr_q = fps_sorted(pos, sptr, self.ratio, draws)
        xq, posq, batchq, curveq = _subset(x, pos, batch, point2curveidx, idx, cache)
        if sptr_q is not None:
            cache.put_sample_ptr(batchq, sptr_q)
        if self.use_fast_knn:
            nbr, _, cnt = ops.frnn_knn(posq, batchq, pos, batch, sptr, self.knn, self.r, want_cnt=True, cache=cache)
        else:
            nbr, cnt = ops.ball_query(posq, batchq, pos, sptr, self.r, 128)
        edges = ops.dense_to_csr(nbr, cnt)
        out = self.conv.forward_edges(x, pos, posq, edges, cache)
        return out, posq, batchq, curveq


def fps_sorted(pos, sptr: ops.SamplePtr, ratio, draws: Draws):
    """fps_pytorch3d (point_ops.py:57-70): K = ceil(len*ratio) in f32, random start per cloud, sorted result."""
    lens = sptr.lengths
    ksel = ops.fps_count(lens, ratio)
    starts = [draws.fps_start(l) for l in lens]
    idx = ops.fps(pos, sptr, ksel, starts)
    off, host = 0, [0]
    for k, l in zip(ksel, lens):
        off += min(k, l)
        host.append(off)
    dev = torch.tensor(host, dtype=torch.int32).to(pos.device, non_blocking=True)
    return idx, ops.SamplePtr(dev, host)


class GlobalSAModule(nn.Module):
    """pointnet2.py:81-116: nn(cat[x,pos]) then per-sample max (or mean)."""

    def __init__(self, nn, **kwargs):
        super().__init__()
        self.nn = nn
        self.pooling = kwargs.get("pooling", "max")
        if self.pooling not in ("max", "mean"):
            raise NotImplementedError(self.pooling)

    def forward(self, x, pos, batch, point2curveidx=None, **kwargs):
        if self.training:
            return T.global_sa_module(self, x, pos, batch, point2curveidx, **kwargs)
        sptr = get_cache(kwargs).sample_ptr(batch)
        h = self.nn(cat_features([x, pos], cache=get_cache(kwargs)))
        out = ops.seg_aggregate(h, None, sptr.dev, sptr.n_samples, self.pooling)
        first = torch.tensor(sptr.host[:-1], dtype=torch.int64).to(pos.device)
        _, posf, batchf, curvef = _subset(None, pos, batch, point2curveidx, first)
        return out, posf, batchf, curvef


def _fp_concat(x, x_skip, pos_skip, with_xyz, cache=None):
    """pointnet2.py:132-139 / :194-201."""
    parts = [x]
    if x_skip is not None:
        parts.append(x_skip)
    if with_xyz:
        parts.append(pos_skip[:, :3])
    return cat_features(parts, cache=cache) if len(parts) > 1 else gemm_input(x, cache)


class FPModule(nn.Module):
    """pointnet2.py:119-143: exact 3-NN inverse-d2 interpolation onto the skip level, concat, MLP."""

    def __init__(self, k, nn, with_xyz=False):
        super().__init__()
        self.k, self.nn, self.with_xyz = k, nn, with_xyz

    def forward(self, x, pos, batch, x_skip, pos_skip, batch_skip, point2curveidx=None,
                point2curveidx_skip=None, **kwargs):
        if self.training:
            return T.fp_module(self, x, pos, batch, x_skip, pos_skip, batch_skip, point2curveidx,
                               point2curveidx_skip, **kwargs)
        cache = get_cache(kwargs)
        sptr = cache.sample_ptr(batch)
        nbr, d2 = ops.knn_exact(pos_skip, batch_skip, pos, sptr, self.k, zero_fill=True, want_d2=True, batch2=batch)
        y = ops.interp_blend(x, nbr, d2, from_d2=True, npl=n_planes(), cache=cache)
        y = mlp_to_features(self.nn, _fp_concat(y, x_skip, pos_skip, self.with_xyz, cache), cache)
        return y, pos_skip, batch_skip, point2curveidx_skip


class CurveSAModule(nn.Module):
    """pointnet2.py:146-181: CurveFPS (or FPS) -> radius_1d_group_subset -> PointNetConv2; the sampled
    indices are the 6th output (consumed by fp-geo)."""

    def __init__(self, ratio, r, nn, curve_fps_arclen=None, use_curve_fps=False, global_nn=None, attend_nn=None,
                 with_xyz=False, aggr_type="max", normalize_radius=False, **kwargs):
        super().__init__()
        self.ratio, self.r = ratio, r
        self.curve_fps_arclen, self.use_curve_fps, self.with_xyz = curve_fps_arclen, use_curve_fps, with_xyz
        self.normalize_radius = r if normalize_radius else None
        self.conv = PointNetConv2(nn, add_self_loops=False, global_nn=global_nn, aggr_type=aggr_type,
                                  attend_nn=attend_nn, normalize_radius=self.normalize_radius)

    def forward(self, x, pos, batch, point2curveidx, draws: Draws = DEFAULT_DRAWS, **kwargs):
        if self.training:
            return T.curve_sa_module(self, x, pos, batch, point2curveidx, draws=draws, **kwargs)
        cache = get_cache(kwargs)
        topo = cache.topology(batch, point2curveidx)
        x = _cat_xyz(x, pos, self.with_xyz, cache)
        if self.use_curve_fps:
            idx, sptr_q = ops.curve_fps(pos, topo, self.curve_fps_arclen, draws.curve_phase())
        else:
            idx, sptr_q = fps_sorted(pos, topo.sptr, self.ratio, draws)
        edges = ops.radius_1d_group(pos, topo, idx, self.r)
        _, posq, batchq, curveq = _subset(None, pos, batch, point2curveidx, idx, cache)
        cache.put_sample_ptr(batchq, sptr_q)
        out = self.conv.forward_edges(x, pos, posq, edges, cache)
        return out, posq, batchq, curveq, None, idx


class CurveFPModule(FPModule):
    """pointnet2.py:184-205: interpolation along the curve onto the fine points."""

    def forward(self, x, idx, x_skip, pos_skip, batch_skip, point2curveidx_skip=None, **kwargs):
        if self.training:
            return T.curve_fp_module(self, x, idx, x_skip, pos_skip, batch_skip, point2curveidx_skip, **kwargs)
        cache = get_cache(kwargs)
        topo = cache.topology(batch_skip, point2curveidx_skip)
        nbr, wgt = ops.knn_1d_superset(pos_skip, topo.curve_ord, idx, self.k)
        y = ops.interp_blend(x, nbr, wgt, from_d2=False, npl=n_planes(), cache=cache)
        y = mlp_to_features(self.nn, _fp_concat(y, x_skip, pos_skip, self.with_xyz, cache), cache)
        return y, pos_skip, batch_skip, point2curveidx_skip


# ----------------------------------------------------------------------------------------------
# dgcnn.py (StaticEdgeConv / SGCNNLayer; DynamicEdgeConv / DGCNNLayer / DGCNNLayerRadius below them)
# ----------------------------------------------------------------------------------------------
class SGCNNLayer(nn.Module):
    """dgcnn.py:130-266.  kNN graph in xyz space.
    dense path  (forward_fast :158-207): FRNN idx (N,K) + self loop; rows [x_j, x_i - x_j] (x_j = 0 for
                missing neighbours); MLP; max with missing rows at -100.
    sparse path (forward_slow :209-246): edge list from FRNN or exact kNN; nn([x_i, x_j - x_i]); scatter max
                or softmax attention.
    In both, the first Linear is decomposed into two per-POINT products P, Q (one GEMM with 2*C1
    outputs over N rows instead of N*(K+1) rows of width 2C); the per-edge first layer is then
    act(P[j] + Q[i]) (ccn_sgcnn_layer1 / ccn_edge_pair_layer1)."""

    def __init__(self, nn, k, aggr="max", r=1.0, num_workers=1, with_xyz=False, attend_nn=None, aggr_type="max",
                 use_sparse_feat_agg=False, use_fast_knn=True, **kwargs):
        super().__init__()
        assert aggr_type in ["max", "attend", "mean", "weighted-sum"]
        self.nn, self.k, self.r, self.with_xyz = nn, k, r, with_xyz
        self.attend_nn, self.aggr_type = attend_nn, aggr_type
        self.use_fast_knn, self.use_sparse_feat_agg = use_fast_knn, use_sparse_feat_agg
        self._dec, self._dec_src = None, None
        self._fused = _ChainCache()

    def _decomposed(self):
        layers = self.nn.layers()
        if self._dec is None or self._dec_src is not layers:
            l0 = layers[0]
            c = l0.in_features // 2
            wa, wb = l0.w[:, :c], l0.w[:, c:]
            zero = torch.zeros_like(l0.b)
            if self.use_sparse_feat_agg:     # [x_i, x_j - x_i]: P (per source j) = Wb, Q (per target i) = Wa - Wb
                wpq, bpq = torch.cat([wb, wa - wb], dim=0), torch.cat([zero, l0.b])
            else:                            # [x_j, x_i - x_j]: P (per neighbour j) = Wa - Wb, Q (per point i) = Wb
                wpq, bpq = torch.cat([wa - wb, wb], dim=0), torch.cat([zero, l0.b])
            self._dec, self._dec_src = (PreparedLinear(wpq.contiguous(), bpq.contiguous()), l0.act, l0.slope), layers
        return self._dec

    def forward(self, x, pos, batch, point2curveidx=None, **kwargs):
        if self.training:
            return T.sgcnn_layer(self, x, pos, batch, point2curveidx, **kwargs)
        cache = get_cache(kwargs)
        sptr = cache.sample_ptr(batch)
        x = _cat_xyz(x, pos, self.with_xyz, cache)
        lin_pq, act0, slope0 = self._decomposed()
        PQ = lin_pq(x)
        if not self.use_sparse_feat_agg and self.aggr_type in ("mean", "attend", "weighted-sum"):
            # dgcnn.py:182-203: mean / sigmoid-weighted (renormalised by clamp(sum w, 1e-3)) / softmax over the VALID rows of the
            # K+1 table (invalid rows are zeroed / weight 0 / pushed to -5e2, i.e. weight exp(-500 - max) = 0): the same
            # reductions over the edge list [self loop, found neighbours]
            nbr, _, cnt = ops.frnn_knn(pos, batch, pos, batch, sptr, self.k, self.r, want_cnt=True, cache=cache)
            n = pos.shape[0]
            ext = torch.cat([torch.arange(n, dtype=torch.int32, device=pos.device).view(n, 1), nbr], dim=1).contiguous()
            cnt = cnt.clone()
            cnt[:n] += 1
            edges = ops.dense_to_csr(ext, cnt)
            npl = n_planes() if self.nn.num_layers > 1 else 0
            msg = ops.edge_pair_layer1(PQ, edges, act0, slope0, npl)
            if self.aggr_type == "mean":
                msg = self.nn.forward_from(msg, 1)
                return ops.seg_aggregate(msg, None, edges.row_ptr, edges.nq, "mean", n_planes(), cache), pos, batch, point2curveidx
            both = npl > 0
            msg = self.nn.forward_from(msg, 1, "both" if both else "f32")
            msg, msg_in = msg if both else (msg, msg)
            mode = "attend" if self.aggr_type == "attend" else "weighted-sum-norm"
            out = ops.seg_aggregate(msg, self.attend_nn(msg_in), edges.row_ptr, edges.nq, mode, n_planes(), cache)
            return out, pos, batch, point2curveidx
        if not self.use_sparse_feat_agg:
            nbr, _, _ = ops.frnn_knn(pos, batch, pos, batch, sptr, self.k, self.r, cache=cache)
            chain = self._fused.get(self.nn.layers(), dense=True) if self.k + 1 <= 128 else None
            if chain is not None:
                # gather -> layer 1 -> remaining layers -> masked max in ONE kernel (csrc/edge_fused.cu)
                return ops.sgcnn_fused(PQ, nbr, act0, slope0, chain, n_planes(), cache), pos, batch, point2curveidx
            h = ops.sgcnn_layer1(PQ, nbr, act0, slope0, n_planes() if self.nn.num_layers > 1 else 0)
            h = self.nn.forward_from(h, 1)
            return ops.group_max_dense(h, nbr, n_planes(), cache), pos, batch, point2curveidx
        if self.use_fast_knn:
            nbr, _, cnt = ops.frnn_knn(pos, batch, pos, batch, sptr, self.k, self.r, want_cnt=True, cache=cache)
        else:
            nbr, _, cnt = ops.knn_exact(pos, batch, pos, sptr, self.k, zero_fill=True, want_d2=False, want_cnt=True)
        edges = ops.dense_to_csr(nbr, cnt)
        npl = n_planes() if self.nn.num_layers > 1 else 0
        sparse_mode = "max" if self.aggr_type == "max" else "attend"     # dgcnn.py:240-244: every other aggr_type takes the softmax branch
        if edges.max_degree <= ops.FUSED_MAX_DEGREE and (sparse_mode == "max" or self.attend_nn is not None):
            layers = self.nn.layers()
            extra = self.attend_nn.layers() if sparse_mode == "attend" else ()
            chain = self._fused.get(layers, extra) if len(layers) >= 2 else None
            if chain is not None:
                out = ops.pointconv_fused(edges, chain, sparse_mode, len(layers) - 2 if extra else -1, act0, slope0,
                                          layers[0].out_features, PQ=PQ, npl=n_planes(), cache=cache)
                return out, pos, batch, point2curveidx
        msg = ops.edge_pair_layer1(PQ, edges, act0, slope0, npl)
        if self.aggr_type == "max":
            msg = self.nn.forward_from(msg, 1)
            out = ops.seg_aggregate(msg, None, edges.row_ptr, edges.nq, "max", n_planes(), cache)
        else:
            if self.attend_nn is None:
                raise UnboundLocalError("attend_nn")                  # the reference dereferences self.attend_nn (dgcnn.py:242)
            both = npl > 0
            msg = self.nn.forward_from(msg, 1, "both" if both else "f32")
            msg, msg_in = msg if both else (msg, msg)
            out = ops.seg_aggregate(msg, self.attend_nn(msg_in), edges.row_ptr, edges.nq, "attend", n_planes(), cache)
        return out, pos, batch, point2curveidx


class _DynamicEdgeConv(nn.Module):
    """DynamicEdgeConv (dgcnn.py:16-95): EdgeConv whose graph is rebuilt from the layer's own INPUT rows - here [x, pos] -
    by a neighbour search in that feature space (knn_ball_group_pytorch3d on D-dimensional rows, point_ops.py:73-111):
    use_knn: fast_knn, i.e. the fixed-radius kNN contract with the 0.25 the reference falls back to when no radius is given
    (point_ops.py:84-87); else ball_query with K = 128.  message = nn([x_i, x_j - x_i]) (:91-92), default aggregation of
    MessagePassing over the target's edges (max / mean / add).  Reached by no shipped configuration: the search is brute
    force (ccn_feat_neighbours); everything behind it is the sparse SGCNN machinery (first Linear decomposed per point,
    ccn_edge_pair_layer1, MLP, ccn_seg_aggregate)."""

    KNN_FALLBACK_RADIUS = 0.25      # point_ops.py:85-87 "Not setting radius for Fast-KNN!"
    BALL_K = 128                    # point_ops.py:81

    def __init__(self, nn, k=None, radius=None, aggr="max", num_workers=1, use_knn=True, with_xyz=False, **kwargs):
        super().__init__()
        if aggr not in ("max", "mean", "add", "sum"):
            raise NotImplementedError("DynamicEdgeConv aggregation %r" % (aggr,))
        self.nn, self.k, self.r, self.aggr = nn, k, radius, aggr
        self.num_workers, self.use_knn, self.with_xyz = num_workers, use_knn, with_xyz
        self._dec, self._dec_src = None, None

    def _decomposed(self):
        layers = self.nn.layers()
        if self._dec is None or self._dec_src is not layers:
            l0 = layers[0]
            c = l0.in_features // 2
            wa, wb = l0.w[:, :c], l0.w[:, c:]
            # nn([x_i, x_j - x_i]) first Linear: P (per source j) = Wb, Q (per target i) = Wa - Wb (+ bias)
            wpq, bpq = torch.cat([wb, wa - wb], dim=0), torch.cat([torch.zeros_like(l0.b), l0.b])
            self._dec, self._dec_src = (PreparedLinear(wpq.contiguous(), bpq.contiguous()), l0.act, l0.slope), layers
        return self._dec

    def search(self, xin, batch, sptr):
        """-> (idx (n, K) int32 -1 padded, counts): the edge list of dgcnn.py:80-85, grouped by query."""
        if self.use_knn:
            return ops.feat_neighbours(xin, batch, xin, sptr, self.k, self.KNN_FALLBACK_RADIUS, "knn")
        return ops.feat_neighbours(xin, batch, xin, sptr, self.BALL_K, self.r, "ball")

    def forward(self, x, pos, batch, point2curveidx=None, **kwargs):
        if x is None and not self.with_xyz:
            raise TypeError("DynamicEdgeConv needs node features (x is None and with_xyz is False)")   # dgcnn.py:68-72 fails on None
        if self.training:
            return T.dgcnn_layer(self, x, pos, batch, point2curveidx, **kwargs)
        cache = get_cache(kwargs)
        sptr = cache.sample_ptr(batch)
        xin = x if not self.with_xyz else (pos if x is None else ops.concat_cols([x, pos]))
        nbr, cnt = self.search(xin, batch, sptr)
        edges = ops.dense_to_csr(nbr, cnt)
        lin_pq, act0, slope0 = self._decomposed()
        PQ = lin_pq(xin)
        npl = n_planes() if self.nn.num_layers > 1 else 0
        msg = self.nn.forward_from(ops.edge_pair_layer1(PQ, edges, act0, slope0, npl), 1)
        mode = "sum" if self.aggr in ("add", "sum") else self.aggr
        out = ops.seg_aggregate(msg, None, edges.row_ptr, edges.nq, mode, n_planes(), cache)
        return out, pos, batch, point2curveidx


class DGCNNLayer(_DynamicEdgeConv):
    """dgcnn.py:98-111: k nearest rows of [x, pos] (within the fast-kNN fallback radius)."""

    def __init__(self, nn, k, aggr="max", num_workers=1, with_xyz=False, **kwargs):
        super().__init__(nn, k, None, aggr, num_workers, use_knn=True, with_xyz=with_xyz)


class DGCNNLayerRadius(_DynamicEdgeConv):
    """dgcnn.py:114-127: the first 128 rows of [x, pos] inside radius r."""

    def __init__(self, nn, r, aggr="max", num_workers=1, with_xyz=False, **kwargs):
        super().__init__(nn, None, r, aggr, num_workers, use_knn=False, with_xyz=with_xyz)


# ----------------------------------------------------------------------------------------------
# mlp.py / skip_connect.py
# ----------------------------------------------------------------------------------------------
class SharedMLP(nn.Module):
    """mlp.py:5-22."""

    def __init__(self, dims, use_bias=False, with_xyz=False, act="leaky_relu", **kwargs):
        super().__init__()
        self.mlp = MLP(dims, dropout=kwargs.get("dropout", 0.0), norm=kwargs.get("norm", "batch_norm"),
                       plain_last=kwargs.get("plain_last", True), act=act, bias=use_bias)
        self.with_xyz = with_xyz

    def forward(self, x, pos, batch, point2curveidx=None, **kwargs):
        if self.training:
            return T.shared_mlp(self, x, pos, batch, point2curveidx, **kwargs)
        cache = kwargs.get("cache")
        return mlp_to_features(self.mlp, _cat_xyz(x, pos, self.with_xyz, cache), cache), pos, batch, point2curveidx


class SkipConnect(nn.Module):
    """skip_connect.py:6-15."""

    def __init__(self, nn, num_skips=1):
        super().__init__()
        self.num_skips, self.nn = num_skips, nn

    def forward(self, xs, pos, batch, point2curveidx=None, **kwargs):
        if self.training:
            return T.skip_connect(self, xs, pos, batch, point2curveidx, **kwargs)
        cache = kwargs.get("cache")
        return mlp_to_features(self.nn, cat_features(list(xs), cache=cache), cache), pos, batch, point2curveidx
