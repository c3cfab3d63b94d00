"""Training-mode bodies of the step modules (modules.py dispatches here when `module.training`).

Same sampling / grouping kernels as the inference path (they carry no gradient); the feature arithmetic goes
through the differentiable ops of autograd.py with BatchNorm in batch-statistics form, so that
`loss.backward()` reproduces the reference's gradients (src/run/*_seg.py training loops).
"""
from __future__ import annotations

import torch

from . import autograd as A
from . import ops
from .nn import DEFAULT_DRAWS, Draws, get_cache

_LEAKY = (ops.ACT_LEAKY, 0.01)


def cat_xyz(x, pos, with_xyz):
    if not with_xyz:
        return x
    return pos if x is None else A.concat([x, pos])


# ---------------------------------------------------------------------------------------------- fast_conv1d.py
def _conv_params(conv):
    w = conv.full_weight().permute(2, 1, 0)          # (k, cin, cout); torch ops on the weights only
    b = conv.bias if conv.bias is not None else torch.zeros(w.shape[2], device=w.device)
    return w, b


def curve_conv_v2(mod, x, pos, batch, curve, **kwargs):
    """fast_conv1d.py:11-75 in training form: BatchNorm statistics over ALL rows of the padded signal."""
    topo = get_cache(kwargs).topology(batch, curve)
    n = pos.shape[0]
    k = mod.kernel_size
    feats = A.curve_feat_prep(x, pos, mod.with_xyz, mod.with_diff, topo.curve_ord)
    pad = (k // 2) * len(mod.conv_modules) if k > 1 else 0
    row2pt = ops.curve_row_map(topo, pad)
    sig = A.pad_rows(feats, row2pt)
    for conv, norm in zip(mod.conv_modules, mod.norm_modules):
        w, b = _conv_params(conv)
        sig = A.bn_act(A.curve_conv(sig, w, b, k), norm, *_LEAKY)
    return A.unpad_rows(sig, row2pt, n), pos, batch, curve


def curve_conv_v1(mod, x, pos, batch, curve, **kwargs):
    """fast_conv1d.py:78-145 in training form: BatchNorm statistics over the valid rows."""
    topo = get_cache(kwargs).topology(batch, curve)
    n = pos.shape[0]
    k = mod.kernel_size
    row2pt = ops.curve_row_map(topo, k // 2 if k > 1 else 0)
    for li, (conv, norm) in enumerate(zip(mod.conv_modules, mod.norm_modules)):
        feats = A.curve_feat_prep(x, pos, mod.with_xyz and li == 0, mod.with_diff, topo.curve_ord)
        w, b = _conv_params(conv)
        z = A.unpad_rows(A.curve_conv(A.pad_rows(feats, row2pt), w, b, k), row2pt, n)
        x = A.bn_act(z, norm, *_LEAKY)
    return x, pos, batch, curve


# ---------------------------------------------------------------------------------------------- point_conv.py
def _first_hidden(mlp, z):
    if len(mlp.lins) > 1 or not mlp.plain_last:
        return A.hidden_tail(mlp, 0, z, dz_planes=False)       # z comes from a per-edge gather layer, not a GEMM
    return A.dropout(z, mlp.dropout[0], mlp.training)


def pointnet_conv(conv, x_src, pos_src, pos_dst, edges: ops.EdgeCSR):
    """PointNetConv2 (point_conv.py:12-93) over CSR edges; first Linear decomposed as in the inference path."""
    mlp = conv.local_nn
    lin0 = mlp.lins[0]
    cx = 0 if x_src is None else x_src.shape[1]
    W = lin0.weight
    assert W.shape[1] == cx + 3
    H = A.linear(x_src, W[:, :cx], None) if cx > 0 else None
    z = A.edge_layer1(H, W[:, cx:].t(), lin0.bias, pos_src, pos_dst, edges, conv.normalize_radius)
    msg = A.mlp_train(mlp, _first_hidden(mlp, z), first=1)
    if conv.aggr_type in ("max", "mean"):
        out = A.seg_aggregate(msg, None, edges.row_ptr, edges.nq, conv.aggr_type)
    else:
        att = A.mlp_train(conv.attend_nn, msg)
        out = A.seg_aggregate(msg, att, edges.row_ptr, edges.nq, conv.aggr_type)
    if conv.global_nn is not None:
        out = A.mlp_train(conv.global_nn, out)
    return out


# ---------------------------------------------------------------------------------------------- pointnet2.py
def sa_module(mod, x, pos, batch, curve=None, draws: Draws = DEFAULT_DRAWS, **kwargs):
    from .modules import _subset, fps_sorted
    cache = get_cache(kwargs)
    sptr = cache.sample_ptr(batch)
    n = pos.shape[0]
    if mod.downsample_type == "random":
        idx = draws.randperm(n)[:int(n * mod.ratio)].sort()[0].to(pos.device)
        sptr_q = None
    elif mod.downsample_type == "curve-fps":
        idx, sptr_q = ops.curve_fps(pos, cache.topology(batch, curve), mod.curve_fps_arclen, draws.curve_phase())
    elif mod.downsample_type == "voxel":
        idx, sptr_q = ops.voxel_fps(pos, batch, mod.voxel_size, draws.voxel_noise(n), sptr.n_samples)
    else:
        idx, sptr_q = fps_sorted(pos, sptr, mod.ratio, draws)
    _, posq, batchq, curveq = _subset(None, pos, batch, curve, idx)
    if sptr_q is not None:
        cache.put_sample_ptr(batchq, sptr_q)
    if mod.use_fast_knn:
        nbr, _, cnt = ops.frnn_knn(posq, batchq, pos, batch, sptr, mod.knn, mod.r, want_cnt=True, cache=cache)
    else:
        nbr, cnt = ops.ball_query(posq, batchq, pos, sptr, mod.r, 128)
    edges = ops.dense_to_csr(nbr, cnt)
    return pointnet_conv(mod.conv, x, pos, posq, edges), posq, batchq, curveq


def global_sa_module(mod, x, pos, batch, curve=None, **kwargs):
    from .modules import _subset
    sptr = get_cache(kwargs).sample_ptr(batch)
    h = A.mlp_train(mod.nn, A.concat([x, pos]))
    out = A.seg_aggregate(h, None, sptr.dev, sptr.n_samples, mod.pooling)
    first = torch.tensor(sptr.host[:-1], dtype=torch.int64).to(pos.device)
    _, posf, batchf, curvef = _subset(None, pos, batch, curve, first)
    return out, posf, batchf, curvef


def _fp_concat(y, x_skip, pos_skip, with_xyz):
    parts = [y]
    if x_skip is not None:
        parts.append(x_skip)
    if with_xyz:
        parts.append(pos_skip[:, :3])
    return A.concat(parts)


def fp_module(mod, x, pos, batch, x_skip, pos_skip, batch_skip, curve=None, curve_skip=None, **kwargs):
    sptr = get_cache(kwargs).sample_ptr(batch)
    nbr, d2 = ops.knn_exact(pos_skip, batch_skip, pos, sptr, mod.k, zero_fill=True, want_d2=True)
    y = A.interp_blend(x, nbr, d2, True)
    return A.mlp_train(mod.nn, _fp_concat(y, x_skip, pos_skip, mod.with_xyz)), pos_skip, batch_skip, curve_skip


def curve_fp_module(mod, x, idx, x_skip, pos_skip, batch_skip, curve_skip=None, **kwargs):
    topo = get_cache(kwargs).topology(batch_skip, curve_skip)
    nbr, wgt = ops.knn_1d_superset(pos_skip, topo.curve_ord, idx, mod.k)
    y = A.interp_blend(x, nbr, wgt, False)
    return A.mlp_train(mod.nn, _fp_concat(y, x_skip, pos_skip, mod.with_xyz)), pos_skip, batch_skip, curve_skip


def curve_sa_module(mod, x, pos, batch, curve, draws: Draws = DEFAULT_DRAWS, **kwargs):
    from .modules import _subset, fps_sorted
    cache = get_cache(kwargs)
    topo = cache.topology(batch, curve)
    x = cat_xyz(x, pos, mod.with_xyz)
    if mod.use_curve_fps:
        idx, sptr_q = ops.curve_fps(pos, topo, mod.curve_fps_arclen, draws.curve_phase())
    else:
        idx, sptr_q = fps_sorted(pos, topo.sptr, mod.ratio, draws)
    edges = ops.radius_1d_group(pos, topo, idx, mod.r)
    _, posq, batchq, curveq = _subset(None, pos, batch, curve, idx)
    cache.put_sample_ptr(batchq, sptr_q)
    return pointnet_conv(mod.conv, x, pos, posq, edges), posq, batchq, curveq, None, idx


# ---------------------------------------------------------------------------------------------- dgcnn.py
def _pad_nbr(nbr, n_pad):
    if n_pad <= 0:
        return nbr
    return torch.cat([nbr, torch.full((n_pad, nbr.shape[1]), -1, dtype=nbr.dtype, device=nbr.device)], dim=0)


def sgcnn_layer(mod, x, pos, batch, curve=None, **kwargs):
    cache = get_cache(kwargs)
    sptr = cache.sample_ptr(batch)
    x = cat_xyz(x, pos, mod.with_xyz)
    mlp = mod.nn
    lin0 = mlp.lins[0]
    W = lin0.weight
    c = W.shape[1] // 2
    wa, wb = W[:, :c], W[:, c:]
    # decomposition of the first Linear (see modules.SGCNNLayer): torch arithmetic on the WEIGHTS only
    wpq = torch.cat([wb, wa - wb], dim=0) if mod.use_sparse_feat_agg else torch.cat([wa - wb, wb], dim=0)
    bpq = torch.cat([torch.zeros_like(lin0.bias), lin0.bias]) if lin0.bias is not None else None
    n_pad = 0
    if not mod.use_sparse_feat_agg:
        # forward_fast runs its MLP - hence its BatchNorm batch statistics - over the B x N_max padded layout of
        # to_batch_padded (dgcnn.py:170-177): every padding point contributes K+1 all-zero input rows.  They are
        # appended here as zero-feature points without neighbours (BatchNorm statistics do not depend on row order).
        lens = sptr.lengths
        n_pad = len(lens) * max(lens) - sum(lens)
        x = A.append_zero_rows(x, n_pad)
    PQ = A.linear(x, wpq, bpq)
    if not mod.use_sparse_feat_agg:
        if mod.aggr_type == "max":
            nbr, _, _ = ops.frnn_knn(pos, batch, pos, batch, sptr, mod.k, mod.r, cache=cache)
            h = A.mlp_train(mlp, _first_hidden(mlp, A.sgcnn_layer1(PQ, _pad_nbr(nbr, n_pad))), first=1)
            return A.group_max_dense(h, nbr), pos, batch, curve
        # mean / attend / weighted-sum (dgcnn.py:182-203): the MLPs - and their batch statistics - run over ALL rows of the
        # padded (B, N_max, K+1) table; the reduction then takes the valid rows of every point (its self loop and its found
        # neighbours, the leading slots of the FRNN table)
        nbr, _, cnt = ops.frnn_knn(pos, batch, pos, batch, sptr, mod.k, mod.r, want_cnt=True, cache=cache)
        n, G = pos.shape[0], mod.k + 1
        h = A.mlp_train(mlp, _first_hidden(mlp, A.sgcnn_layer1(PQ, _pad_nbr(nbr, n_pad))), first=1)
        att_all = A.mlp_train(mod.attend_nn, h) if mod.aggr_type in ("attend", "weighted-sum") else None
        ext = torch.cat([torch.arange(n, dtype=torch.int32, device=pos.device).view(n, 1), nbr], dim=1).contiguous()
        cnt = cnt.clone()
        cnt[:n] += 1
        edges = ops.dense_to_csr(ext, cnt)
        slot = torch.arange(edges.n_edges, device=pos.device, dtype=torch.int64) - edges.row_ptr[:n].to(torch.int64)[edges.erow.to(torch.int64)]
        rows = edges.erow.to(torch.int64) * G + slot                    # dense row of every valid (point, slot): index plumbing
        msg = A.gather_rows(h, rows)
        if mod.aggr_type == "mean":
            return A.seg_aggregate(msg, None, edges.row_ptr, edges.nq, "mean"), pos, batch, curve
        att = A.gather_rows(att_all, rows)
        mode = "attend" if mod.aggr_type == "attend" else "weighted-sum-norm"
        return A.seg_aggregate(msg, att, edges.row_ptr, edges.nq, mode), pos, batch, curve
    if mod.use_fast_knn:
        nbr, _, cnt = ops.frnn_knn(pos, batch, pos, batch, sptr, mod.k, mod.r, want_cnt=True, cache=cache)
    else:
        nbr, _, cnt = ops.knn_exact(pos, batch, pos, sptr, mod.k, zero_fill=True, want_d2=False, want_cnt=True)
    edges = ops.dense_to_csr(nbr, cnt)
    msg = A.mlp_train(mlp, _first_hidden(mlp, A.edge_pair(PQ, edges)), first=1)
    if mod.aggr_type == "max":
        out = A.seg_aggregate(msg, None, edges.row_ptr, edges.nq, "max")
    else:                                # dgcnn.py:240-244: everything but 'max' takes the softmax branch
        if mod.attend_nn is None:
            raise UnboundLocalError("attend_nn")                  # the reference dereferences self.attend_nn (dgcnn.py:242)
        out = A.seg_aggregate(msg, A.mlp_train(mod.attend_nn, msg), edges.row_ptr, edges.nq, "attend")
    return out, pos, batch, curve


def dgcnn_layer(mod, x, pos, batch, curve=None, **kwargs):
    """DynamicEdgeConv in training form (dgcnn.py:64-92): the graph comes from the detached input rows (index selection
    carries no gradient in the reference either), the message MLP runs in training mode over the edge list."""
    cache = get_cache(kwargs)
    sptr = cache.sample_ptr(batch)
    x = cat_xyz(x, pos, mod.with_xyz)
    nbr, cnt = mod.search(x.detach(), batch, sptr)
    edges = ops.dense_to_csr(nbr, cnt)
    mlp = mod.nn
    lin0 = mlp.lins[0]
    W = lin0.weight
    c = W.shape[1] // 2
    wa, wb = W[:, :c], W[:, c:]
    wpq = torch.cat([wb, wa - wb], dim=0)
    bpq = torch.cat([torch.zeros_like(lin0.bias), lin0.bias]) if lin0.bias is not None else None
    PQ = A.linear(x, wpq, bpq)
    msg = A.mlp_train(mlp, _first_hidden(mlp, A.edge_pair(PQ, edges)), first=1)
    mode = "sum" if mod.aggr in ("add", "sum") else mod.aggr
    return A.seg_aggregate(msg, None, edges.row_ptr, edges.nq, mode), pos, batch, curve


# ---------------------------------------------------------------------------------------------- mlp.py / skip_connect.py
def shared_mlp(mod, x, pos, batch, curve=None, **kwargs):
    return A.mlp_train(mod.mlp, cat_xyz(x, pos, mod.with_xyz)), pos, batch, curve


def skip_connect(mod, xs, pos, batch, curve=None, **kwargs):
    return A.mlp_train(mod.nn, A.concat(list(xs))), pos, batch, curve
