"""Training step: differentiable versions of the path's ops.

Every forward AND backward body is one or two calls into libccn_b200.so (csrc/train_ops.cu and the forward
kernels); torch.autograd is used only as the tape (which backward runs when, summing the gradients of a tensor
that is consumed twice).  The reference gets the same gradients from autograd over cuBLAS / cuDNN /
torch_scatter / pytorch3d (`loss.backward()`, src/run/shapenet_seg.py:186-188, nuscenes_seg.py:229-252).

BatchNorm runs in training mode here (batch statistics, running statistics updated in place), so nothing is
folded into the weights; the first layer of every edge MLP is still applied in its decomposed per-point form.
"""
from __future__ import annotations

from typing import List, Optional

import torch
from torch.autograd import Function

from . import ops
from .native import call

_ACTS = {"relu": (ops.ACT_RELU, 0.0), "leaky_relu": (ops.ACT_LEAKY, 0.01), "leakyrelu": (ops.ACT_LEAKY, 0.01)}


def _precision() -> str:
    from . import nn as _nn
    return _nn.get_precision()


def _zeros(shape, like, dtype=torch.float32):
    return torch.zeros(shape, dtype=dtype, device=like.device)


def _empty(shape, like, dtype=torch.float32):
    return torch.empty(shape, dtype=dtype, device=like.device)


# ----------------------------------------------------------------------------------------------
# Linear
# ----------------------------------------------------------------------------------------------
def linear_wgrad(dy: torch.Tensor, x: torch.Tensor, want_bias: bool):
    dy, lddy = ops._rows(dy)
    x, ldx = ops._rows(x)
    M, N = dy.shape
    K = x.shape[1]
    dW = _zeros((N, K), dy)
    db = _zeros((N,), dy) if want_bias else None
    call("ccn_linear_wgrad", dy, lddy, x, ldx, M, N, K, dW, K, db if db is not None else 0)
    return dW, db


def _tc_wgrad_ok(N: int, K: int) -> bool:
    return N >= 16 and K >= 16


def wgrad_planes(dyp: "ops.Planes", xp: "ops.Planes", row_off: int = 0) -> torch.Tensor:
    """dW (N,K) = dY^T X[. + row_off] on tcgen05 from activation planes (ccn_linear_wgrad_tc)."""
    M, N = dyp.shape
    K = xp.shape[1]
    dW = _zeros((N, K), dyp.buf)
    call("ccn_linear_wgrad_tc", dyp.buf, dyp.kpad, dyp.plane_stride, xp.buf, xp.kpad, xp.plane_stride, M, N, K,
         int(row_off), xp.m, 3 if dyp.n_planes == 2 else 1, dW, K)
    return dW


def col_sum(dy: torch.Tensor) -> torch.Tensor:
    dy, ld = ops._rows(dy)
    out = _zeros((dy.shape[1],), dy)
    call("ccn_col_sum", dy, ld, dy.shape[0], dy.shape[1], out)
    return out


class _Linear(Function):
    """y = x W^T + b.  fp32 mode: FP32-pipe kernels.  Tensor-core modes: x is packed into split-bf16 planes once; the
    planes serve the forward GEMM and - kept for backward instead of x - the weight gradient; dY is packed once in
    backward and serves both dX = dY W and dW = dY^T X (all three GEMMs on tcgen05)."""

    @staticmethod
    def forward(ctx, x, w, b):
        ctx.has_bias = b is not None
        ctx.mode = mode = _precision()    # backward runs after ModelBase.forward has restored the global mode
        wd = w.detach().contiguous()
        bd = None if b is None else b.detach()
        ctx.xp = None
        ctx.had_xp = False
        if mode == "fp32" or x.shape[0] == 0:
            ctx.save_for_backward(x, w)
            return ops.linear(x, wd, bd, ops.ACT_NONE)
        xp = _planes_of(x, _npl(mode)) or ops.pack_planes([x], _npl(mode))
        if _tc_wgrad_ok(wd.shape[0], wd.shape[1]):
            ctx.xp = xp
            ctx.had_xp = True
            ctx.save_for_backward(None, w)
        else:
            ctx.save_for_backward(x, w)
        return ops.linear_planes(xp, ops.tc_weight_param(w), bd, ops.ACT_NONE, 0.0, "f32")

    @staticmethod
    def backward(ctx, dy):
        x, w = ctx.saved_tensors
        dx = dW = db = None
        need_w = ctx.needs_input_grad[1] or (ctx.has_bias and ctx.needs_input_grad[2])
        if ctx.mode == "fp32" or dy.shape[0] == 0:
            if ctx.needs_input_grad[0]:
                dx = ops.linear(dy, w.detach().t().contiguous(), None, ops.ACT_NONE)        # dX = dY W
            if need_w:
                dW, db = linear_wgrad(dy, x, ctx.has_bias)
            return dx, dW, db
        dyp = _planes_of(dy, _npl(ctx.mode)) or ops.pack_planes([dy], _npl(ctx.mode))
        if ctx.needs_input_grad[0]:
            # dX = dY W straight from the forward weight's planes (MN-major B tiles): no transpose, no second split
            dx = ops.linear_planes_bt(dyp, ops.tc_weight_param(w))
        if need_w:
            if ctx.had_xp and ctx.xp is None:
                raise RuntimeError("second backward through a tensor-core Linear: its input planes were released after the first "
                                   "(retain_graph=True is not supported on this path)")
            if ctx.xp is not None:
                dW = wgrad_planes(dyp, ctx.xp)
                db = _colsum_of(dy) if ctx.has_bias else None
                if ctx.has_bias and db is None:
                    db = col_sum(dy)
            else:
                dW, db = linear_wgrad(dy, x, ctx.has_bias)
        ctx.xp = None
        return dx, dW, db


def linear(x, w, b=None):
    return _Linear.apply(x, w, b)


# ----------------------------------------------------------------------------------------------
# BatchNorm (+ activation), dropout
# ----------------------------------------------------------------------------------------------
def _npl(mode: str) -> int:
    return {"fp32": 0, "bf16x3": 2, "bf16": 1}[mode]


def _planes_of(t: torch.Tensor, npl: int):
    """Activation planes a producer kernel attached to `t` (same values, same shape), else None."""
    pl = getattr(t, "_ccn_planes", None)
    if pl is not None and pl.n_planes == npl and pl.shape == tuple(t.shape) and _tag_ok(t):
        return pl
    return None


def _tag(t: torch.Tensor):
    """Stamp of the tensor contents the attached planes / column sums were computed from: an in-place accumulation into the
    same TensorImpl (a second gradient contribution) bumps the version and invalidates them."""
    t._ccn_tag = (t._version, t.data_ptr())


def _tag_ok(t: torch.Tensor) -> bool:
    return getattr(t, "_ccn_tag", None) == (t._version, t.data_ptr())


def _colsum_of(t: torch.Tensor):
    cs = getattr(t, "_ccn_colsum", None)
    return cs if cs is not None and _tag_ok(t) else None


class _BNAct(Function):
    """act(BatchNorm(z)).  On the tensor-core modes the forward kernel ALSO writes y as activation planes and the
    backward kernel writes dz as planes plus the column sums of dz; they ride on the returned tensors
    (`_ccn_planes`, `_ccn_colsum`) so that the Linear behind / in front needs no pack and no bias-gradient pass."""

    @staticmethod
    def forward(ctx, z, gamma, beta, bn, act, slope, batch_stats, dz_planes=True):
        z, ldz = ops._rows(z)
        M, C = z.shape
        y = _empty((M, C), z)
        npl = _npl(_precision()) if M > 0 else 0
        ctx.npl = npl if dz_planes else 0
        yp = ops.alloc_planes(M, C, npl, z) if npl else None
        pargs = (yp.buf, yp.kpad, yp.plane_stride, npl) if npl else (0, 0, 0, 0)
        if gamma is None:
            call("ccn_bn_apply", z, ldz, M, C, 0, 0, 0, 0, act, float(slope), y, C, *pargs)
            mean = invstd = None
        else:
            if batch_stats:
                mean, invstd = _empty((C,), z), _empty((C,), z)
                sums = _empty((2 * C,), z, torch.float64)
                track = bn.track_running_stats and bn.running_mean is not None
                if bn.momentum is None and track:
                    # torch's cumulative moving average (momentum=None -> 1 / num_batches_tracked) is not built
                    raise NotImplementedError("BatchNorm1d(momentum=None) (cumulative moving average) is not supported")
                mom = bn.momentum if bn.momentum is not None else 0.0
                call("ccn_bn_stats", z, ldz, M, C, float(bn.eps), float(mom), sums, mean, invstd,
                     bn.running_mean if track else 0, bn.running_var if track else 0)
                if track:
                    # the kernel updated the running statistics through raw pointers: bump their version counters so that
                    # the folded-weight caches (nn.MLP._key, modules._CurveConvBase._layers) see the change
                    torch.autograd.graph.increment_version(bn.running_mean)
                    torch.autograd.graph.increment_version(bn.running_var)
                    if bn.num_batches_tracked is not None:
                        bn.num_batches_tracked.add_(1)
            else:
                mean = bn.running_mean
                invstd = torch.rsqrt(bn.running_var + bn.eps)           # C values: weight preparation
            call("ccn_bn_apply", z, ldz, M, C, mean, invstd, gamma.detach(), beta.detach(), act, float(slope), y, C,
                 *pargs)
        ctx.save_for_backward(z, gamma, beta, mean, invstd)
        ctx.cfg = (act, float(slope), bool(batch_stats))
        if yp is not None:
            y._ccn_planes = yp
            _tag(y)
        return y

    @staticmethod
    def backward(ctx, dy):
        z, gamma, beta, mean, invstd = ctx.saved_tensors
        act, slope, batch_stats = ctx.cfg
        dy, lddy = ops._rows(dy)
        z, ldz = ops._rows(z)
        M, C = z.shape
        dz = _empty((M, C), z)
        npl = ctx.npl
        dzp = ops.alloc_planes(M, C, npl, z) if npl else None
        pargs = (dzp.buf, dzp.kpad, dzp.plane_stride, npl) if npl else (0, 0, 0, 0)
        if gamma is None:
            call("ccn_bn_backward", dy, lddy, z, ldz, M, C, 0, 0, 0, 0, act, slope, 0, 0, dz, C, 0, 0, 0, *pargs)
        else:
            dgamma, dbeta, dzsum = _empty((C,), z), _empty((C,), z), _empty((C,), z)
            sums = _empty((2 * C,), z, torch.float64)
            call("ccn_bn_backward", dy, lddy, z, ldz, M, C, mean, invstd, gamma.detach(), beta.detach(), act, slope,
                 int(batch_stats), sums, dz, C, dgamma, dbeta, dzsum, *pargs)
            dz._ccn_colsum = dzsum
        if dzp is not None:
            dz._ccn_planes = dzp
        _tag(dz)
        if gamma is None:
            return dz, None, None, None, None, None, None, None
        return dz, dgamma, dbeta, None, None, None, None, None


def bn_act(z, bn: Optional[torch.nn.BatchNorm1d], act: int, slope: float, dz_planes: bool = True):
    """act(BatchNorm1d(z)) (bn None: act(z)); batch statistics iff bn.training.  dz_planes=False: the producer of z is
    not a GEMM (per-edge gather layers), so backward need not emit dz as planes."""
    if bn is None:
        if act == ops.ACT_NONE:
            return z
        return _BNAct.apply(z, None, None, None, act, slope, False, dz_planes)
    return _BNAct.apply(z, bn.weight, bn.bias, bn, act, slope, bn.training, dz_planes)


class _Dropout(Function):
    @staticmethod
    def forward(ctx, x, u, p):
        x, ldx = ops._rows(x)
        y = _empty(x.shape, x)
        call("ccn_dropout", x, ldx, u, x.shape[0], x.shape[1], float(p), y, x.shape[1])
        ctx.save_for_backward(u)
        ctx.p = float(p)
        return y

    @staticmethod
    def backward(ctx, dy):
        (u,) = ctx.saved_tensors
        dy, ld = ops._rows(dy)
        dx = _empty(dy.shape, dy)
        call("ccn_dropout", dy, ld, u, dy.shape[0], dy.shape[1], ctx.p, dx, dy.shape[1])
        return dx, None, None


def dropout(x, p: float, training: bool):
    """F.dropout: the uniform draws come from torch's CUDA generator, the masking kernel is ours."""
    if not training or p <= 0.0:
        return x
    u = torch.rand(x.shape, dtype=torch.float32, device=x.device)
    return _Dropout.apply(x, u, p)


def mlp_train(mlp, x, first: int = 0):
    """torch_geometric.nn.MLP.forward in training form: Linear -> BatchNorm1d (batch statistics) -> act -> dropout
    per hidden layer, plain last Linear (+ its dropout) - layers `first`.. only."""
    act, slope = _ACTS[mlp.act_name]
    nl = len(mlp.lins)
    for i in range(first, nl):
        lin = mlp.lins[i]
        hidden = i < nl - 1 or not mlp.plain_last
        x = linear(x, lin.weight, lin.bias)
        if hidden:
            x = hidden_tail(mlp, i, x)
        else:
            x = dropout(x, mlp.dropout[i], mlp.training)
    return x


def hidden_tail(mlp, i: int, z, dz_planes: bool = True):
    """BatchNorm -> act -> dropout of hidden layer i applied to its pre-activation z."""
    act, slope = _ACTS[mlp.act_name]
    norm = mlp.norms[i] if i < len(mlp.norms) else None
    bn = norm.module if norm is not None and hasattr(norm, "module") else None
    z = bn_act(z, bn, act, slope, dz_planes)
    return dropout(z, mlp.dropout[i], mlp.training)


# ----------------------------------------------------------------------------------------------
# gathers / scatters / concatenation
# ----------------------------------------------------------------------------------------------
class _GatherRows(Function):
    """x[idx] for int64 idx (down-sampling subsets, per-sample broadcast)."""

    @staticmethod
    def forward(ctx, x, idx):
        ctx.save_for_backward(idx)
        ctx.n_src = x.shape[0]
        return ops.gather_rows(x, idx)

    @staticmethod
    def backward(ctx, dy):
        (idx,) = ctx.saved_tensors
        dy, ld = ops._rows(dy)
        dx = _zeros((ctx.n_src, dy.shape[1]), dy)
        call("ccn_scatter_add_rows_i64", dy, ld, idx, idx.shape[0], dy.shape[1], dx, dy.shape[1])
        return dx, None


def gather_rows(x, idx):
    if not x.requires_grad:
        return ops.gather_rows(x, idx)
    return _GatherRows.apply(x, ops._i64(idx))


def _gather_i32(src, idx):
    src, lds = ops._rows(src)
    out = _empty((idx.shape[0], src.shape[1]), src)
    call("ccn_gather_rows_i32", src, lds, idx, idx.shape[0], src.shape[1], out, src.shape[1])
    return out


def _scatter_add_i32(src, idx, n_dst):
    src, lds = ops._rows(src)
    out = _zeros((n_dst, src.shape[1]), src)
    call("ccn_scatter_add_rows_i32", src, lds, idx, idx.shape[0], src.shape[1], out, src.shape[1])
    return out


class _PadRows(Function):
    """sig[r] = row2pt[r] >= 0 ? x[row2pt[r]] : 0  (the zero rows of fast_conv1d.py:48-61)."""

    @staticmethod
    def forward(ctx, x, row2pt):
        ctx.save_for_backward(row2pt)
        ctx.n = x.shape[0]
        return _gather_i32(x, row2pt)

    @staticmethod
    def backward(ctx, dy):
        (row2pt,) = ctx.saved_tensors
        return _scatter_add_i32(dy, row2pt, ctx.n), None


class _UnpadRows(Function):
    """x[row2pt[r]] = sig[r] for the valid rows (sig[valid], fast_conv1d.py:75)."""

    @staticmethod
    def forward(ctx, sig, row2pt, n):
        ctx.save_for_backward(row2pt)
        return _scatter_add_i32(sig, row2pt, n)

    @staticmethod
    def backward(ctx, dy):
        (row2pt,) = ctx.saved_tensors
        return _gather_i32(dy, row2pt), None, None


def pad_rows(x, row2pt):
    return _PadRows.apply(x, row2pt)


def unpad_rows(sig, row2pt, n):
    return _UnpadRows.apply(sig, row2pt, n)


class _AppendZeroRows(Function):
    """[x ; 0] with n_pad zero rows (to_batch_padded's padding points, point_ops.py:263-284, as extra rows)."""

    @staticmethod
    def forward(ctx, x, n_pad):
        x, ld = ops._rows(x)
        n, c = x.shape
        ctx.n = n
        out = _zeros((n + n_pad, c), x)
        call("ccn_gather_rows_f32", x, ld, 0, n, c, out, c)
        return out

    @staticmethod
    def backward(ctx, dy):
        return dy[:ctx.n], None


def append_zero_rows(x, n_pad: int):
    return x if n_pad <= 0 else _AppendZeroRows.apply(x, n_pad)


class _Concat(Function):
    @staticmethod
    def forward(ctx, *parts):
        ctx.widths = [p.shape[1] for p in parts]
        return ops.concat_cols(list(parts))

    @staticmethod
    def backward(ctx, dy):
        outs, off = [], 0
        for i, w in enumerate(ctx.widths):
            outs.append(dy[:, off:off + w] if ctx.needs_input_grad[i] else None)   # views: our kernels take a stride
            off += w
        return tuple(outs)


def concat(parts: List[torch.Tensor]):
    parts = [p for p in parts if p is not None]
    return parts[0] if len(parts) == 1 else _Concat.apply(*parts)


# ----------------------------------------------------------------------------------------------
# per-edge first layers (pre-activation: BatchNorm needs the raw Linear output)
# ----------------------------------------------------------------------------------------------
class _EdgeLayer1(Function):
    """z_e = H[col_e] + Wp (pos_src[col_e] - pos_dst[row_e]) / r + b     (point_conv.py:60-69, first Linear)."""

    @staticmethod
    def forward(ctx, H, wp_t, b, pos_src, pos_dst, edges, radius):
        ctx.edges, ctx.radius = edges, radius
        ctx.n_src = pos_src.shape[0]
        ctx.save_for_backward(pos_src, pos_dst)
        ctx.has = (H is not None, b is not None)
        return ops.edge_layer1(None if H is None else H.detach(), wp_t.detach().contiguous(),
                               None if b is None else b.detach(), pos_src, pos_dst, edges, radius, ops.ACT_NONE)

    @staticmethod
    def backward(ctx, dz):
        pos_src, pos_dst = ctx.saved_tensors
        edges = ctx.edges
        dz, ld = ops._rows(dz)
        E, c = dz.shape
        dH = dwp = db = None
        if ctx.has[0] and ctx.needs_input_grad[0]:
            dH = _scatter_add_i32(dz, edges.ecol, ctx.n_src)
        if ctx.needs_input_grad[1] or (ctx.has[1] and ctx.needs_input_grad[2]):
            rel = _empty((max(E, 1), 3), dz)
            call("ccn_edge_rel_pos", ops._f32(pos_src), ops._f32(pos_dst), edges.erow, edges.ecol, E,
                 float(ctx.radius or 0.0), rel)
            dW, db = linear_wgrad(dz, rel[:E], ctx.has[1])          # (c, 3)
            dwp = dW.t()
        return dH, dwp, db, None, None, None, None


def edge_layer1(H, wp_t, b, pos_src, pos_dst, edges, radius):
    """wp_t: (3, C) = W[:, cx:].T"""
    return _EdgeLayer1.apply(H, wp_t, b, pos_src, pos_dst, edges, radius)


class _EdgePair(Function):
    """z_e = P[col_e] + Q[row_e], PQ = [P | Q]            (dgcnn.py:227-228, first Linear, sparse path)."""

    @staticmethod
    def forward(ctx, PQ, edges):
        ctx.edges, ctx.n = edges, PQ.shape[0]
        return ops.edge_pair_layer1(PQ, edges, ops.ACT_NONE)

    @staticmethod
    def backward(ctx, dz):
        edges = ctx.edges
        dz, ld = ops._rows(dz)
        E, c = dz.shape
        dPQ = _zeros((ctx.n, 2 * c), dz)
        call("ccn_scatter_add_rows_i32", dz, ld, edges.ecol, E, c, dPQ, 2 * c)
        call("ccn_scatter_add_rows_i32", dz, ld, edges.erow, E, c, dPQ[:, c:], 2 * c)
        return dPQ, None


def edge_pair(PQ, edges):
    return _EdgePair.apply(PQ, edges)


class _SgcnnLayer1(Function):
    """rows (i, 0) = P[i] + Q[i], (i, 1+k) = P[nbr[i][k]] + Q[i] (P term dropped for missing neighbours)
    (dgcnn.py:166-173, first Linear, dense path)."""

    @staticmethod
    def forward(ctx, PQ, nbr):
        ctx.save_for_backward(nbr)
        ctx.n = PQ.shape[0]
        return ops.sgcnn_layer1(PQ, nbr, ops.ACT_NONE)

    @staticmethod
    def backward(ctx, dz):
        (nbr,) = ctx.saved_tensors
        dz = dz.contiguous()
        n, K = nbr.shape
        G, c = K + 1, dz.shape[1]
        dPQ = _zeros((n, 2 * c), dz)
        call("ccn_group_sum_rows", dz, c, n, G, c, dPQ[:, c:], 2 * c, 0)               # dQ[i] = sum over the slots
        slots = torch.cat([_iota32(n, dz).view(n, 1), nbr], dim=1).contiguous()        # source row of every slot
        call("ccn_scatter_add_rows_i32", dz, c, slots, n * G, c, dPQ, 2 * c)            # dP[j] += dz[(i, slot)]
        return dPQ, None


_IOTA = {}


def _iota32(n, like):
    key = (n, like.device)
    t = _IOTA.get(key)
    if t is None:
        if len(_IOTA) > 64:
            _IOTA.clear()
        t = _IOTA[key] = torch.arange(n, dtype=torch.int32, device=like.device)
    return t


def sgcnn_layer1(PQ, nbr):
    return _SgcnnLayer1.apply(PQ, nbr)


# ----------------------------------------------------------------------------------------------
# aggregation, interpolation
# ----------------------------------------------------------------------------------------------
class _SegAggregate(Function):
    @staticmethod
    def forward(ctx, msg, att, row_ptr, nq, mode):
        ctx.save_for_backward(msg, att, row_ptr)
        ctx.cfg = (nq, mode)
        return ops.seg_aggregate(msg, att, row_ptr, nq, mode)

    @staticmethod
    def backward(ctx, dout):
        msg, att, row_ptr = ctx.saved_tensors
        nq, mode = ctx.cfg
        dout, ldo = ops._rows(dout)
        msg, ldm = ops._rows(msg)
        E, c = msg.shape
        dmsg = _zeros((E, c), msg) if E else _empty((0, c), msg)     # rows outside every segment stay zero
        datt = None
        lda = 0
        if att is not None:
            att, lda = ops._rows(att)
            datt = _zeros((E, c), msg)
        call("ccn_seg_aggregate_bwd", msg, ldm, att if att is not None else 0, lda, row_ptr, nq, c, ops.AGG[mode],
             dout, ldo, dmsg, c, datt if datt is not None else 0, c)
        return dmsg, datt, None, None, None


def seg_aggregate(msg, att, row_ptr, nq, mode):
    return _SegAggregate.apply(msg, att, row_ptr, nq, mode)


class _GroupMaxDense(Function):
    @staticmethod
    def forward(ctx, h, nbr):
        ctx.save_for_backward(h, nbr)
        return ops.group_max_dense(h, nbr)

    @staticmethod
    def backward(ctx, dout):
        h, nbr = ctx.saved_tensors
        h, ldf = ops._rows(h)
        dout, ldo = ops._rows(dout)
        n, K = nbr.shape
        c = h.shape[1]
        dh = _empty(h.shape, h) if h.shape[0] == n * (K + 1) else _zeros(h.shape, h)   # rows of padded points: no gradient
        call("ccn_group_max_dense_bwd", h, ldf, nbr, n, K, c, dout, ldo, dh, c)
        return dh, None


def group_max_dense(h, nbr):
    return _GroupMaxDense.apply(h, nbr)


class _InterpBlend(Function):
    @staticmethod
    def forward(ctx, x, nbr, wgt, from_d2):
        ctx.save_for_backward(nbr, wgt)
        ctx.cfg = (x.shape[0], bool(from_d2))
        return ops.interp_blend(x, nbr, wgt, from_d2)

    @staticmethod
    def backward(ctx, dout):
        nbr, wgt = ctx.saved_tensors
        n_src, from_d2 = ctx.cfg
        dout, ldo = ops._rows(dout)
        n, knn = nbr.shape
        c = dout.shape[1]
        dx = _zeros((n_src, c), dout)
        call("ccn_interp_blend_bwd", nbr, wgt, int(from_d2), n, knn, c, dout, ldo, dx, c)
        return dx, None, None, None


def interp_blend(x, nbr, wgt, from_d2):
    return _InterpBlend.apply(x, nbr, wgt, from_d2)


# ----------------------------------------------------------------------------------------------
# curve convolution
# ----------------------------------------------------------------------------------------------
class _CurveFeatPrep(Function):
    @staticmethod
    def forward(ctx, x, pos, with_xyz, with_diff, curve_ord):
        ctx.save_for_backward(x, curve_ord)
        ctx.cfg = (bool(with_xyz), bool(with_diff))
        return ops.curve_feat_prep(x, pos, with_xyz, with_diff, curve_ord)

    @staticmethod
    def backward(ctx, dout):
        x, curve_ord = ctx.saved_tensors
        with_xyz, with_diff = ctx.cfg
        x, ldx = ops._rows(x)
        dout, ldo = ops._rows(dout)
        n, cx = x.shape
        dx = _empty((n, cx), x)
        call("ccn_curve_feat_prep_bwd", x, ldx, cx, int(with_xyz), int(with_diff), curve_ord, n, dout, ldo, dx, cx)
        return dx, None, None, None, None


def curve_feat_prep(x, pos, with_xyz, with_diff, curve_ord):
    if x is None or not x.requires_grad:
        return ops.curve_feat_prep(x, pos, with_xyz, with_diff, curve_ord)
    return _CurveFeatPrep.apply(x, pos, with_xyz, with_diff, curve_ord)


class _CurveConv(Function):
    """z = conv1d('same', zero padded) of the padded signal (rows, cin) with w (k, cin, cout) + b; no norm, no act.
    fp32 mode: the FP32-pipe tap kernel (ccn_curve_conv_layer).  Tensor-core modes: forward, data gradient and weight
    gradient are tcgen05 GEMMs over the signal's activation planes - a tap's row shift is a TMA coordinate."""

    @staticmethod
    def forward(ctx, sig, w, b, ksize):
        sig, _ = ops._rows(sig)
        ctx.ksize = ksize
        ctx.mode = mode = _precision()
        rows = sig.shape[0]
        wd = w.detach().contiguous()
        ctx.tc = mode != "fp32" and rows > 0 and wd.shape[1] >= 16 and wd.shape[2] >= 16
        if ctx.tc:
            sigp = _planes_of(sig, _npl(mode)) or ops.pack_planes([sig], _npl(mode))
            ctx.sigp = sigp
            ctx.save_for_backward(None, w)
            return ops.conv_planes(sigp, wd, b.detach())
        ctx.save_for_backward(sig, w)
        return ops.curve_conv_layer(sig, wd, b.detach(), _iota32(1, sig), rows, rows, ksize, False, False, slope=1.0)

    @staticmethod
    def backward(ctx, dz):
        sig, w = ctx.saved_tensors
        k = ctx.ksize
        dz_in, dz = dz, dz.contiguous()
        rows, cout = dz.shape
        cin = w.shape[1]
        dsig = dw = db = None
        half = k // 2
        need_w = ctx.needs_input_grad[1] or ctx.needs_input_grad[2]
        wt = torch.flip(w.detach(), dims=[0]).transpose(1, 2).contiguous()           # (k, cout, cin): weight prep
        if ctx.tc:
            dzp = _planes_of(dz_in, _npl(ctx.mode)) or ops.pack_planes([dz], _npl(ctx.mode))
            if ctx.needs_input_grad[0]:
                dsig = ops.conv_planes(dzp, wt, None)
            if need_w:
                # one tcgen05 weight-gradient GEMM per tap: the tap's row shift is a TMA coordinate offset
                dw = torch.stack([wgrad_planes(dzp, ctx.sigp, tap - half).t() for tap in range(k)], dim=0)   # (k, cin, cout)
                db = col_sum(dz)
            ctx.sigp = None
            return dsig, dw, db, None
        if ctx.needs_input_grad[0]:
            dsig = ops.curve_conv_layer(dz, wt, _zeros((cin,), dz), _iota32(1, dz), rows, rows, k, False, False, slope=1.0)
        if need_w:
            dw = _zeros((k, cin, cout), dz)
            db = _zeros((cout,), dz)
            for tap in range(k):
                off = tap - half                  # z[r] reads sig[r + off]
                lo, hi = max(0, -off), min(rows, rows - off)
                if hi <= lo:
                    continue
                dwt = _zeros((cout, cin), dz)
                call("ccn_linear_wgrad", dz[lo:hi], cout, sig[lo + off:hi + off], cin, hi - lo, cout, cin, dwt, cin,
                     db if tap == half else 0)
                dw[tap] = dwt.t()
        return dsig, dw, db, None


def curve_conv(sig, w, b, ksize):
    return _CurveConv.apply(sig, w, b, ksize)


# ----------------------------------------------------------------------------------------------
# loss
# ----------------------------------------------------------------------------------------------
class _NllLoss(Function):
    @staticmethod
    def forward(ctx, logits, target, weight, ignore_index, perpoint=None):
        logits, ld = ops._rows(logits)
        n, C = logits.shape
        target = ops._i64(target)
        lse = _empty((n,), logits)
        acc = _empty((2,), logits, torch.float64)
        loss = _empty((1,), logits)
        call("ccn_nll_loss_fwd", logits, ld, target, weight if weight is not None else 0, n, C, int(ignore_index), lse,
             acc, loss, perpoint if perpoint is not None else 0)
        ctx.save_for_backward(logits, target, weight, lse, acc)
        ctx.ignore = int(ignore_index)
        return loss.view(())

    @staticmethod
    def backward(ctx, g):
        logits, target, weight, lse, acc = ctx.saved_tensors
        logits, ld = ops._rows(logits)
        n, C = logits.shape
        d = _empty((n, C), logits)
        call("ccn_nll_loss_bwd", logits, ld, target, weight if weight is not None else 0, n, C, ctx.ignore, lse, acc,
             g.reshape(1).contiguous().float(), d, C)
        return d, None, None, None, None


def nll_loss(logits, target, weight=None, ignore_index=-100, perpoint_out=None):
    """F.nll_loss(F.log_softmax(logits, dim=1), target, weight=weight, ignore_index=ignore_index) (mean).
    perpoint_out: optional f32 (n) buffer that receives the unweighted per-point NLL (reduction="none")."""
    return _NllLoss.apply(logits, target, weight, ignore_index, perpoint_out)
