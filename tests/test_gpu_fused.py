"""GPU parity of the fused edge-MLP kernels (csrc/edge_fused.cu: ccn_sgcnn_fused, ccn_pointconv_fused): against a plain
torch fp64 statement of the reference arithmetic (dgcnn.py:166-189, point_conv.py:60-93) within the precision budget, and
BIT FOR BIT against the unfused chain of kernels they replace (same MMA order, same bias / activation / reduction order)."""
import pytest
import torch

from golden_utils import rel_err

pytestmark = pytest.mark.gpu
DEV = "cuda"
TOL = {"bf16x3": 2e-5, "bf16": 1.5e-2}


def _act(x, act, slope=0.01):
    if act == 1:
        return x.clamp(min=0)
    if act == 2:
        return torch.where(x > 0, x, x * slope)
    return x


def _mlp_layers(dims, acts, seed, bias=True):
    gen = torch.Generator().manual_seed(seed)
    out = []
    for (k, n), a in zip(zip(dims[:-1], dims[1:]), acts):
        w = (torch.randn(n, k, generator=gen) / k ** 0.5).to(DEV)
        b = (torch.randn(n, generator=gen) * 0.1).to(DEV) if bias else None
        out.append((w, b, a, 0.01))
    return out


def _set(precision):
    from curvecloudnet_b200 import nn
    nn.set_precision(precision)
    return nn.n_planes()


@pytest.mark.parametrize("precision", ["bf16x3", "bf16"])
@pytest.mark.parametrize("n,K,dims", [(1000, 20, [64, 64, 64]), (37, 20, [128, 128]), (5000, 20, [256, 256]),
                                       (777, 9, [64, 128, 192, 256]), (4, 31, [128, 64]), (2500, 20, [64, 64]),
                                       (12180 // 21, 20, [512, 512]), (174, 20, [1024, 1024]), (3000, 20, [128, 192]),
                                       (900, 5, [256, 64])])
def test_sgcnn_fused(precision, n, K, dims):
    from curvecloudnet_b200 import nn, ops
    from curvecloudnet_b200.nn import ForwardCache, PreparedLinear
    gen = torch.Generator().manual_seed(n + K)
    c1 = dims[0]
    PQ = torch.randn(n, 2 * c1, generator=gen).to(DEV)
    idx = torch.randint(0, n, (n, K), generator=gen, dtype=torch.int32)
    # -1 padded tails (missing neighbours), some rows with no neighbour at all
    cnt = torch.randint(0, K + 1, (n,), generator=gen)
    idx[torch.arange(K).view(1, K) >= cnt.view(n, 1)] = -1
    idx = idx.to(DEV)
    acts = [2] * (len(dims) - 2) + [0]                       # hidden LeakyReLU, plain last (PyG MLP plain_last)
    layers = _mlp_layers(dims, acts, seed=5)
    try:
        npl = _set(precision)
        chain = ops.ChainWeights(layers, npl)
        cache = ForwardCache()
        got = ops.sgcnn_fused(PQ, idx, 1, 0.01, chain, npl, cache)
        planes = cache.planes_for(got, npl)
        # (1) the unfused chain: layer-1 rows -> Linear launches -> masked group max
        h = ops.sgcnn_layer1(PQ, idx, 1, 0.01, npl)
        for i, (w, b, a, s) in enumerate(layers):
            h = PreparedLinear(w, b, a, s)(h, "f32" if i == len(layers) - 1 else "planes")
        want_unfused = ops.group_max_dense(h, idx)
        assert torch.equal(got, want_unfused), "fused and unfused chains differ by %.3e" % rel_err(got.cpu(), want_unfused.cpu())
        assert torch.equal(planes.to_f32(), ops.pack_planes([got], npl).to_f32())
    finally:
        nn.set_precision("fp32")
    # (2) the reference arithmetic in fp64: rows [x_j ; x_i - x_j] folded into P[j] + Q[i] (dgcnn.py:166-173), MLP, -100 mask, max
    P, Q = PQ[:, :c1].double(), PQ[:, c1:].double()
    full = torch.cat([torch.arange(n, device=DEV, dtype=torch.int64).view(n, 1), idx.long()], dim=1)       # self loop first
    pj = torch.where((full >= 0).unsqueeze(-1), P[full.clamp(min=0)], torch.zeros((), dtype=torch.float64, device=DEV))
    x = _act(pj + Q.unsqueeze(1), 1)
    for w, b, a, s in layers:
        x = _act(x @ w.double().t() + b.double(), a, s)
    x = torch.where((full >= 0).unsqueeze(-1), x, torch.full_like(x, -100.0))
    want = x.max(dim=1).values
    err = rel_err(got.cpu(), want.cpu())
    assert err <= TOL[precision], "fused SGCNN step off by %.3e (%s)" % (err, precision)


def _random_csr(nq, n_src, max_deg, seed, empties=True):
    from curvecloudnet_b200 import ops
    gen = torch.Generator().manual_seed(seed)
    deg = torch.randint(0 if empties else 1, max_deg + 1, (nq,), generator=gen)
    if empties and nq > 10:
        deg[3:9] = 0                                            # a run of targets without edges
        deg[-2:] = 0                                            # ... and trailing ones
    row_ptr = torch.zeros(nq + 1, dtype=torch.int64)
    row_ptr[1:] = torch.cumsum(deg, 0)
    E = int(row_ptr[-1])
    erow = torch.repeat_interleave(torch.arange(nq), deg)
    ecol = torch.randint(0, n_src, (E,), generator=gen)
    return ops.EdgeCSR(row_ptr.to(torch.int32).to(DEV), erow.to(torch.int32).to(DEV), ecol.to(torch.int32).to(DEV), E, nq, max_deg)


def _segment_ref(msg, att, edges, mode):
    """fp64 statement of point_conv.py:71-93 over a CSR edge list."""
    nq, C = edges.nq, msg.shape[1]
    out = torch.zeros((nq, C), dtype=torch.float64, device=msg.device)
    rp = edges.row_ptr.tolist()
    for qi in range(nq):
        a, b = rp[qi], rp[qi + 1]
        if b == a:
            continue
        m = msg[a:b]
        if mode == "max":
            out[qi] = m.max(dim=0).values
        elif mode == "mean":
            out[qi] = m.mean(dim=0)
        elif mode == "weighted-sum":
            out[qi] = (m * torch.sigmoid(att[a:b])).sum(dim=0)
        else:
            w = torch.exp(att[a:b] - att[a:b].max(dim=0).values)
            out[qi] = (m * (w / (w.sum(dim=0) + 1e-16))).sum(dim=0)
    return out


@pytest.mark.parametrize("precision", ["bf16x3", "bf16"])
@pytest.mark.parametrize("mode", ["max", "mean", "attend", "weighted-sum"])
@pytest.mark.parametrize("nq,n_src,max_deg,dims,att_hidden,radius", [
    (700, 3000, 9, [64, 128, 192, 256], 256, 0.02),      # sa-geo shape: curve groups, normalised radius, attend [C, C, C]
    (900, 2500, 32, [128, 128, 128], 64, None),          # sa: FRNN K = 32, attend [C, C/2, C]
    (50, 400, 32, [256, 256], 128, None),
    (13, 40, 64, [64, 64], 64, 0.5)])
def test_pointconv_fused(precision, mode, nq, n_src, max_deg, dims, att_hidden, radius):
    from curvecloudnet_b200 import nn, ops
    from curvecloudnet_b200.nn import ForwardCache, PreparedLinear
    edges = _random_csr(nq, n_src, max_deg, seed=nq)
    gen = torch.Generator().manual_seed(7)
    c1, C = dims[0], dims[-1]
    H = torch.randn(n_src, c1, generator=gen).to(DEV)
    wp = torch.randn(3, c1, generator=gen).to(DEV)
    b0 = (torch.randn(c1, generator=gen) * 0.1).to(DEV)
    pos_src = torch.rand(n_src, 3, generator=gen).to(DEV)
    pos_dst = torch.rand(nq, 3, generator=gen).to(DEV)
    local = _mlp_layers(dims, [2] * (len(dims) - 2) + [0], seed=11)
    att = mode in ("attend", "weighted-sum")
    attend = _mlp_layers([C, att_hidden, C], [2, 0], seed=12) if att else []
    msg_layer = len(local) - 1 if att else -1
    try:
        npl = _set(precision)
        chain = ops.ChainWeights(local + attend, npl)
        cache = ForwardCache()
        got = ops.pointconv_fused(edges, chain, mode, msg_layer, 2, 0.01, c1, H=H, wp=wp, bias0=b0, pos_src=pos_src,
                                  pos_dst=pos_dst, radius=radius, npl=npl, cache=cache)
        # (1) the unfused chain of kernels
        h = ops.edge_layer1(H, wp, b0, pos_src, pos_dst, edges, radius, 2, 0.01, npl)
        for i, (w, b, a, s) in enumerate(local):
            last = i == len(local) - 1
            h = PreparedLinear(w, b, a, s)(h, ("both" if att else "f32") if last else "planes")
        if att:
            msg, a_in = h
            for i, (w, b, a, s) in enumerate(attend):
                a_in = PreparedLinear(w, b, a, s)(a_in, "f32" if i == len(attend) - 1 else "planes")
            want_unfused = ops.seg_aggregate(msg, a_in, edges.row_ptr, edges.nq, mode)
        else:
            want_unfused = ops.seg_aggregate(h, None, edges.row_ptr, edges.nq, mode)
        assert torch.equal(got, want_unfused), "fused and unfused chains differ by %.3e" % rel_err(got.cpu(), want_unfused.cpu())
        assert torch.equal(cache.planes_for(got, npl).to_f32(), ops.pack_planes([got], npl).to_f32())
    finally:
        nn.set_precision("fp32")
    # (2) the reference arithmetic in fp64 (point_conv.py:60-93)
    j, i = edges.ecol.long(), edges.erow.long()
    rel = (pos_src[j] - pos_dst[i]).double()
    if radius:
        rel = rel / radius
    x = _act(H[j].double() + rel @ wp.double() + b0.double(), 2)
    for w, b, a, s in local:
        x = _act(x @ w.double().t() + b.double(), a, s)
    msg64, att64 = x, x
    for w, b, a, s in attend:
        att64 = _act(att64 @ w.double().t() + b.double(), a, s)
    want = _segment_ref(msg64, att64, edges, mode)
    err = rel_err(got.cpu(), want.cpu())
    # the segment softmax amplifies the operand rounding (see tests/test_gpu_fullsize_oracle.py STEP_TOL)
    tol = TOL[precision] * (10 if mode == "attend" else 3 if att else 1)
    assert err <= tol, "fused PointNetConv step (%s) off by %.3e (%s)" % (mode, err, precision)


@pytest.mark.parametrize("precision", ["bf16x3", "bf16"])
@pytest.mark.parametrize("mode", ["max", "attend"])
def test_sparse_sgcnn_fused(precision, mode):
    """Sparse StaticEdgeConv (dgcnn.py:209-246): first layer act(P[j] + Q[i]) over an edge list, MLP, scatter max / attention."""
    from curvecloudnet_b200 import nn, ops
    from curvecloudnet_b200.nn import ForwardCache, PreparedLinear
    n, K, dims = 1500, 30, [128, 128, 128]
    edges = _random_csr(n, n, K, seed=3, empties=False)
    gen = torch.Generator().manual_seed(9)
    c1, C = dims[0], dims[-1]
    PQ = torch.randn(n, 2 * c1, generator=gen).to(DEV)
    local = _mlp_layers(dims, [1] * (len(dims) - 2) + [0], seed=21)
    att = mode == "attend"
    attend = _mlp_layers([C, C, C], [2, 0], seed=22) if att else []
    try:
        npl = _set(precision)
        chain = ops.ChainWeights(local + attend, npl)
        got = ops.pointconv_fused(edges, chain, mode, len(local) - 1 if att else -1, 1, 0.01, c1, PQ=PQ, npl=npl, cache=ForwardCache())
        h = ops.edge_pair_layer1(PQ, edges, 1, 0.01, npl)
        for i, (w, b, a, s) in enumerate(local):
            last = i == len(local) - 1
            h = PreparedLinear(w, b, a, s)(h, ("both" if att else "f32") if last else "planes")
        if att:
            msg, a_in = h
            for i, (w, b, a, s) in enumerate(attend):
                a_in = PreparedLinear(w, b, a, s)(a_in, "f32" if i == len(attend) - 1 else "planes")
            want = ops.seg_aggregate(msg, a_in, edges.row_ptr, edges.nq, mode)
        else:
            want = ops.seg_aggregate(h, None, edges.row_ptr, edges.nq, mode)
        assert torch.equal(got, want), "fused and unfused chains differ by %.3e" % rel_err(got.cpu(), want.cpu())
    finally:
        nn.set_precision("fp32")


@pytest.mark.parametrize("precision", ["bf16x3", "bf16"])
@pytest.mark.parametrize("M,segs,dims", [(5000, [64], [64, 64, 17]), (700, [256], [128, 128, 64]), (3001, [64, 64], [128, 128]),
                                          (129, [32, 128], [128, 64]), (40000, [64, 32, 3], [128, 128]), (1, [128, 128], [256, 128])])
def test_mlp_chain_equals_consecutive_linears(precision, M, segs, dims):
    """ccn_mlp_chain (2+ layers of a shared per-point MLP in one launch, activation tile resident between the layers) against
    consecutive ccn_linear_planes launches: bit-identical fp32 rows and planes; segmented inputs, last width 17, M = 1."""
    from curvecloudnet_b200 import nn, ops
    from curvecloudnet_b200.nn import PreparedLinear
    gen = torch.Generator().manual_seed(M)
    parts = [torch.randn(M, k, generator=gen).to(DEV) for k in segs]
    layers = _mlp_layers([sum(segs)] + dims, [2] * (len(dims) - 1) + [0], seed=31)
    try:
        npl = _set(precision)
        A = ops.PlanesCat([ops.pack_planes([t], npl) for t in parts]) if len(parts) > 1 else ops.pack_planes(parts, npl)
        chain = ops.ChainWeights(layers, npl, first_seg_widths=segs)
        y, yp = ops.mlp_chain(A, chain, "both")
        lins = [PreparedLinear(w, b, a, s) for w, b, a, s in layers]
        h = A
        for i, lin in enumerate(lins):
            if i == 0 and len(parts) > 1:
                h = ops.linear_planes(h, lin.tc(tuple(segs)), lin.b, lin.act, lin.slope, "planes")
            else:
                h = lin(h, "both" if i == len(lins) - 1 else "planes")
        want, wantp = h
        assert torch.equal(y, want), "chain differs from the per-layer launches by %.3e" % rel_err(y.cpu(), want.cpu())
        assert torch.equal(yp.buf, wantp.buf)
        only = ops.mlp_chain(A, chain, "f32")
        assert torch.equal(only, want)
    finally:
        nn.set_precision("fp32")
