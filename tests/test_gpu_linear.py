"""GPU parity of the TMA/tcgen05 Linear kernel (ccn_linear_planes) and the planes pack kernel against float64
torch arithmetic.  Tolerances: split-bf16 x3 -> 1e-4 budget (we require 3e-5), bf16 -> 2e-2 (BASELINE.json)."""
import pytest
import torch

pytestmark = pytest.mark.gpu

SHAPES = [  # (M, N, K): resident big-M, streaming, ragged N/K, tiny M, one tile per CTA, wide K
    (40000, 64, 64), (30000, 256, 256), (20011, 128, 67), (20000, 192, 128), (5695, 256, 131), (1933, 512, 259),
    (174, 2048, 1027), (580, 1024, 2051), (33, 17, 64), (1, 64, 35), (129, 40, 8), (3654, 1024, 1024),
]


def _ref(x, w, b, act):
    y = x.double() @ w.double().t() + (b.double() if b is not None else 0)
    if act == 1:
        y = y.clamp(min=0)
    elif act == 2:
        y = torch.where(y > 0, y, y * 0.01)
    return y


@pytest.mark.parametrize("mode", ["bf16x3", "bf16"])
@pytest.mark.parametrize("shape", SHAPES)
def test_linear_planes_matches_float64(shape, mode):
    from curvecloudnet_b200 import nn as cnn, ops
    M, N, K = shape
    g = torch.Generator().manual_seed(M + N + K)
    x = torch.randn(M, K, generator=g).cuda()
    w = (torch.randn(N, K, generator=g) / K ** 0.5).cuda()
    b = torch.randn(N, generator=g).cuda()
    prev = cnn.get_precision()
    cnn.set_precision(mode)
    try:
        npl = cnn.n_planes()
        tol = 3e-5 if mode == "bf16x3" else 2e-2
        lin = cnn.PreparedLinear(w, b, act=2, slope=0.01)
        xp = ops.pack_planes([x], npl)
        assert xp.buf.shape == (npl, M, (K + 15) // 16 * 16)
        assert float(xp.buf[:, :, K:].float().abs().max() if xp.kpad > K else 0.0) == 0.0
        back = xp.to_f32()
        assert float((back - x).abs().max() / x.abs().max()) <= (2e-5 if npl == 2 else 8e-3)
        want = _ref(x, w, b, 2)
        scale = float(want.abs().max())
        y, yp = lin(xp, "both")
        torch.cuda.synchronize()
        assert float((y.double() - want).abs().max()) / scale <= tol
        assert float((yp.to_f32().double() - want).abs().max()) / scale <= tol * 1.5 + (8e-3 if npl == 1 else 0)
        if yp.kpad > N:
            assert float(yp.buf[:, :, N:].float().abs().max()) == 0.0
        y2 = lin(x)                                   # fp32 in -> pack -> GEMM -> fp32 out
        assert torch.equal(y2, y)
        lin0 = cnn.PreparedLinear(w, None, act=0)
        y3 = lin0(xp)
        assert float((y3.double() - _ref(x, w, None, 0)).abs().max()) / scale <= tol
    finally:
        cnn.set_precision(prev)


def test_pack_concat_and_gather():
    from curvecloudnet_b200 import ops
    g = torch.Generator().manual_seed(3)
    a = torch.randn(1000, 37, generator=g).cuda()
    b = torch.randn(1000, 3, generator=g).cuda()
    c = torch.randn(50, 64, generator=g).cuda()
    ic = torch.randint(0, 50, (1000,), generator=g).cuda()
    wide = torch.randn(1000, 80, generator=g).cuda()
    pl = ops.pack_planes([a, wide[:, 5:17], b, c], 2, [None, None, None, ic])
    want = torch.cat([a, wide[:, 5:17], b, c[ic]], dim=1)
    assert pl.shape == (1000, 37 + 12 + 3 + 64)
    assert float((pl.to_f32() - want).abs().max()) <= 2e-5 * float(want.abs().max())
    sel = torch.randint(0, 1000, (321,), generator=g).cuda()
    pl = ops.pack_planes([a, b], 1, [sel, sel])
    assert float((pl.to_f32() - torch.cat([a, b], 1)[sel]).abs().max()) <= 8e-3 * float(a.abs().max())


@pytest.mark.parametrize("mode", ["bf16x3", "bf16"])
@pytest.mark.parametrize("widths", [(64, 3), (256, 256, 3), (1024, 1024, 1024), (32, 3), (64, 32, 3), (100, 7, 129, 3)])
def test_linear_over_concatenated_segments(widths, mode):
    """[A_0 | A_1 | ...] W^T with every block its own TMA source == Linear over the materialised concatenation."""
    from curvecloudnet_b200 import nn as cnn, ops
    M, N = 3000, 192
    g = torch.Generator().manual_seed(sum(widths))
    parts = [torch.randn(M, k, generator=g).cuda() for k in widths]
    K = sum(widths)
    w = (torch.randn(N, K, generator=g) / K ** 0.5).cuda()
    b = torch.randn(N, generator=g).cuda()
    prev = cnn.get_precision()
    cnn.set_precision(mode)
    try:
        cache = cnn.ForwardCache()
        lin = cnn.PreparedLinear(w, b, act=1)
        a = cnn.cat_features(parts, cache=cache)
        assert isinstance(a, ops.PlanesCat) and a.shape == (M, K)
        y = lin(a)
        want = _ref(torch.cat(parts, 1), w, b, 1)
        tol = 3e-5 if mode == "bf16x3" else 2e-2
        assert float((y.double() - want).abs().max() / want.abs().max()) <= tol
        y2 = lin(torch.cat(parts, 1))                 # packed single-block path
        assert float((y2 - y).abs().max() / want.abs().max()) <= (1e-6 if mode == "bf16x3" else 1e-6)
        # planes are made once per tensor and forward
        assert cnn.cat_features(parts, cache=cache).segs[0] is a.segs[0]
    finally:
        cnn.set_precision(prev)


def test_producers_write_planes_identical_to_pack():
    """ccn_*_both / ccn_subset_gather(pos_planes): the planes a producer writes beside its fp32 result are bit-identical to
    ccn_pack_planes of that result, and the fp32 result equals the plain entry point's."""
    from curvecloudnet_b200 import ops
    from curvecloudnet_b200.nn import ForwardCache
    g = torch.Generator().manual_seed(21)
    n, K, c = 3001, 9, 64
    feat = torch.randn(n * (K + 1), c, generator=g).cuda()
    idx = torch.randint(-1, n, (n, K), generator=g, dtype=torch.int32).cuda()
    for npl in (1, 2):
        cache = ForwardCache()
        a, b = ops.group_max_dense(feat, idx), ops.group_max_dense(feat, idx, npl, cache)
        assert torch.equal(a, b) and len(cache.planes) == 1
        assert torch.equal(cache.planes_for(b, npl).buf, ops.pack_planes([a], npl).buf)
        # segmented aggregation (max and attention) over a ragged CSR
        cnt = torch.randint(0, 12, (n,), generator=g)
        row_ptr = torch.cat([torch.zeros(1, dtype=torch.int64), torch.cumsum(cnt, 0)]).to(torch.int32).cuda()
        E = int(cnt.sum())
        msg, att = torch.randn(E, c, generator=g).cuda(), torch.randn(E, c, generator=g).cuda()
        for mode, at in (("max", None), ("attend", att), ("mean", None)):
            cache = ForwardCache()
            a, b = ops.seg_aggregate(msg, at, row_ptr, n, mode), ops.seg_aggregate(msg, at, row_ptr, n, mode, npl, cache)
            assert torch.equal(a, b) and len(cache.planes) == 1
            assert torch.equal(cache.planes_for(b, npl).buf, ops.pack_planes([a], npl).buf)
        # interpolation
        x = torch.randn(700, c, generator=g).cuda()
        nbr = torch.randint(0, 700, (n, 3), generator=g, dtype=torch.int32).cuda()
        d2 = torch.rand(n, 3, generator=g).cuda()
        cache = ForwardCache()
        a, b = ops.interp_blend(x, nbr, d2, True), ops.interp_blend(x, nbr, d2, True, npl, cache)
        assert torch.equal(a, b) and torch.equal(cache.planes_for(b, npl).buf, ops.pack_planes([a], npl).buf)
        # positions of a down-sampled level
        pos = torch.randn(n, 3, generator=g).cuda()
        batch = torch.zeros(n, dtype=torch.int64).cuda()
        sel = torch.arange(0, n, 3).cuda()
        cache = ForwardCache()
        pq, _, _ = ops.subset_gather(pos, batch, None, sel, npl, cache)
        assert torch.equal(pq, pos[sel]) and torch.equal(cache.planes_for(pq, npl).buf, ops.pack_planes([pq], npl).buf)
    # widths the float4 / 16-column planes path does not cover fall back to the plain kernel and register nothing
    cache = ForwardCache()
    odd = torch.randn(n * (K + 1), 24, generator=g).cuda()
    assert torch.equal(ops.group_max_dense(odd, idx, 2, cache), ops.group_max_dense(odd, idx)) and not cache.planes
