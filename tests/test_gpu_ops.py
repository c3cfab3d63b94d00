"""GPU parity of the op-level kernels (through the C-ABI) against the golden vectors produced by the real
reference sources (oracle/make_golden.py) and against the CPU oracle on fresh seeded inputs.
Integer / index results: bit-exact.  Float results: relative 1e-4 (BASELINE.json north_star)."""
import pytest
import torch

from golden_utils import FixedDraws, load, rel_err

pytestmark = pytest.mark.gpu

OPS = load("ops.pt")
CASES = [c for c in OPS if c != "get_curves"]
TOL = 1e-4


def dev(t):
    return None if t is None else t.cuda()


def _local_dense(po, case):
    g = OPS[case]
    return g, dev(g["pos"]), dev(g["batch"]), dev(g["curve"])


@pytest.mark.parametrize("case", CASES)
def test_csr_helpers(case):
    from curvecloudnet_b200 import point_ops as po
    g, pos, batch, curve = _local_dense(po, case)
    assert torch.equal(po.curveidx_local2global(curve, batch).cpu(), g["glob"])
    assert torch.equal(po.batch2ptr(batch).cpu(), g["ptr"])
    assert torch.equal(po.batch2ptr(batch, True).cpu(), g["ptr_ends"])
    padded, mask = po.to_batch_padded(pos, batch)
    assert torch.equal(mask.cpu(), g["padded_mask"])
    assert torch.allclose(padded.sum(dim=1).cpu(), g["padded_sum"], rtol=1e-5, atol=1e-5)


@pytest.mark.parametrize("case", CASES)
def test_samplers(case):
    from curvecloudnet_b200 import point_ops as po
    g, pos, batch, curve = _local_dense(po, case)
    idx = po.CurveFPS(g["curvefps_arclen"])(pos, batch, curve, draws=FixedDraws(u=g["curvefps_u"]))
    assert torch.equal(idx.cpu(), g["curvefps_idx"])
    idx = po.VoxelFPS(g["voxel_size"])(pos, batch, draws=FixedDraws(noise=g["voxel_noise"]))
    assert torch.equal(idx.cpu(), g["voxel_idx"])
    idx = po.fps_pytorch3d(pos, batch, g["fps_ratio"], draws=FixedDraws(starts=g["fps_starts"].tolist()))
    assert torch.equal(idx.cpu(), g["fps_idx"])


@pytest.mark.parametrize("case", CASES)
def test_groupings(case):
    from curvecloudnet_b200 import point_ops as po
    g, pos, batch, curve = _local_dense(po, case)
    idx, sub = dev(g["curvefps_idx"]), dev(g["voxel_idx"])
    row, col = po.radius_1d_group_subset(pos, idx, curve, batch, g["r1d_radius"])
    assert torch.equal(row.cpu(), g["r1d_row"]) and torch.equal(col.cpu(), g["r1d_col"])
    row, col = po.knn_1d_group_superset(pos, idx, curve, batch, 3)
    assert torch.equal(row.cpu(), g["s1d_row"]) and torch.equal(col.cpu(), g["s1d_col"])
    r = g["knn_radius"]
    psub, bsub = pos[sub].contiguous(), batch[sub].contiguous()
    dense, _, _ = po.knn_ball_group_pytorch3d(psub, pos, bsub, batch, "knn", radius=r, knn=16, return_dense=True)
    assert torch.equal(dense.cpu(), g["frnn_dense"])
    row, col = po.knn_ball_group_pytorch3d(psub, pos, bsub, batch, "knn", radius=r, knn=16)
    assert torch.equal(row.cpu(), g["frnn_row"]) and torch.equal(col.cpu(), g["frnn_col"])
    row, col = po.knn_ball_group_pytorch3d(psub, pos, bsub, batch, "ball-group", radius=r)
    assert torch.equal(row.cpu(), g["ball_row"]) and torch.equal(col.cpu(), g["ball_col"])
    row, col = po.knn_ball_group_pytorch3d(pos, psub, batch, bsub, "knn", knn=3, accel_knn=False)
    assert torch.equal(row.cpu(), g["knn3_row"]) and torch.equal(col.cpu(), g["knn3_col"])


@pytest.mark.parametrize("case", CASES)
def test_float_ops(case):
    from curvecloudnet_b200 import modules as M
    from curvecloudnet_b200 import point_ops as po
    from curvecloudnet_b200.synthetic import randomize_model_
    g, pos, batch, curve = _local_dense(po, case)
    idx, sub = dev(g["curvefps_idx"]), dev(g["voxel_idx"])
    feat = dev(g["feat"])
    assert rel_err(po.compute_feature_diffs(feat, curve, batch).cpu(), g["feat_diffs"]) <= 1e-6
    psub, bsub = pos[sub].contiguous(), batch[sub].contiguous()
    y = po.knn_interpolate_pytorch3d(dev(g["feat_sub"]), psub, pos, bsub, batch, k=3)
    assert rel_err(y.cpu(), g["interp3"]) <= TOL
    y = po.knn_interpolate_1D_pytorch3d(dev(g["feat_idx"]), idx, pos, batch, curve, 3)
    assert rel_err(y.cpu(), g["interp1d"]) <= TOL
    for ver, cls in (("v2", M.SymmetricCurve1DConvV2), ("v1", M.SymmetricCurve1DConvFastV1)):
        torch.manual_seed(3)
        mod = cls([9, 16, 12], g["conv_%s_k" % ver], with_xyz=True, with_diff=True)
        randomize_model_(mod, 55)
        mod = mod.cuda().eval()
        out = mod(feat, pos, batch, curve)[0]
        assert rel_err(out.cpu(), g["conv_%s_out" % ver]) <= TOL, ver


def test_linear_matches_torch_fp32():
    """The fp32 Linear kernel against plain torch fp32 (float64 accumulation as the yardstick)."""
    from curvecloudnet_b200 import ops
    gen = torch.Generator().manual_seed(0)
    for (m, n, k, act) in [(1, 5, 3, 0), (130, 64, 67, 1), (777, 129, 259, 2), (40000, 64, 38, 1),
                           (300, 1024, 2051, 1), (5000, 256, 256, 0)]:
        a = torch.randn(m, k, generator=gen)
        w = torch.randn(n, k, generator=gen) / k ** 0.5
        b = torch.randn(n, generator=gen)
        ref = a.double() @ w.double().t() + b.double()
        if act == 1:
            ref = ref.clamp(min=0)
        elif act == 2:
            ref = torch.where(ref > 0, ref, ref * 0.01)
        y = ops.linear(a.cuda(), w.cuda(), b.cuda(), act, 0.01).cpu()
        assert rel_err(y, ref.float()) <= 2e-6, (m, n, k, act)
        # strided A (a column slice) and no bias
        a2 = torch.randn(m, k + 5, generator=gen)
        y2 = ops.linear(a2.cuda()[:, 2:2 + k], w.cuda(), None).cpu()
        assert rel_err(y2, (a2[:, 2:2 + k].double() @ w.double().t()).float()) <= 2e-6


def test_random_clouds_against_oracle():
    """Fresh seeded inputs (not in the fixtures): neighbour search and samplers vs the CPU oracle, incl.
    duplicate points (distance ties -> lowest index), K larger than the cloud and ragged batches."""
    from curvecloudnet_b200 import ops
    from oracle import tp_ops as tp
    gen = torch.Generator().manual_seed(5)
    lens2, lens1 = [700, 1, 333, 40], [50, 3, 120, 7]
    p2 = torch.rand(sum(lens2), 3, generator=gen)
    p2[10:20] = p2[0:10]                      # exact duplicates
    p2 = (p2 * 64).round() / 64                 # many exact distance ties
    p1 = (torch.rand(sum(lens1), 3, generator=gen) * 64).round() / 64
    b2 = torch.cat([torch.full((l,), i, dtype=torch.int64) for i, l in enumerate(lens2)])
    b1 = torch.cat([torch.full((l,), i, dtype=torch.int64) for i, l in enumerate(lens1)])
    s2 = ops.sample_ptr(b2.cuda())
    off2 = [0] + list(torch.tensor(lens2).cumsum(0))
    P1, P2 = max(lens1), max(lens2)
    pp1, pp2 = torch.zeros(4, P1, 3), torch.zeros(4, P2, 3)
    o1 = 0
    for b in range(4):
        pp1[b, :lens1[b]] = p1[o1:o1 + lens1[b]]
        pp2[b, :lens2[b]] = p2[off2[b]:off2[b] + lens2[b]]
        o1 += lens1[b]
    l1t, l2t = torch.tensor(lens1), torch.tensor(lens2)

    def flat(dense, fill_local):
        rows = []
        for b in range(4):
            blk = dense[b, :lens1[b]].clone()
            blk = torch.where(blk >= 0, blk + off2[b], blk) if fill_local else blk + off2[b]
            rows.append(blk)
        return torch.cat(rows)

    for K, r in [(8, 0.25), (32, 0.11), (1, 0.5)]:
        _, want = tp.frnn_grid_points(pp1, pp2, l1t, l2t, K, r)
        got, _, cnt = ops.frnn_knn(p1.cuda(), b1.cuda(), p2.cuda(), b2.cuda(), s2, K, r, want_cnt=True)
        assert torch.equal(got.cpu().long(), flat(want, True)), (K, r)
        assert torch.equal(cnt[:p1.shape[0]].cpu().long(), (flat(want, True) >= 0).sum(1))
    for K in (3, 20):
        _, want = tp.knn_points(pp1, pp2, l1t, l2t, K)
        got, _ = ops.knn_exact(p1.cuda(), b1.cuda(), p2.cuda(), s2, K, zero_fill=True)
        assert torch.equal(got.cpu().long(), flat(want, False)), K
    want = tp.ball_query(pp1, pp2, l1t, l2t, 16, 0.2)
    got, _ = ops.ball_query(p1.cuda(), b1.cuda(), p2.cuda(), s2, 0.2, 16)
    assert torch.equal(got.cpu().long(), flat(want, True))
    # FPS with ties, every start, K > len
    idx = tp.sample_farthest_points(pp2, l2t, torch.tensor([70, 5, 333, 13]), [3, 0, 100, 39])
    want = torch.cat([idx[b][idx[b] >= 0] + off2[b] for b in range(4)]).sort()[0]
    got = ops.fps(p2.cuda(), s2, [70, 5, 333, 13], [3, 0, 100, 39])
    assert torch.equal(got.cpu(), want)


def test_no_cpu_path():
    from curvecloudnet_b200 import native, ops
    with pytest.raises((native.CcnError, TypeError, RuntimeError)):
        ops.linear(torch.zeros(4, 4), torch.zeros(4, 4), None)


@pytest.mark.parametrize("mode,tol", [(3, 2e-5), (1, 1e-2)])
def test_linear_tensor_core(mode, tol):
    """tcgen05 Linear (ccn_linear_tc) vs float64: x3 split-bf16 is fp32-class, plain bf16 within its budget."""
    from curvecloudnet_b200 import ops
    gen = torch.Generator().manual_seed(1)
    shapes = [(1, 5, 3, 0), (130, 64, 67, 1), (777, 129, 259, 2), (40000, 64, 38, 1), (300, 1024, 2051, 1),
              (5000, 256, 256, 0), (129, 17, 64, 0), (4096, 192, 128, 2), (257, 32, 640, 1)]
    for (m, n, k, act) in shapes:
        a = torch.randn(m, k, generator=gen)
        w = torch.randn(n, k, generator=gen) / k ** 0.5
        b = torch.randn(n, generator=gen)
        ref = a.double() @ w.double().t() + b.double()
        if act == 1:
            ref = ref.clamp(min=0)
        elif act == 2:
            ref = torch.where(ref > 0, ref, ref * 0.01)
        tw = ops.prepare_tc_weight(w.cuda())
        y = ops.linear_tc(a.cuda(), tw, b.cuda(), act, 0.01, mode).cpu()
        assert torch.isfinite(y).all(), (m, n, k, act)
        assert rel_err(y, ref.float()) <= tol, (m, n, k, act, rel_err(y, ref.float()))
        a2 = torch.randn(m, k + 5, generator=gen)          # unaligned, strided A; no bias
        y2 = ops.linear_tc(a2.cuda()[:, 2:2 + k], tw, None, 0, 0.01, mode).cpu()
        assert rel_err(y2, (a2[:, 2:2 + k].double() @ w.double().t()).float()) <= tol, (m, n, k, "strided")


@pytest.mark.parametrize("K", [1, 2, 3, 4])
def test_small_k_exact_knn_equals_warp_kernel_and_oracle(K):
    """ccn_knn_exact has two kernels (K <= 4: 8 lanes per query over smem tiles; K > 4: warp per query); both must
    give the oracle's (d2, index) total order, including duplicate points, ragged clouds and clouds shorter than K."""
    from curvecloudnet_b200 import ops
    from oracle import tp_ops as tp
    g = torch.Generator().manual_seed(40 + K)
    lens2 = [700, 2, 2500, 0, 40]
    lens1 = [300, 5, 1000, 3, 77]
    p2 = torch.rand(sum(lens2), 3, generator=g)
    p2[10:20] = p2[0:10]                                     # exact duplicates -> index tie-break
    p1 = torch.rand(sum(lens1), 3, generator=g)
    p1[0:5] = p2[0:5]
    b1 = torch.cat([torch.full((n,), i, dtype=torch.int64) for i, n in enumerate(lens1)])
    b2 = torch.cat([torch.full((n,), i, dtype=torch.int64) for i, n in enumerate(lens2)])
    host = [0]
    for n in lens2:
        host.append(host[-1] + n)
    sp = ops.SamplePtr(torch.tensor(host, dtype=torch.int32).cuda(), host)
    idx, d2, cnt = ops.knn_exact(p1.cuda(), b1.cuda(), p2.cuda(), sp, K, zero_fill=False, want_d2=True, want_cnt=True)
    idx5, d25 = ops.knn_exact(p1.cuda(), b1.cuda(), p2.cuda(), sp, 5, zero_fill=False, want_d2=True)
    assert torch.equal(idx.cpu(), idx5.cpu()[:, :K])
    valid = idx.cpu() >= 0
    assert torch.equal(d2.cpu()[valid], d25.cpu()[:, :K][valid])
    # oracle (pytorch3d.knn_points restatement) on padded clouds
    off1, off2 = 0, 0
    for n1_, n2_ in zip(lens1, lens2):
        if n1_ and n2_:
            od, oi = tp.knn_points(p1[off1:off1 + n1_][None], p2[off2:off2 + n2_][None], torch.tensor([n1_]), torch.tensor([n2_]), K)
            kk = min(K, n2_)
            got = idx.cpu()[off1:off1 + n1_, :kk].long() - off2
            assert torch.equal(got, oi[0, :, :kk])
            assert torch.equal(cnt.cpu()[off1:off1 + n1_].long(), torch.full((n1_,), kk))
        elif n1_:
            assert bool((idx.cpu()[off1:off1 + n1_] == -1).all())
        off1 += n1_
        off2 += n2_


# ---------------------------------------------------------------------------------------------------------------
# a3: curve grouping of a raw sweep (_get_curves) - bit-exact against the golden vector of the REAL reference method
# and against the oracle on full synthetic sweeps (32 and 64 beams), empty / single-point / single-beam inputs
# ---------------------------------------------------------------------------------------------------------------
def test_get_curves_matches_reference_golden():
    from curvecloudnet_b200 import ops
    g = OPS["get_curves"]
    reorder, inv, cid, pts, counts = ops.get_curves(g["points"].cuda(), g["beam"].cuda(), 0.08, 32)
    assert counts.tolist()[1] == 0
    assert torch.equal(inv.cpu(), g["inv_reorder"])
    assert torch.equal(cid.cpu(), g["curve_idxs"])
    assert torch.equal(pts.cpu(), g["sorted_points"])
    assert counts.tolist()[0] == int(g["curve_idxs"].max()) + 1


@pytest.mark.parametrize("seed,beams,az", [(5, 32, 1085), (6, 64, 1875)])
def test_get_curves_full_sweep_matches_oracle(seed, beams, az):
    from curvecloudnet_b200 import point_ops
    from curvecloudnet_b200.synthetic import raycast_driving_scan
    from oracle import ccn_point_ops as po
    raw = raycast_driving_scan(seed, beams=beams, azimuth_steps=az)
    pts = torch.from_numpy(raw.points)
    beam = torch.from_numpy(raw.beam_ids).to(torch.int64)
    refl = torch.from_numpy(raw.reflectance)[:, None].float()
    labels = torch.arange(pts.shape[0], dtype=torch.int64)
    reorder, inv, cid = po.get_curves(pts, beam, 0.08)
    p2, cid2, lab2, ref2, inv2 = point_ops.get_curves(pts.cuda(), beam.cuda(), labels.cuda(), refl.cuda())
    assert torch.equal(inv2.cpu(), inv) and torch.equal(cid2.cpu(), cid)
    assert torch.equal(p2.cpu(), pts[reorder]) and torch.equal(lab2.cpu(), labels[reorder])
    assert torch.equal(ref2.cpu(), refl[reorder])
    # size-independent properties: reorder is a permutation sorted by beam, stable inside a beam, ids are non-decreasing
    b_sorted = beam[lab2.cpu()]
    assert bool(torch.all(b_sorted[1:] >= b_sorted[:-1]))
    same = b_sorted[1:] == b_sorted[:-1]
    assert bool(torch.all(lab2.cpu()[1:][same] > lab2.cpu()[:-1][same]))
    assert bool(torch.all(cid2.cpu()[1:] - cid2.cpu()[:-1] >= 0)) and int(cid2[0]) == 0


def test_get_curves_edge_cases():
    from curvecloudnet_b200 import ops
    from oracle import ccn_point_ops as po
    for n, nb in ((0, 4), (1, 4), (300, 1)):
        gen = torch.Generator().manual_seed(n)
        pts = torch.rand(n, 3, generator=gen) * 20
        beam = torch.randint(0, nb, (n,), generator=gen)
        reorder, inv, cid, sp, counts = ops.get_curves(pts.cuda(), beam.cuda(), 0.08, nb)
        if n:
            r2, i2, c2 = po.get_curves(pts, beam, 0.08)
            assert torch.equal(inv.cpu(), i2) and torch.equal(cid.cpu(), c2) and torch.equal(reorder.cpu(), r2)
        assert counts.tolist()[1] == 0
    bad = ops.get_curves(torch.rand(10, 3).cuda(), torch.full((10,), 9).cuda(), 0.08, 4)[4]
    assert bad.tolist()[1] == 1


@pytest.mark.parametrize("precision", ["fp32", "bf16x3"])
@pytest.mark.parametrize("aggr", ["max", "mean", "attend", "weighted-sum"])
def test_dense_sgcnn_aggregations_match_oracle(aggr, precision):
    """StaticEdgeConv.forward_fast with every aggregation of dgcnn.py:180-203 (the shipped configs use 'max' only): our
    SGCNNLayer against the oracle's restatement, same weights, B = 2 so that padded rows and a sample boundary occur."""
    from curvecloudnet_b200 import modules as M
    from curvecloudnet_b200 import nn as N
    from curvecloudnet_b200.synthetic import driving_scan_batch, randomize_model_
    from oracle import ccn_modules as OM
    from oracle.tp_nn import MLP as OMLP
    data = driving_scan_batch([1, 2], subsample=900)
    gen = torch.Generator().manual_seed(4)
    x = torch.randn(data.pos.shape[0], 13, generator=gen)
    dims = [2 * (13 + 3), 64, 64]
    att = aggr in ("attend", "weighted-sum")
    want_mod = OM.SGCNNLayer(OMLP(dims), 9, r=0.12, with_xyz=True, aggr_type=aggr,
                             attend_nn=OMLP([64, 64, 64], act="leaky_relu") if att else None)
    randomize_model_(want_mod, 77)
    want_mod.eval()
    with torch.no_grad():
        want = want_mod(x, data.pos, data.batch)[0]
    mod = M.SGCNNLayer(N.MLP(dims), 9, r=0.12, with_xyz=True, aggr_type=aggr,
                       attend_nn=N.MLP([64, 64, 64], act="leaky_relu") if att else None)
    mod.load_state_dict(want_mod.state_dict(), strict=True)
    mod = mod.cuda().eval()
    N.set_precision(precision)
    try:
        with torch.no_grad():
            got = mod(x.cuda(), data.pos.cuda(), data.batch.cuda(), cache=N.ForwardCache())[0]
    finally:
        N.set_precision("fp32")
    assert got.shape == want.shape
    assert rel_err(got.cpu(), want) <= (1e-4 if precision == "fp32" else 3e-4), aggr


@pytest.mark.parametrize("K", [1, 3, 4])
def test_knn_grid_equals_brute_force(K):
    """ccn_knn_grid (hashed grid + shells + brute-force pass for the open queries) against ccn_knn_exact and the oracle's
    knn_points: identical indices and distances - surface-like clouds, duplicates, clouds smaller than K (zero fill),
    queries far outside the sources, ragged batches."""
    from curvecloudnet_b200 import ops
    from oracle import tp_ops as tp
    gen = torch.Generator().manual_seed(K)
    for trial in range(4):
        lens2 = [3000, 2, 1500][:1 + trial % 3] if trial else [5000]
        lens1 = [int(l * 1.7) + 5 for l in lens2]
        p2 = torch.cat([torch.rand(l, 3, generator=gen) * torch.tensor([6.0, 6.0, 0.3]) for l in lens2])
        p1 = torch.cat([torch.rand(l, 3, generator=gen) * torch.tensor([9.0, 6.0, 0.3]) - 1.0 for l in lens1])
        if trial == 1:
            p2[100:200] = p2[0:100]                       # duplicates: ties by lowest index
            p2 = (p2 * 16).round() / 16
        b2 = torch.repeat_interleave(torch.arange(len(lens2)), torch.tensor(lens2))
        b1 = torch.repeat_interleave(torch.arange(len(lens1)), torch.tensor(lens1))
        s2 = ops.sample_ptr(b2.cuda())
        a = ops.knn_exact(p1.cuda(), b1.cuda(), p2.cuda(), s2, K, use_grid=False, want_cnt=True)
        g = ops.knn_exact(p1.cuda(), b1.cuda(), p2.cuda(), s2, K, batch2=b2.cuda(), use_grid=True, want_cnt=True)
        assert torch.equal(a[0], g[0]) and torch.equal(a[1], g[1]) and torch.equal(a[2][:p1.shape[0]], g[2][:p1.shape[0]]), trial
        # the oracle, cloud by cloud (global indices, zero fill -> first point of the cloud)
        off1 = off2 = 0
        for l1, l2 in zip(lens1, lens2):
            dd, ii = tp.knn_points(p1[off1:off1 + l1][None], p2[off2:off2 + l2][None], torch.tensor([l1]), torch.tensor([l2]), K)
            assert torch.equal(g[0][off1:off1 + l1].cpu().long(), ii[0] + off2), trial
            off1 += l1
            off2 += l2


def _ragged_segments(nq, max_deg, gen):
    """CSR rows with empty segments, single-row segments and a long one; returns (row_ptr i32, erow i64)."""
    deg = torch.randint(0, max_deg + 1, (nq,), generator=gen)
    deg[0], deg[nq // 2], deg[nq - 1] = 0, 1, 4 * max_deg          # an empty first, a single, a long last segment
    row_ptr = torch.zeros(nq + 1, dtype=torch.int64)
    row_ptr[1:] = deg.cumsum(0)
    erow = torch.repeat_interleave(torch.arange(nq), deg)
    return row_ptr.int(), erow


@pytest.mark.parametrize("c", [7, 20, 64])                          # scalar kernel, float4 kernel, float4 + planes
@pytest.mark.parametrize("mode", ["max", "mean", "sum", "attend", "weighted-sum", "weighted-sum-norm"])
def test_seg_aggregate_against_oracle_scatter(mode, c):
    """ccn_seg_aggregate(_both) against the oracle's torch_scatter / torch_geometric restatements on a ragged edge list
    (point_conv.py:81-92, dgcnn.py:190-197, 240-244): empty segments give 0, 'attend' is softmax over the target's rows
    with the 1e-16 of torch_geometric.utils.softmax.  Sums run in row order on both sides; fp32 tolerance 1e-5."""
    from curvecloudnet_b200 import nn as N
    from curvecloudnet_b200 import ops
    from oracle import tp_ops as T
    gen = torch.Generator().manual_seed(100 + c)
    nq = 301
    row_ptr, erow = _ragged_segments(nq, 9, gen)
    E = erow.shape[0]
    msg = torch.randn(E, c, generator=gen)
    att = 3.0 * torch.randn(E, c, generator=gen)
    if mode == "max":
        want = T.scatter_max(msg, erow, 0, None, nq)[0]
    elif mode == "mean":
        want = T.scatter_mean(msg, erow, 0, None, nq)
    elif mode == "sum":
        want = T.scatter_add(msg, erow, 0, None, nq)
    elif mode == "attend":
        want = T.scatter_add(msg * T.segment_softmax(att, erow, nq), erow, 0, None, nq)
    elif mode == "weighted-sum":
        want = T.scatter_add(msg * torch.sigmoid(att), erow, 0, None, nq)
    else:
        w = torch.sigmoid(att)
        tot = T.scatter_add(w, erow, 0, None, nq).clamp(min=1e-3)
        want = T.scatter_add(msg * (w / tot[erow]), erow, 0, None, nq)
    a = att.cuda() if mode in ("attend", "weighted-sum", "weighted-sum-norm") else None
    got = ops.seg_aggregate(msg.cuda(), a, row_ptr.cuda(), nq, mode)
    assert torch.equal(got[0].cpu(), torch.zeros(c)), "empty segment"
    if mode == "max":
        assert torch.equal(got.cpu(), want)
    else:
        assert rel_err(got.cpu(), want) <= 1e-5, mode
    if c % 16 == 0:
        cache = N.ForwardCache()
        both = ops.seg_aggregate(msg.cuda(), a, row_ptr.cuda(), nq, mode, npl=2, cache=cache)
        assert torch.equal(both, got)
        pl = cache.planes_for(both, 2)
        assert torch.equal(pl.buf, ops.pack_planes([got], 2).buf)
    # a column slice of a wider buffer (leading dimension > c) gives the same rows
    wide = torch.zeros(E, c + 12)
    wide[:, 4:4 + c] = msg
    got_ld = ops.seg_aggregate(wide.cuda()[:, 4:4 + c], a, row_ptr.cuda(), nq, mode)
    assert torch.equal(got_ld, got)


@pytest.mark.parametrize("c", [6, 64])
def test_group_max_dense_and_interp_blend_against_formula(c):
    """ccn_group_max_dense: max over the K+1 rows of every point with the rows of missing neighbours (idx < 0) counted as
    -100 (dgcnn.py:187-189).  ccn_interp_blend: sum w x / sum w with w = 1 / clamp(d2, 1e-16) (point_ops.py:330-341) or
    the given weights, missing neighbours skipped.  Both against plain fp32 torch on the CPU."""
    from curvecloudnet_b200 import ops
    gen = torch.Generator().manual_seed(7 + c)
    n, K = 517, 9
    idx = torch.randint(0, n, (n, K), generator=gen).int()
    idx[torch.rand(n, K, generator=gen) < 0.3] = -1
    idx[5] = -1
    feat = 40.0 * torch.randn(n * (K + 1), c, generator=gen) - 90.0       # many rows below -100: the mask value matters
    f3 = feat.view(n, K + 1, c).clone()
    f3[:, 1:][idx < 0] = -100.0
    want = f3.max(dim=1)[0]
    got = ops.group_max_dense(feat.cuda(), idx.cuda())
    assert torch.equal(got.cpu(), want)

    ns, knn = 133, 3
    x = torch.randn(ns, c, generator=gen)
    nbr = torch.randint(0, ns, (n, knn), generator=gen).int()
    nbr[torch.rand(n, knn, generator=gen) < 0.2] = -1
    nbr[:, 0].clamp_(min=0)                                               # at least one neighbour per row
    d2 = torch.rand(n, knn, generator=gen) ** 4
    d2[3, 0] = 0.0                                                        # coincident point: the 1e-16 clamp
    valid = (nbr >= 0).float()
    for from_d2, given in ((True, d2), (False, d2 + 0.1)):
        w = (1.0 / given.clamp(min=1e-16) if from_d2 else given) * valid
        xs = x[nbr.clamp(min=0).long()]                                   # (n, knn, c)
        num = torch.zeros(n, c)
        den = torch.zeros(n)
        for s in range(knn):                                              # the kernel's order of additions
            num = num + xs[:, s] * w[:, s:s + 1]
            den = den + w[:, s]
        want = num / den[:, None]
        got = ops.interp_blend(x.cuda(), nbr.cuda(), given.cuda(), from_d2)
        assert rel_err(got.cpu(), want) <= 1e-6, from_d2


def test_unused_grouping_helpers_match_reference():
    """knn_1d_group_subset (point_ops.py:114-140) and group_first_k_values (:384-428) - called by no shipped configuration -
    against the outputs of the REAL functions (tests/golden/helpers.pt): index results bit-exact, values bit-exact (copies);
    plus the layout helpers dense2padded_pyg / pytorch3didx2pyg / is_sorted of the same file."""
    from curvecloudnet_b200 import point_ops as po
    g = load("helpers.pt")
    pos, batch, curve = dev(g["pos"]), dev(g["batch"]), dev(g["curve"])
    for knn, rec in g["knn_1d"].items():
        row, col = po.knn_1d_group_subset(pos, dev(g["idxs"]), curve, batch, int(knn))
        assert row.dtype == torch.int64 and torch.equal(row.cpu(), rec["row"]) and torch.equal(col.cpu(), rec["col"]), knn
    ids = dev(g["first_k_ids"])
    for tag, rec in g["first_k"].items():
        vals = dev(g["first_k_v1"] if tag.startswith("v1") else g["first_k_v2"])
        k = dev(rec["k"]) if torch.is_tensor(rec["k"]) else rec["k"]
        grouped, mask = po.group_first_k_values(vals, ids, k)
        assert grouped.shape == rec["grouped"].shape, tag
        assert torch.equal(mask.cpu(), rec["mask"]) and torch.equal(grouped.cpu(), rec["grouped"]), tag
    xp, mask, lengths, offsets = po.dense2padded_pyg(pos, batch)
    lens = torch.bincount(g["batch"])
    assert xp.shape == (2, int(lens.max()), 3) and torch.equal(lengths.cpu(), lens) and torch.equal(offsets.cpu(), lens.cumsum(0)[:-1])
    assert torch.equal(xp[mask].cpu(), g["pos"]) and float(xp[~mask].abs().sum()) == 0.0
    local = torch.arange(int(lens.max())).view(1, -1, 1).repeat(2, 1, 1).cuda()
    flat = po.pytorch3didx2pyg(local, mask.unsqueeze(-1), offsets)
    assert torch.equal(flat.cpu(), torch.arange(g["pos"].shape[0]))
    assert bool(po.is_sorted(batch)) and bool(po.is_sorted(torch.tensor([3, 1, 2]).cuda()))     # upstream quirk: always True
