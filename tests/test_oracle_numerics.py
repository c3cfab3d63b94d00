"""Pins for the explicit float formulas the oracle uses where an index depends on them
(oracle/ccn_point_ops.py header) and property tests of the third-party restatements
against naive loops."""
import numpy as np
import pytest
import torch

from oracle import ccn_point_ops as po
from oracle import tp_ops as tp


def test_norm3_is_torch_linalg_norm():
    g = torch.Generator().manual_seed(0)
    v = torch.randn(50000, 3, generator=g) * torch.rand(50000, 1, generator=g) * 3
    assert torch.equal(po.norm3(v), torch.linalg.norm(v, dim=-1))


def test_cumsum_is_sequential_f64():
    g = torch.Generator().manual_seed(1)
    v = torch.rand(100000, generator=g) * 0.01
    assert torch.equal(po.cumsum_f32(v), torch.cumsum(v, 0))
    assert torch.equal(torch.cumsum(v, 0), torch.from_numpy(np.cumsum(v.double().numpy())).float())


def test_scalar_over_tensor_is_reciprocal_times_scalar():
    g = torch.Generator().manual_seed(2)
    t = torch.rand(10000, generator=g) * 0.01 + 1e-4
    assert torch.equal(0.02 / t, torch.reciprocal(t) * 0.02)


def test_fps_count_is_float32():
    # 3000 * 0.3 evaluates to 900.00006 in f32 -> 901 (point_ops.py:63)
    assert int(po.fps_count(torch.tensor([3000]), 0.3)) == int(np.ceil(np.float32(3000) * np.float32(0.3)))


def _naive_knn(p1, p2, K, r2=None, strict=False):
    out = []
    for a in p1:
        d = [float(((a[0] - b[0]) * (a[0] - b[0]) + (a[1] - b[1]) * (a[1] - b[1])) + (a[2] - b[2]) * (a[2] - b[2]))
             for b in p2]
        order = sorted(range(len(p2)), key=lambda j: (d[j], j))
        if r2 is not None:
            order = [j for j in order if (d[j] < r2 if strict else d[j] <= r2)]
        out.append((order[:K] + [-1] * K)[:K])
    return torch.tensor(out)


@pytest.mark.parametrize("seed", [0, 1])
def test_frnn_contract_vs_naive(seed):
    g = torch.Generator().manual_seed(seed)
    p2 = torch.rand(150, 3, generator=g)
    p2[10] = p2[3]                      # duplicate -> distance tie -> lowest index first
    p1 = p2[:40].clone()
    r = 0.3
    _, idx = tp.frnn_grid_points(p1[None], p2[None], torch.tensor([40]), torch.tensor([150]), 8, r)
    r2 = float(torch.tensor(r, dtype=torch.float32) * torch.tensor(r, dtype=torch.float32))
    assert torch.equal(idx[0], _naive_knn(p1.numpy().astype(np.float32), p2.numpy().astype(np.float32), 8, r2))


def test_knn_points_and_ball_query_vs_naive():
    g = torch.Generator().manual_seed(3)
    p2 = torch.rand(120, 3, generator=g)
    p1 = torch.rand(30, 3, generator=g)
    _, idx = tp.knn_points(p1[None], p2[None], torch.tensor([30]), torch.tensor([120]), 5)
    assert torch.equal(idx[0], _naive_knn(p1.numpy(), p2.numpy(), 5))
    bq = tp.ball_query(p1[None], p2[None], torch.tensor([30]), torch.tensor([120]), 6, 0.3)[0]
    r2 = np.float32(0.3) * np.float32(0.3)
    for i in range(30):
        d = ((p1[i, 0] - p2[:, 0]) ** 2 + (p1[i, 1] - p2[:, 1]) ** 2) + (p1[i, 2] - p2[:, 2]) ** 2
        want = [j for j in range(120) if d[j] < r2][:6]
        assert bq[i].tolist() == (want + [-1] * 6)[:6]


def test_fps_vs_naive():
    g = torch.Generator().manual_seed(4)
    p = torch.rand(1, 200, 3, generator=g)
    idx = tp.sample_farthest_points(p, torch.tensor([200]), torch.tensor([20]), [7])[0].tolist()
    sel, closest = [7], np.full(200, np.inf, dtype=np.float32)
    pn = p[0].numpy()
    for _ in range(19):
        q = pn[sel[-1]]
        d = ((pn[:, 0] - q[0]) ** 2 + (pn[:, 1] - q[1]) ** 2) + (pn[:, 2] - q[2]) ** 2
        closest = np.minimum(closest, d.astype(np.float32))
        sel.append(int(np.argmax(closest)))
    assert idx == sel


def test_scatter_semantics():
    src = torch.tensor([[1.0], [5.0], [5.0], [-2.0]])
    index = torch.tensor([0, 2, 2, 0])
    val, arg = tp.scatter_max(src, index, 0, None, 4)
    assert val.flatten().tolist() == [1.0, 0.0, 5.0, 0.0] and arg.flatten().tolist() == [0, 4, 1, 4]
    val, arg = tp.scatter_min(src, index, 0, None, 3)
    assert val.flatten().tolist() == [-2.0, 0.0, 5.0] and arg.flatten().tolist() == [3, 4, 1]
    assert tp.scatter_mean(src, index, 0, None, 3).flatten().tolist() == [-0.5, 0.0, 5.0]
    sm = tp.segment_softmax(torch.tensor([[0.0], [1.0], [1.0], [0.0]]), index)
    assert torch.allclose(sm.flatten(), torch.tensor([0.5, 0.5, 0.5, 0.5]))


def test_c_neighbour_search_equals_the_torch_statements():
    """oracle/nbr_search.c (grid-pruned, multi-threaded) against the O(P1*P2) torch statements of tp_ops it accelerates:
    FRNN (d2 <= r2, (d2, idx) order), ball_query (first K by index, strict), knn_points (exact, zero fill) - bit-exact,
    incl. duplicate points, quantised coordinates (distance ties), queries outside the cloud and K > cloud size."""
    import os
    from oracle import tp_ops
    so = os.path.join(os.path.dirname(os.path.abspath(tp_ops.__file__)), "_build", "liboracle_nbr.so")
    if not os.path.exists(so):
        pytest.skip("oracle/_build/liboracle_nbr.so not built (make -C oracle)")
    gen = torch.Generator().manual_seed(0)

    def both(fn, *a):
        prev = tp_ops.SEARCH_IMPL
        try:
            tp_ops.SEARCH_IMPL = "torch"
            x = fn(*a)
            tp_ops.SEARCH_IMPL = "c"
            y = fn(*a)
        finally:
            tp_ops.SEARCH_IMPL = prev
        return x, y

    for trial in range(4):
        n1, n2 = 700, 1500
        p2 = torch.rand(1, n2, 3, generator=gen) * torch.tensor([4.0, 4.0, 0.5])
        p2[0, 100:200] = p2[0, 0:100]
        if trial % 2:
            p2 = (p2 * 32).round() / 32
        p1 = p2[:, :n1].clone() if trial < 2 else torch.rand(1, n1, 3, generator=gen) * 5 - 0.5
        l1, l2 = torch.tensor([n1]), torch.tensor([n2])
        for K, r in ((20, 0.1), (32, 0.3), (5, 0.01)):
            a, b = both(tp_ops.frnn_grid_points, p1, p2, l1, l2, K, r)
            assert torch.equal(a[1], b[1]) and torch.equal(a[0], b[0]), (trial, K, r)
        for K in (3, 20):
            a, b = both(tp_ops.knn_points, p1, p2, l1, l2, K)
            assert torch.equal(a[1], b[1]) and torch.equal(a[0], b[0]), (trial, K)
        for K, r in ((128, 0.3), (16, 0.2)):
            a, b = both(tp_ops.ball_query, p1, p2, l1, l2, K, r)
            assert torch.equal(a, b), (trial, K, r)
    p2, p1 = torch.rand(1, 2, 3, generator=gen), torch.rand(1, 5, 3, generator=gen)
    a, b = both(tp_ops.knn_points, p1, p2, torch.tensor([5]), torch.tensor([2]), 3)
    assert torch.equal(a[1], b[1]) and torch.equal(a[0], b[0])
