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
