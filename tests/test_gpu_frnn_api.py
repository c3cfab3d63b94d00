"""The two public FRNN entry points the reference calls by name (third_party/FRNN/frnn/frnn.py:210-225, 351-397; call sites
src/models/utils/point_ops.py:459, src/models/modules/dgcnn.py:172) as drop-ins of curvecloudnet_b200.frnn: same signature,
same (dists, idxs, nn, grid) result, checked against the oracle's statement of the result contract (tp_ops.frnn_grid_points,
itself pinned against the compiled reference FRNNBruteForceCPU) and against the reference-generated golden tables."""
import pytest
import torch

from golden_utils import load

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("K,r", [(8, 0.12), (20, 0.3), (32, 0.05)])
def test_frnn_grid_points_matches_oracle(K, r):
    from curvecloudnet_b200 import frnn
    from oracle import tp_ops as tp
    gen = torch.Generator().manual_seed(K)
    N, P1, P2 = 3, 400, 900
    p1 = torch.rand(N, P1, 3, generator=gen)
    p2 = torch.rand(N, P2, 3, generator=gen)
    p2[:, 50:80] = p2[:, 20:50]                               # duplicate points: distance ties -> lowest index
    l1 = torch.tensor([400, 250, 1])
    l2 = torch.tensor([900, 3, 600])
    want_d, want_i = tp.frnn_grid_points(p1, p2, l1, l2, K, r)
    dists, idxs, nn, grid = frnn.frnn_grid_points(p1.cuda(), p2.cuda(), l1.cuda(), l2.cuda(), K, r, return_nn=True)
    for b in range(N):
        n = int(l1[b])
        assert torch.equal(idxs[b, :n].cpu(), want_i[b, :n])
        assert torch.equal(dists[b, :n].cpu(), want_d[b, :n])
        assert bool((idxs[b, n:] == -1).all())
    want_nn = tp.frnn_gather(p2, want_i.clamp(min=-1))
    for b in range(N):
        n = int(l1[b])
        assert torch.equal(nn[b, :n].cpu(), want_nn[b, :n])
    # the grid handed back in: same result, no rebuild
    d2, i2, _, g2 = frnn.frnn_grid_points(p1.cuda(), p2.cuda(), l1.cuda(), l2.cuda(), K, r, grid=grid)
    assert torch.equal(i2, idxs) and torch.equal(d2, dists) and g2 is not None


def test_frnn_gather_matches_reference_semantics():
    from curvecloudnet_b200 import frnn
    gen = torch.Generator().manual_seed(3)
    x = torch.randn(2, 50, 7, generator=gen)
    idxs = torch.randint(-1, 50, (2, 30, 5), generator=gen)
    out = frnn.frnn_gather(x.cuda(), idxs.cuda())
    ref = x[:, :, None].expand(-1, -1, 5, -1).gather(1, idxs.clamp(min=0)[:, :, :, None].expand(-1, -1, -1, 7)).clone()
    ref[(idxs < 0)[:, :, :, None].expand(-1, -1, -1, 7)] = 0.0       # frnn.py:384-395
    assert torch.equal(out.cpu(), ref)


def test_frnn_grid_points_matches_reference_golden():
    """tests/golden/ops.pt `frnn_dense`: the table the REAL point_ops.knn_ball_group_pytorch3d produced over the shims."""
    from curvecloudnet_b200 import frnn
    from curvecloudnet_b200 import point_ops as po
    g = load("ops.pt")
    case = next(c for c in g if c != "get_curves")
    rec = g[case]
    pos, batch, sub = rec["pos"].cuda(), rec["batch"].cuda(), rec["voxel_idx"].cuda()
    psub, bsub = pos[sub].contiguous(), batch[sub].contiguous()
    p1, m1 = po.to_batch_padded(psub, bsub)
    p2, m2 = po.to_batch_padded(pos, batch)
    _, idxs, _, _ = frnn.frnn_grid_points(p1, p2, m1.sum(1), m2.sum(1), 16, rec["knn_radius"])
    want = rec["frnn_dense"]
    assert torch.equal(idxs[:, :want.shape[1]].cpu(), want)


@pytest.mark.parametrize("dim", [2, 5, 17])
def test_frnn_grid_points_other_dimensions(dim):
    """frnn_grid_points on D != 3 points (upstream's N-D kernels, grid.h:27-34; what DynamicEdgeConv feeds it): same result
    contract as D = 3 - d2 <= r^2, ascending by (d2, index), -1 padded - against the oracle's D-dimensional statement."""
    from curvecloudnet_b200 import frnn
    from oracle import tp_ops as T
    gen = torch.Generator().manual_seed(dim)
    lens = torch.tensor([61, 40])
    pts = 0.3 * torch.randn(2, 61, dim, generator=gen)
    pts[0, 7] = pts[0, 3]
    K, r = 9, 0.22 * dim ** 0.5
    want_d, want_i = T.frnn_grid_points(pts, pts, lens, lens, K, r)
    d, i, nn, grid = frnn.frnn_grid_points(pts.cuda(), pts.cuda(), lens.cuda(), lens.cuda(), K, r, return_nn=True)
    assert torch.equal(i.cpu(), want_i) and torch.equal(d.cpu(), want_d) and grid is None
    g = T.frnn_gather(pts, want_i, lens)
    assert torch.equal(nn.cpu(), g)
    with pytest.raises(ValueError, match="dimension mismatch"):
        frnn.frnn_grid_points(pts.cuda(), pts[:, :, :dim - 1].contiguous().cuda(), lens.cuda(), lens.cuda(), K, r)


def test_frnn_grid_points_radius_per_cloud():
    """r as an (N,) tensor (frnn.py:316-323; no reference call site passes one): every cloud searched with its own radius."""
    from curvecloudnet_b200 import frnn
    from oracle import tp_ops as T
    gen = torch.Generator().manual_seed(3)
    lens = torch.tensor([200, 150, 90])
    pts = torch.rand(3, 200, 3, generator=gen)
    r = torch.tensor([0.08, 0.2, 0.12])
    want_d, want_i = T.frnn_grid_points(pts, pts, lens, lens, 8, r)
    d, i, _, grid = frnn.frnn_grid_points(pts.cuda(), pts.cuda(), lens.cuda(), lens.cuda(), 8, r.cuda())
    assert torch.equal(i.cpu(), want_i) and torch.equal(d.cpu(), want_d) and grid is None
