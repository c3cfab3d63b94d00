"""CPU tests of the scanning simulator's host logic and of its oracle (no GPU): the line end-point draws against the
reference's own (tests/golden/simulator.pt), the Rodrigues formula, the C rasteriser oracle on analytic views."""
import math
import os

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_host_draws_equal_the_reference():
    from curvecloudnet_b200.simulator import ScanLineGenerator
    for case in torch.load(os.path.join(ROOT, "tests", "golden", "simulator.pt"))["cases"]:
        B = len(case["kinds"])
        gen = ScanLineGenerator(case["res"], "cpu", "linear", case["direction"], case["density"])
        assert gen.ts == int(2 ** 0.5 * case["res"] * case["density"])
        np.random.seed(case["seed"])
        if case["direction"] == "random":
            got = gen._sample_linear_border_points(gen.NLINES, B)
            assert np.issubdtype(got.dtype, np.integer)
        elif case["direction"] == "parallel":
            got = gen._sample_parallel_border_points(0, gen.NLINES, B)
            assert np.array_equal(gen._sample_parallel_border_points(1, gen.NLINES, B), case["border1"].numpy())
        else:
            got = gen._sample_grid_border_points(0, gen.NLINES, B)
            assert np.array_equal(gen._sample_grid_border_points(1, gen.NLINES, B), case["border1"].numpy())
        assert got.shape == (gen.NLINES * B, 2, 2) and np.array_equal(got.astype(np.float64), case["border0"].numpy())


def test_rodrigues():
    from curvecloudnet_b200.simulator import rodrigues
    assert np.allclose(rodrigues([0, 0, 0]), np.eye(3))
    a = 0.7
    assert np.allclose(rodrigues([0, 0, a]), [[math.cos(a), -math.sin(a), 0], [math.sin(a), math.cos(a), 0], [0, 0, 1]])
    R = rodrigues([0.3, -1.1, 0.4])
    assert np.allclose(R @ R.T, np.eye(3)) and abs(np.linalg.det(R) - 1) < 1e-12
    v = np.array([0.3, -1.1, 0.4])
    assert np.allclose(R @ v, v)                       # the axis is fixed


def test_raster_oracle_on_a_sphere_and_a_plane():
    from curvecloudnet_b200.synthetic import synthetic_mesh
    from oracle import ccn_simulator as osim
    res = 128
    R, t = torch.eye(3), torch.tensor([0.0, 0.0, 1.0])
    v, f = synthetic_mesh("sphere", 0, 64)
    coords, normals, mask, depth = osim.rasterize(v, f, R, t, res, return_depth=True)
    # silhouette of a sphere of radius 0.28 at distance 1: sin(alpha) = 0.28 -> tan(alpha) = 0.2917 -> NDC radius 2 tan = 0.583
    frac = mask.float().mean().item()
    assert abs(frac - math.pi * (0.2917 * res) ** 2 / res ** 2) < 0.01
    assert abs(depth[res // 2, res // 2].item() - 0.72) < 2e-3 and depth[0, 0].item() == 1e10
    r = coords[mask].norm(dim=-1)
    assert float((r - 0.28).abs().max()) < 2e-3        # scene-frame points lie on the (faceted) sphere
    outward = (torch.nn.functional.normalize(coords[mask], dim=-1) * normals[mask]).sum(-1)
    assert float(outward.abs().min()) > 0.95           # face normals are radial (the sign follows the kept winding)
    assert torch.equal(coords[~mask], (torch.zeros(3) - t).expand_as(coords[~mask]))
    # one big quad at z = 0.5 in the scene (camera depth 1.5): constant depth, both windings present -> exactly one is kept
    v = torch.tensor([[-2.0, -2.0, 0.5], [2.0, -2.0, 0.5], [2.0, 2.0, 0.5], [-2.0, 2.0, 0.5]])
    f = torch.tensor([[0, 1, 2], [0, 2, 3], [0, 2, 1], [0, 3, 2]], dtype=torch.int32)
    coords, normals, mask, depth = osim.rasterize(v, f, R, t, 64, return_depth=True)
    assert bool(mask.all()) and float((depth - 1.5).abs().max()) < 1e-6
    assert float((normals[..., 2].abs() - 1).abs().max()) < 1e-6 and len(torch.unique(normals[..., 2].sign())) == 1
    # pixel (i, j) looks along ((j + .5) / res - .5, (i + .5) / res - .5, 1): x of the hit = 1.5 * that
    j = torch.arange(64).float()
    assert float((coords[10, :, 0] - 1.5 * ((j + 0.5) / 64 - 0.5)).abs().max()) < 1e-5
