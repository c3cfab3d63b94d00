"""Generate tests/golden/simulator.pt by running the REAL reference python of the scanning simulator on CPU.

TEST INFRASTRUCTURE, build container only (needs /root/reference; the committed fixture travels to the GPU box instead).
Usage:  python -m oracle.make_golden_sim

What runs from the reference tree, unmodified: scanning_simulator/utils/scanlines.py (ScanLineGenerator), curve_cloud.py
(CurveClouds: curve split, anti-alias, compute_curvature_new) and curvature_approx.py.  Their third-party imports resolve to
oracle/shims (torch_scatter is imported but unused without intersections; torch_geometric.nn.knn = torch_cluster.knn restated
by brute force).  Two documented deviations: `np.int` (removed in numpy >= 1.24, used by the 'parallel' / 'grid' directions,
scanlines.py:106,122) is set to `int`; CurveClouds is built per mesh with with_intersections=True (sampling.py:85).
The rasterised views the lines run over come from oracle/sim_raster.c (the reference rasteriser is CUDA-only; it is compared
with ours directly on the GPU box, tests/test_gpu_simulator.py).
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
OUT = os.path.join(ROOT, "tests", "golden", "simulator.pt")


def main():
    for p in (ROOT, os.path.join(ROOT, "oracle", "shims"), REF):
        if p not in sys.path:
            sys.path.insert(0, p)
    if not hasattr(np, "int"):
        np.int = int
    from scanning_simulator.utils.curve_cloud import CurveClouds
    from scanning_simulator.utils.scanlines import ScanLineGenerator
    from curvecloudnet_b200.simulator import rodrigues
    from curvecloudnet_b200.synthetic import synthetic_mesh
    from oracle import ccn_simulator as osim

    def views(kinds, res, ry):
        R = torch.tensor(rodrigues([0.6, 0, 0]) @ rodrigues([0, ry, 0]) @ rodrigues([0, 0, 0.2]), dtype=torch.float32)
        t = torch.tensor([0.0, 0.0, 1.0])
        out = [osim.rasterize(*synthetic_mesh(k, s, 32), R, t, res) for s, k in enumerate(kinds)]
        return torch.stack([o[0] for o in out]), torch.stack([o[1] for o in out]), torch.stack([o[2] for o in out])

    cases = []
    specs = [("random", ["blob", "torus", "table"], 512, 400, 1.0, 11, True),      # several meshes, one attempt
             ("random", ["sphere", "torus"], 96, 2500, 0.25, 12, False),                      # many attempts: N close to the visible pixels
             ("random", ["table", "blob"], 384, 64, 0.5, 13, True),                 # the cut inside the first attempt, half density
             ("parallel", ["blob", "torus"], 400, 500, 1.0, 14, True),
             ("grid", ["torus", "blob"], 512, 700, 1.0, 15, True),
             ("random", ["blob", "torus"], 640, 1500, 1.0, 16, True)]    # more points than one 1024-thread round of the scans
    for direction, kinds, res, N, density, seed, clouds in specs:
        coords, normals, masks = views(kinds, res, 0.4 * seed)
        B = len(kinds)
        np.random.seed(seed)
        torch.manual_seed(seed)
        gen = ScanLineGenerator(res, "cpu", "linear", direction, density)
        ok, uv = gen.generate_scan_lines(B, N, masks)
        assert ok, (direction, kinds)
        # the first attempt's line end points again, for the CPU test of the host-side draws
        np.random.seed(seed)
        g2 = ScanLineGenerator(res, "cpu", "linear", direction, density)
        border0 = (g2._sample_linear_border_points(g2.NLINES, B) if direction == "random" else
                   g2._sample_parallel_border_points(0, g2.NLINES, B) if direction == "parallel" else
                   g2._sample_grid_border_points(0, g2.NLINES, B))
        border1 = None if direction == "random" else (g2._sample_parallel_border_points(1, g2.NLINES, B) if direction == "parallel"
                                                       else g2._sample_grid_border_points(1, g2.NLINES, B))
        case = {"border0": border0.cpu().double(), "border1": border1.cpu().double() if border1 is not None else None,
                "direction": direction, "kinds": kinds, "res": res, "N": N, "density": density, "seed": seed,
                "uv": uv.to(torch.int16), "lines": int(uv[:, :, 2].max()) + 1}
        if clouds:       # (the reference's curvature estimate throws when no curve of a mesh has curvature_knn points)
            bi = torch.arange(B).repeat_interleave(N)
            pts = coords[bi, uv[:, :, 0].flatten(), uv[:, :, 1].flatten()].view(B, N, 3)
            nrm = normals[bi, uv[:, :, 0].flatten(), uv[:, :, 1].flatten()].view(B, N, 3)
            # one mesh per CurveClouds, as every caller in the reference does (generate_shapenet.py:69,83: a list of ONE model):
            # with several meshes its batchdense2curvepadded (curve_cloud.py:200-214) offsets the curve ids by the INCLUSIVE
            # running curve count, so the ids of a mesh with fewer curves than its predecessor collide with the predecessor's
            # with_intersections=True as sampling.py:85 passes it (the result is only ever visualised): it works for one mesh
            # only - the second scatter_min of curve_cloud.py:146 cannot broadcast its index for B > 1
            ccs = [CurveClouds(pts[b:b + 1], nrm[b:b + 1], uv[b:b + 1, :, :2], uv[b:b + 1, :, 2], anti_alias=True,
                               with_intersections=True, curvature_knn=10) for b in range(B)]
            cat = lambda name: torch.cat([getattr(c, name) for c in ccs], dim=0)
            case.update({"curve_idx": cat("point2curve_idxs").to(torch.int16), "points": cat("points"), "normals": cat("normals"),
                         "curvature": cat("curvature"), "gradients": cat("gradients") if N <= 500 else None,
                         "intersections": [c.intersections.to(torch.int16) for c in ccs],
                         "intersection_batch": [c.intersection2batch_idxs.to(torch.int16) for c in ccs]})
        cases.append(case)
        print(direction, kinds, "res", res, "N", N, "lines", case["lines"],
              "curves %d" % (int(case["curve_idx"].max()) + 1) if clouds else "")
    torch.save({"cases": cases}, OUT)
    print("wrote", OUT, os.path.getsize(OUT) // 1024, "KiB")


if __name__ == "__main__":
    main()
