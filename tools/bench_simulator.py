"""Dev tool: device time of the scanning simulator stages (CUDA events, 20 repetitions, medians) on synthetic meshes at the
reference's default settings (1024 x 1024 raster, 2048 points), with the reference's own rasteriser kernels
(oracle/_ref/ref_mpr_cuda.so) timed beside ours.   usage: python tools/bench_simulator.py"""
import importlib.util
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from curvecloudnet_b200 import simulator as sim
from curvecloudnet_b200.synthetic import synthetic_mesh

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def timed(fn, reps=20):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


ref = None
path = os.path.join(ROOT, "oracle", "_ref", "ref_mpr_cuda.so")
if os.path.exists(path):
    spec = importlib.util.spec_from_file_location("ref_mpr_cuda", path)
    ref = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref)

res = 1024
R = torch.tensor(sim.rodrigues([0.6, 0, 0]) @ sim.rodrigues([0, 2.0, 0]), dtype=torch.float32).cuda()
t = torch.tensor([0.0, 0.0, 1.0]).cuda()
print("rasterise %d x %d (ours: project + z-buffer + resolve to the scene frame; reference: its two kernels only)" % (res, res))
for kind, detail in (("blob", 64), ("blob", 256), ("torus", 128), ("table", 0)):
    v, f = synthetic_mesh(kind, 1, detail)
    v, f = v.cuda(), f.cuda()
    ours = timed(lambda: sim.rasterize(v, f, R, t, res))
    line = "  %-6s %7d faces   ours %8.1f us" % (kind, f.shape[0], ours)
    if ref is not None:
        cam = v @ R.T + t.view(1, 3)
        K = torch.tensor([[res, 0, res // 2], [0, res, res // 2], [0, 0, 1]], device="cuda", dtype=torch.float32)
        ndc = cam @ K.T
        ndc[:, :2] /= ndc[:, 2:] + 1e-9
        ndc[:, :2] = 2 * ndc[:, :2] / res - 1
        filt = torch.ones(v.shape[0], dtype=torch.uint8, device="cuda")
        ndc, fi, cam = ndc.contiguous(), f.int().contiguous(), cam.contiguous()
        theirs = timed(lambda: ref.estimate_normals(ndc, fi, cam, filt, res, res))
        line += "   reference kernels %8.1f us   x%.1f" % (theirs, theirs / ours)
    print(line)

meshes = [synthetic_mesh(k, s, 96) for s, k in enumerate(["blob", "torus", "table", "sphere"] * 4)]
verts, faces = [m[0].cuda() for m in meshes], [m[1].cuda() for m in meshes]
sampler = sim.SampleMeshes(2048, "cuda", resolution=res, camera_view=(1.0, 0.6, 2.0, 0.0))
coords, normals, masks = sampler._render_meshes(verts, faces)
B = len(meshes)


def lines():
    np.random.seed(0)
    torch.manual_seed(0)
    return sampler.scan_lines_gen.generate_scan_lines(B, 2048, masks, coords, normals)


ok, uv, pts, nrm = lines()
import time
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(5):
    lines()
torch.cuda.synchronize()
print("scan lines, B = %d meshes x 2048 points: %.2f ms wall per call (host draws of 100 x B lines included)" % (B, (time.perf_counter() - t0) / 5 * 1e3))
print("curve clouds (split + 2 x anti-alias + curvature), B = %d x 2048: %.1f us" %
      (B, timed(lambda: sim.CurveClouds(pts, nrm, uv[:, :, :2], uv[:, :, 2]))))
t0 = time.perf_counter()
for _ in range(3):
    sampler.meshes2points(verts, faces)
torch.cuda.synchronize()
dt = (time.perf_counter() - t0) / 3
print("SampleMeshes.meshes2points, %d meshes: %.1f ms wall = %.0f curve clouds / s" % (B, dt * 1e3, B / dt))
