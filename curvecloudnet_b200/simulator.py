"""Scanning simulator on the GPU: triangle meshes -> curve clouds, the generator of the reference's
simulated ShapeNet / Kortx datasets.  Host-side mirror of the reference interface - same names, argument meaning and random
draws - over the C-ABI kernels of csrc/sim_raster.cu and csrc/sim_scan.cu (no CPU path: tensors must be CUDA tensors):

    rasterize(vertices, faces, R, t, res)             scanning_simulator/utils/minimal_rasterizer.py:14-51
    ScanLineGenerator(resolution, device, ...)        scanning_simulator/utils/scanlines.py:4-150
    CurveClouds(points, normals, uv, curve_idxs, ...) scanning_simulator/utils/curve_cloud.py:8-190
    SampleMeshes(npoints, device, ...)                scanning_simulator/utils/sampling.py:13-85

No graphics API is involved: the rasteriser is a z-buffer of 64-bit atomicMin keys in plain CUDA.
"""
from __future__ import annotations

import math

import numpy as np
import torch

from . import native
from .native import call


def rodrigues(rvec) -> np.ndarray:
    """cv2.Rodrigues(rvec)[0] for a rotation vector (sampling.py:34-36 builds the camera rotation from three of them):
    R = cos(th) I + (1 - cos(th)) r r^T + sin(th) [r]_x, th = |rvec|, r = rvec / th  (float64)."""
    rvec = np.asarray(rvec, dtype=np.float64).reshape(3)
    th = float(np.linalg.norm(rvec))
    if th < np.finfo(np.float64).eps:
        return np.eye(3)
    r = rvec / th
    c, s = math.cos(th), math.sin(th)
    K = np.array([[0.0, -r[2], r[1]], [r[2], 0.0, -r[0]], [-r[1], r[0], 0.0]])
    return c * np.eye(3) + (1.0 - c) * np.outer(r, r) + s * K


def rasterize(vertices: torch.Tensor, faces: torch.Tensor, R: torch.Tensor, t: torch.Tensor, res: int = 1024,
              return_depth: bool = False):
    """minimal_rasterizer.rasterize: -> coords (res, res, 3), normals (res, res, 3) in the scene frame, mask (res, res) bool.
    vertices (nv, >=3) float, faces (nf, 3) int, R (3, 3), t (3,) or (1, 3): CUDA tensors."""
    if not vertices.is_cuda:
        raise native.CcnError("simulator.rasterize: CUDA tensors only (no CPU path exists)")
    dev = vertices.device
    v = vertices[:, :3].to(torch.float32).contiguous()
    f = faces.to(device=dev, dtype=torch.int32).contiguous()
    Rm = R.to(device=dev, dtype=torch.float32).contiguous()
    tv = t.to(device=dev, dtype=torch.float32).reshape(3).contiguous()
    nv, nf = v.shape[0], f.shape[0]
    cam = torch.empty((max(nv, 1), 3), dtype=torch.float32, device=dev)
    ndc = torch.empty((max(nv, 1), 3), dtype=torch.float32, device=dev)
    keys = torch.empty((res * res,), dtype=torch.int64, device=dev)
    lst = torch.empty((nf + 1,), dtype=torch.int32, device=dev)
    coords = torch.empty((res, res, 3), dtype=torch.float32, device=dev)
    normals = torch.empty((res, res, 3), dtype=torch.float32, device=dev)
    mask = torch.empty((res, res), dtype=torch.uint8, device=dev)
    depth = torch.empty((res, res), dtype=torch.float32, device=dev) if return_depth else None
    call("ccn_sim_rasterize", v, nv, f, nf, Rm, tv, res, cam, ndc, keys, lst, coords, normals, mask,
         depth if depth is not None else 0)
    if return_depth:
        return coords, normals, mask.view(torch.bool), depth
    return coords, normals, mask.view(torch.bool)
