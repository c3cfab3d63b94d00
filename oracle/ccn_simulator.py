"""TEST INFRASTRUCTURE - CPU oracle of the scanning simulator.  Only tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline leg may import this; the product path (curvecloudnet_b200/simulator.py) never does.

  rasterize()   -> oracle/sim_raster.c (plain C restatement of minimal_rasterizer.py:14-51 + rasterizer_kernel.cu:80-465);
                   pinned on the GPU box against the reference's own CUDA kernels compiled into oracle/_ref/ref_mpr_cuda.so
                   (tests/test_gpu_simulator.py); the reference has no CPU rasteriser.
"""
from __future__ import annotations

import ctypes
import os

import numpy as np
import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_lib = None


def _load():
    global _lib
    if _lib is None:
        path = os.path.join(_HERE, "_build", "liboracle_sim.so")
        if not os.path.exists(path):
            raise RuntimeError("oracle/_build/liboracle_sim.so is missing: run `make -C oracle` (__graft_entry__.build() does)")
        _lib = ctypes.CDLL(path)
        _lib.oracle_sim_rasterize.restype = ctypes.c_int
        _lib.oracle_sim_rasterize.argtypes = [ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p,
                                              ctypes.c_void_p, ctypes.c_int32, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                              ctypes.c_void_p]
    return _lib


def rasterize(vertices, faces, R, t, res=1024, return_depth=False):
    """CPU tensors in, CPU tensors out: coords (res, res, 3), normals (res, res, 3), mask (res, res) bool [, depth]."""
    v = vertices[:, :3].detach().cpu().to(torch.float32).contiguous()
    f = faces.detach().cpu().to(torch.int32).contiguous()
    Rm = R.detach().cpu().to(torch.float32).contiguous()
    tv = t.detach().cpu().to(torch.float32).reshape(3).contiguous()
    coords = torch.empty((res, res, 3), dtype=torch.float32)
    normals = torch.empty((res, res, 3), dtype=torch.float32)
    mask = torch.empty((res, res), dtype=torch.uint8)
    depth = torch.empty((res, res), dtype=torch.float32)
    rc = _load().oracle_sim_rasterize(v.data_ptr(), v.shape[0], f.data_ptr(), f.shape[0], Rm.data_ptr(), tv.data_ptr(), res,
                                      coords.data_ptr(), normals.data_ptr(), mask.data_ptr(), depth.data_ptr())
    if rc != 0:
        raise RuntimeError("oracle_sim_rasterize failed (%d)" % rc)
    if return_depth:
        return coords, normals, mask.bool(), depth
    return coords, normals, mask.bool()
