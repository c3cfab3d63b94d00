"""Drop-in replacements of the two public entry points of the reference's vendored FRNN package
(third_party/FRNN/frnn/frnn.py): `frnn_grid_points` (:210-347, called at src/models/utils/point_ops.py:459) and
`frnn_gather` (:351-397, called at src/models/modules/dgcnn.py:172), same signatures and result conventions, on the
sm_100a kernels of libccn_b200 (hashed grid + warp top-K, csrc/search_ops.cu; row gather, csrc/agg_ops.cu).

Result contract kept: up to K points of points2 with d2 <= r*r (find_nbrs.cu:490), ascending by distance, idxs (N, P1, K)
int64 LOCAL to each cloud and -1 padded, dists -1 padded; ties in distance by lowest index (the reference's in-cell order
is atomic dependent, insert_points.cu:109).  D = 3 only (the path never uses the 2-D variant); no backward (the path
consumes indices only).  `grid`: the reference lets a caller pass the returned grid back in to skip the rebuild for the
same points2 / r; the returned `_GRID` here carries our hashed grid and is reused the same way.
"""
from __future__ import annotations

from collections import namedtuple
from typing import Optional, Union

import torch

from . import native, ops

_GRID = namedtuple("_GRID", "sorted_points2 pc2_grid_off sorted_points2_idxs grid_params")


def _flatten(points, lengths):
    """(N, P, 3) padded + lengths -> ragged (n, 3), batch ids (n), host lengths."""
    N, P, _ = points.shape
    if lengths is None:
        lens = [P] * N
    else:
        lens = [int(v) for v in lengths.tolist()]
    if all(l == P for l in lens):
        flat = points.reshape(N * P, points.shape[2])
    else:
        flat = torch.cat([points[b, :lens[b]] for b in range(N)])
    batch = torch.repeat_interleave(torch.arange(N, device=points.device), torch.tensor(lens, device=points.device))
    return ops._f32(flat.contiguous()), batch, lens


def frnn_grid_points(points1: torch.Tensor, points2: torch.Tensor, lengths1: Optional[torch.Tensor] = None,
                     lengths2: Optional[torch.Tensor] = None, K: int = -1, r: Union[float, torch.Tensor] = -1,
                     grid: Optional[_GRID] = None, return_nn: bool = False, return_sorted: bool = True,
                     radius_cell_ratio: float = 2.0):
    """-> (dists (N, P1, K) f32, idxs (N, P1, K) int64, nn (N, P1, K, 3) or None, grid)."""
    if not (points1.is_cuda and points2.is_cuda):
        raise TypeError("frnn_grid_points only supports CUDA tensors")
    if points1.shape[0] != points2.shape[0]:
        raise ValueError("points1 and points2 must have the same batch dimension")
    if points1.shape[2] != points2.shape[2]:
        raise ValueError("dimension mismatch: points1 of dimension %d while points2 of dimension %d"
                         % (points1.shape[2], points2.shape[2]))                      # frnn.py:286-289
    if K <= 0 or K > 32:
        raise ValueError("K must be in [1, 32]")                      # the reference kernels are instantiated for K <= 32
    N, P1, _ = points1.shape
    if torch.is_tensor(r):
        rv = [float(v) for v in r.reshape(-1).tolist()]
        if len(rv) not in (1, N):
            raise ValueError("r: one radius, or one per cloud")                              # frnn.py:319
        if any(v != rv[0] for v in rv):
            # a radius per cloud (frnn.py:316-323; no reference call site uses it): one search per cloud at the boundary
            l1 = [P1] * N if lengths1 is None else [int(v) for v in lengths1.tolist()]
            l2 = [points2.shape[1]] * N if lengths2 is None else [int(v) for v in lengths2.tolist()]
            dists = torch.full((N, P1, K), -1.0, dtype=torch.float32, device=points1.device)
            idxs = torch.full((N, P1, K), -1, dtype=torch.int64, device=points1.device)
            for b in range(N):
                if l1[b] == 0 or l2[b] == 0:
                    continue
                d, i, _, _ = frnn_grid_points(points1[b:b + 1, :l1[b]].contiguous(), points2[b:b + 1, :l2[b]].contiguous(), None, None,
                                              K, rv[b], None, False, return_sorted, radius_cell_ratio)
                dists[b, :l1[b]], idxs[b, :l1[b]] = d[0], i[0]
            return dists, idxs, (frnn_gather(points2, idxs, lengths2) if return_nn else None), None
        r = rv[0]
    p1, b1, l1 = _flatten(points1, lengths1)
    gp = grid.grid_params if grid is not None and isinstance(grid.grid_params, dict) else None
    if gp is not None and gp["key"] == (tuple(points2.shape), float(r)):
        # the caller hands the grid of the same points2 / r back (frnn.py:329-332): no rebuild
        p2, b2, l2, s2, cache = gp["p2"], gp["b2"], gp["l2"], gp["s2"], gp["cache"]
    else:
        from .nn import ForwardCache
        p2, b2, l2 = _flatten(points2, lengths2)
        s2 = ops.sample_ptr(b2) if p2.shape[0] else None
        cache = ForwardCache()
    dists = torch.full((N, P1, K), -1.0, dtype=torch.float32, device=points1.device)
    idxs = torch.full((N, P1, K), -1, dtype=torch.int64, device=points1.device)
    if p1.shape[0] and p2.shape[0]:
        if points1.shape[2] == 3:
            idx, d2, _ = ops.frnn_knn(p1, b1, p2, b2, s2, K, float(r), want_d2=True, cache=cache)
        else:
            # D != 3 (upstream's N-D kernels, grid.h:27-34; reached only by DynamicEdgeConv): brute force in feature space,
            # same result contract (d2 <= r^2, ascending by (d2, index)); no grid object comes back
            idx, _, d2 = ops.feat_neighbours(p1, b1, p2, s2, K, float(r), "knn", want_d2=True)
        off1 = off2 = 0
        for b in range(N):
            blk = idx[off1:off1 + l1[b]].to(torch.int64)
            dd = d2[off1:off1 + l1[b]]
            ok = blk >= 0
            idxs[b, :l1[b]] = torch.where(ok, blk - off2, blk)
            dists[b, :l1[b]] = torch.where(ok, dd, torch.full_like(dd, -1.0))
            off1 += l1[b]
            off2 += l2[b]
    g = None
    if cache.grids:
        cells, cell_off, sorted_, _ = next(iter(cache.grids.values()))
        g = _GRID(sorted_points2=sorted_, pc2_grid_off=cell_off, sorted_points2_idxs=None,
                  grid_params={"key": (tuple(points2.shape), float(r)), "cache": cache, "cells": cells, "p2": p2, "b2": b2,
                               "l2": l2, "s2": s2})
    nn = frnn_gather(points2, idxs, lengths2) if return_nn else None
    return dists, idxs, nn, g


def frnn_gather(x: torch.Tensor, idxs: torch.Tensor, lengths: Optional[torch.Tensor] = None):
    """x (N, M, U), idxs (N, L, K) from frnn_grid_points -> (N, L, K, U): x_out[n, l, k] = x[n, idxs[n, l, k]], zeros where
    idxs < 0 (frnn.py:351-397)."""
    if not x.is_cuda:
        raise TypeError("frnn_gather only supports CUDA tensors")
    N, M, U = x.shape
    _N, L, K = idxs.shape
    if N != _N:
        raise ValueError("x and idxs must have same batch dimension")
    flat = ops._f32(x.reshape(N * M, U).contiguous())
    off = (torch.arange(N, device=x.device, dtype=torch.int64) * M).view(N, 1, 1)
    g = torch.where(idxs >= 0, idxs + off, idxs).reshape(-1).to(torch.int32).contiguous()     # index plumbing of the boundary
    out = ops.gather_rows_i32(flat, g)                                                          # -1 rows -> zeros
    return out.view(N, L, K, U)
