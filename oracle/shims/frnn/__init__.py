"""shim: frnn (vendored third_party/FRNN needs CUDA <= 11; result contract only)"""
from oracle import tp_ops as _t


def frnn_grid_points(points1, points2, lengths1=None, lengths2=None, K=-1, r=-1, grid=None,
                     return_nn=False, return_sorted=True, radius_cell_ratio=2.0):
    dists, idxs = _t.frnn_grid_points(points1, points2, lengths1, lengths2, K, r)
    nn = _t.frnn_gather(points2, idxs, lengths2) if return_nn else None
    return dists, idxs, nn, None


frnn_gather = _t.frnn_gather
