"""Restatement (CPU, plain torch) of the third-party arithmetic on the hot path.

TEST INFRASTRUCTURE - see oracle/__init__.py.  None of these packages is vendored
in /root/reference, so each function states the published behaviour of the pinned
upstream version and cites the reference CALL SITE it serves.

Determinism rules (where upstream is silent or device dependent):
  * squared distances are ((dx*dx)+(dy*dy))+(dz*dz) in fp32, every operation
    rounded separately (no FMA) - the CUDA kernels use __fmul_rn/__fadd_rn;
  * every selection is by the total order (d2, index): ties -> lowest index;
  * segment sums run in index order.
"""
from __future__ import annotations

import torch

_I64_MAX = torch.iinfo(torch.int64).max

# ----------------------------------------------------------------------------------
# C restatement of the neighbour queries (oracle/nbr_search.c, built by oracle/Makefile into
# oracle/_build/liboracle_nbr.so): same result contract as the torch statements below, a cell grid
# prunes the candidates so that full-size scans finish in seconds.  SEARCH_IMPL = "torch" forces the
# O(P1*P2) torch statements (tests/test_oracle_numerics.py compares the two).
# ----------------------------------------------------------------------------------
SEARCH_IMPL = "auto"        # "auto" (C when built) | "c" | "torch"
_nbr_lib = None


def _nbr():
    global _nbr_lib
    if SEARCH_IMPL == "torch":
        return None
    if _nbr_lib is None:
        import ctypes
        import os
        path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_build", "liboracle_nbr.so")
        if not os.path.exists(path):
            if SEARCH_IMPL == "c":
                raise RuntimeError("%s not built (make -C oracle)" % path)
            _nbr_lib = False
        else:
            lib = ctypes.CDLL(path)
            P, Q, I, F = ctypes.c_void_p, ctypes.c_int64, ctypes.c_int, ctypes.c_float
            lib.oracle_frnn.argtypes = [P, Q, P, Q, I, F, P, P, I]
            lib.oracle_ball.argtypes = [P, Q, P, Q, I, F, P, I]
            lib.oracle_knn.argtypes = [P, Q, P, Q, I, P, P, I]
            _nbr_lib = lib
    return _nbr_lib or None


def _c_search(kind, q, pts, K, r=0.0):
    """One cloud through nbr_search.c -> (idx int64 (n1,K) -1 padded, d2 f32 (n1,K) -1 padded | None)."""
    lib = _nbr()
    q = q.contiguous().float()
    pts = pts.contiguous().float()
    n1, n2 = q.shape[0], pts.shape[0]
    idx = torch.empty((n1, K), dtype=torch.int64)
    d2 = torch.empty((n1, K), dtype=torch.float32) if kind != "ball" else None
    th = max(1, torch.get_num_threads())
    if kind == "frnn":
        rc = lib.oracle_frnn(q.data_ptr(), n1, pts.data_ptr(), n2, K, float(r), idx.data_ptr(), d2.data_ptr(), th)
    elif kind == "ball":
        rc = lib.oracle_ball(q.data_ptr(), n1, pts.data_ptr(), n2, K, float(r), idx.data_ptr(), th)
    else:
        rc = lib.oracle_knn(q.data_ptr(), n1, pts.data_ptr(), n2, K, idx.data_ptr(), d2.data_ptr(), th)
    if rc:
        raise RuntimeError("oracle nbr_search: error %d" % rc)
    return idx, d2


# ----------------------------------------------------------------------------------
# helpers
# ----------------------------------------------------------------------------------
def sqdist3(a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    """((dx*dx)+(dy*dy))+(dz*dz), broadcasting over leading dims, each op rounded."""
    dx = a[..., 0] - b[..., 0]
    dy = a[..., 1] - b[..., 1]
    dz = a[..., 2] - b[..., 2]
    return (dx * dx + dy * dy) + dz * dz


def _sqdist(a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    """Generic-D version of sqdist3 (sequential over d, no FMA)."""
    D = a.shape[-1]
    if D == 3:
        return sqdist3(a, b)
    acc = None
    for d in range(D):
        t = a[..., d] - b[..., d]
        t = t * t
        acc = t if acc is None else acc + t
    return acc


def _packed_keys(d2: torch.Tensor, ncols: int) -> torch.Tensor:
    """int64 key = (bits(d2) << 32) | column, monotone in (d2, col) for d2 >= 0."""
    bits = d2.contiguous().view(torch.int32).to(torch.int64)
    cols = torch.arange(ncols, dtype=torch.int64, device=d2.device)
    return (bits << 32) | cols


def _smallest_k(d2: torch.Tensor, valid: torch.Tensor, K: int):
    """Row-wise K smallest by (d2, col); returns (cols, d2, count), col=-1 padded."""
    P1, P2 = d2.shape
    keys = _packed_keys(d2, P2)
    keys = torch.where(valid, keys, torch.full_like(keys, _I64_MAX))
    k = min(K, P2)
    top = torch.topk(keys, k, dim=1, largest=False, sorted=True).values if k > 0 else keys[:, :0]
    ok = top != _I64_MAX
    cols = torch.where(ok, top & 0xFFFFFFFF, torch.full_like(top, -1))
    dist = torch.where(ok, (top >> 32).to(torch.int32).view(torch.float32), torch.zeros_like(top, dtype=torch.float32))
    if k < K:
        cols = torch.cat([cols, cols.new_full((P1, K - k), -1)], dim=1)
        dist = torch.cat([dist, dist.new_zeros((P1, K - k))], dim=1)
    return cols, dist, ok.sum(dim=1)


# ----------------------------------------------------------------------------------
# pytorch3d.ops (@stable, unpinned - setup.sh:23)
# ----------------------------------------------------------------------------------
def sample_farthest_points(points, lengths, K, start_idxs):
    """pytorch3d.ops.sample_farthest_points, called at src/models/utils/point_ops.py:64.

    points (N,P,3) f32, lengths (N,) int, K (N,) number of samples per cloud,
    start_idxs (N,) the first selected point (the reference draws it with
    torch.randint, here it is an explicit input).  Iteration: closest = min(closest,
    |p - p_last|^2) (init +inf), next = argmax(closest), ties -> lowest index.
    Returns idx (N, max K) int64, -1 padded, in selection order.
    """
    N = points.shape[0]
    K = [int(k) for k in (K.tolist() if torch.is_tensor(K) else K)]
    lengths = [int(l) for l in (lengths.tolist() if torch.is_tensor(lengths) else lengths)]
    start_idxs = [int(s) for s in (start_idxs.tolist() if torch.is_tensor(start_idxs) else start_idxs)]
    maxk = max(K) if N else 0
    out = torch.full((N, maxk), -1, dtype=torch.int64)
    for n in range(N):
        L = lengths[n]
        kn = min(K[n], L)
        if kn <= 0:
            continue
        p = points[n, :L]
        closest = torch.full((L,), float("inf"), dtype=points.dtype)
        sel = start_idxs[n]
        out[n, 0] = sel
        for i in range(1, kn):
            closest = torch.minimum(closest, sqdist3(p, p[sel]))
            sel = int(torch.argmax(closest))  # first maximal value
            out[n, i] = sel
    return out


def ball_query(p1, p2, lengths1, lengths2, K, radius):
    """pytorch3d.ops.ball_query, called at point_ops.py:81 with K=128.

    For every p1: the first K points of p2 IN INDEX ORDER with d2 < radius^2
    (strict); -1 padded; NOT distance sorted.  Returns idx (N,P1,K) int64.
    """
    N, P1, _ = p1.shape
    r2 = torch.tensor(radius, dtype=torch.float32) * torch.tensor(radius, dtype=torch.float32)
    out = torch.full((N, P1, K), -1, dtype=torch.int64)
    for n in range(N):
        l1, l2 = int(lengths1[n]), int(lengths2[n])
        if l1 == 0 or l2 == 0:
            continue
        if p1.shape[-1] == 3 and _nbr() is not None:
            out[n, :l1] = _c_search("ball", p1[n, :l1], p2[n, :l2], K, float(torch.tensor(radius, dtype=torch.float32)))[0]
            continue
        for s in range(0, l1, 2048):
            q = p1[n, s:min(s + 2048, l1)]
            inside = _sqdist(q[:, None, :], p2[n, None, :l2, :]) < r2     # any D (sqdist3 for D = 3)
            rank = torch.cumsum(inside, dim=1)  # 1-based ordinal among hits
            keep = inside & (rank <= K)
            rows, cols = torch.nonzero(keep, as_tuple=True)
            out[n, s + rows, rank[rows, cols] - 1] = cols
    return out


def knn_points(p1, p2, lengths1, lengths2, K):
    """pytorch3d.ops.knn_points, called at point_ops.py:91 (-> :331, dgcnn.py:222).

    Exact K nearest by (d2, index), ascending.  Where lengths2 < K the tail (and all
    padded p1 rows) is 0-filled: idx 0, dist 0 (NOT -1; upstream docs).  Returns
    (dists (N,P1,K) f32, idx (N,P1,K) int64).
    """
    N, P1, _ = p1.shape
    idx = torch.zeros((N, P1, K), dtype=torch.int64)
    dists = torch.zeros((N, P1, K), dtype=torch.float32)
    for n in range(N):
        l1, l2 = int(lengths1[n]), int(lengths2[n])
        if l1 == 0 or l2 == 0:
            continue
        if p1.shape[-1] == 3 and _nbr() is not None:
            cols, dd = _c_search("knn", p1[n, :l1], p2[n, :l2], K)
            idx[n, :l1] = torch.where(cols >= 0, cols, torch.zeros_like(cols))
            dists[n, :l1] = torch.where(cols >= 0, dd, torch.zeros_like(dd))
            continue
        for s in range(0, l1, 2048):
            e = min(s + 2048, l1)
            d2 = _sqdist(p1[n, s:e, None, :], p2[n, None, :l2, :])
            cols, dd, _ = _smallest_k(d2, torch.ones_like(d2, dtype=torch.bool), K)
            idx[n, s:e] = torch.where(cols >= 0, cols, torch.zeros_like(cols))
            dists[n, s:e] = dd
    return dists, idx


# ----------------------------------------------------------------------------------
# FRNN (vendored: third_party/FRNN) - result contract of frnn_grid_points
# ----------------------------------------------------------------------------------
def frnn_grid_points(points1, points2, lengths1, lengths2, K, r):
    """Result contract of frnn.frnn_grid_points (third_party/FRNN/frnn/frnn.py:210-347,
    kernel third_party/FRNN/frnn/csrc/grid/find_nbrs.cu:418-504), called at
    point_ops.py:459: up to K points of points2 with d2 <= r*r (non-strict, find_nbrs.cu:490),
    ascending by d2, idx -1 padded, dists -1 padded.  The grid is an implementation
    detail; in-cell order in the reference comes from atomicAdd (insert_points.cu:109),
    so ties are ordered by lowest index here.  r: float or (N,) tensor.
    """
    N, P1, _ = points1.shape
    idx = torch.full((N, P1, K), -1, dtype=torch.int64)
    dists = torch.full((N, P1, K), -1.0, dtype=torch.float32)
    if not torch.is_tensor(r):
        r = torch.full((N,), float(r), dtype=torch.float32)
    r = r.to(torch.float32).reshape(-1)
    if r.numel() == 1:
        r = r.expand(N)
    for n in range(N):
        l1, l2 = int(lengths1[n]), int(lengths2[n])
        if l1 == 0 or l2 == 0:
            continue
        r2 = r[n] * r[n]
        if points1.shape[-1] == 3 and _nbr() is not None:
            idx[n, :l1], dists[n, :l1] = _c_search("frnn", points1[n, :l1], points2[n, :l2], K, float(r[n]))
            continue
        for s in range(0, l1, 2048):
            e = min(s + 2048, l1)
            d2 = _sqdist(points1[n, s:e, None, :], points2[n, None, :l2, :])
            cols, dd, cnt = _smallest_k(d2, d2 <= r2, K)
            idx[n, s:e] = cols
            dists[n, s:e] = torch.where(cols >= 0, dd, torch.full_like(dd, -1.0))
    return dists, idx


def frnn_gather(x, idxs, lengths=None):
    """frnn.frnn_gather (third_party/FRNN/frnn/frnn.py:351-397): out[n,l,k] = x[n, idxs[n,l,k]],
    zeros where idxs < 0."""
    N, P2, U = x.shape
    safe = idxs.clamp(min=0)
    out = torch.gather(x[:, :, None, :].expand(-1, -1, idxs.shape[2], -1), 1,
                       safe[:, :, :, None].expand(-1, -1, -1, U))
    out = out.clone()
    out[(idxs < 0)[..., None].expand_as(out)] = 0.0
    return out


# ----------------------------------------------------------------------------------
# torch_scatter 2.1.1
# ----------------------------------------------------------------------------------
def _out_size(index, dim_size):
    if dim_size is not None:
        return int(dim_size)
    return int(index.max()) + 1 if index.numel() else 0


def _expand_index(index, src):
    if index.dim() == src.dim():
        return index
    view = [-1] + [1] * (src.dim() - 1)
    return index.view(view).expand_as(src)


def scatter_add(src, index, dim=0, out=None, dim_size=None):
    """torch_scatter.scatter_add along dim 0; sums in index order; size = dim_size or max+1
    (call sites point_conv.py:87,92, point_ops.py:338-339,352-353)."""
    assert dim in (0, -2) or src.dim() == 1
    size = _out_size(index, dim_size)
    res = torch.zeros((size,) + tuple(src.shape[1:]), dtype=src.dtype)
    return res.scatter_add_(0, _expand_index(index, src), src)


def scatter_mean(src, index, dim=0, out=None, dim_size=None):
    """torch_scatter.scatter_mean: sum / max(count, 1)."""
    size = _out_size(index, dim_size)
    s = scatter_add(src, index, 0, None, size)
    cnt = torch.zeros(size, dtype=src.dtype).scatter_add_(0, index, torch.ones_like(index, dtype=src.dtype))
    cnt = cnt.clamp(min=1)
    return s / cnt.view([-1] + [1] * (src.dim() - 1))


def _scatter_ext(src, index, dim_size, is_max):
    size = _out_size(index, dim_size)
    idx = _expand_index(index, src)
    init = float("-inf") if is_max else float("inf")
    val = torch.full((size,) + tuple(src.shape[1:]), init, dtype=src.dtype)
    val = val.scatter_reduce(0, idx, src, "amax" if is_max else "amin", include_self=True)
    # arg = lowest source row attaining the extremum; empty segment -> src.size(0)
    E = src.shape[0]
    rows = torch.arange(E, dtype=torch.int64).view([-1] + [1] * (src.dim() - 1)).expand_as(src)
    hit = src == val.gather(0, idx)
    cand = torch.where(hit, rows, torch.full_like(rows, E))
    arg = torch.full(val.shape, E, dtype=torch.int64).scatter_reduce(0, idx, cand, "amin", include_self=True)
    val = torch.where(arg == E, torch.zeros_like(val), val)  # empty segments -> 0
    return val, arg


def scatter_max(src, index, dim=0, out=None, dim_size=None):
    """torch_scatter.scatter_max -> (values, argmax); empty segment = (0, src.size(dim)).
    Call sites point_conv.py:81, dgcnn.py:240, pointnet2.py:95-98."""
    return _scatter_ext(src, index, dim_size, True)


def scatter_min(src, index, dim=0, out=None, dim_size=None):
    """torch_scatter.scatter_min -> (values, argmin); ties -> lowest row (fps_ops.py:59).  dim = 1 of a matrix (the second
    reduction of curve_cloud.py:146; a 1-D index broadcasts over the rows): the same along the transposed matrix."""
    if src.dim() == 2 and dim in (1, -1):
        if index.dim() != 1 or index.shape[0] != src.shape[1]:
            raise RuntimeError("The expanded size of the tensor (%d) must match the existing size (%d) at non-singleton "
                               "dimension 1" % (src.shape[1], index.shape[0]))      # torch_scatter.utils.broadcast's failure
        val, arg = _scatter_ext(src.t().contiguous(), index, dim_size, False)
        return val.t().contiguous(), arg.t().contiguous()
    return _scatter_ext(src, index, dim_size, False)


# ----------------------------------------------------------------------------------
# torch_geometric 2.3.0 utilities
# ----------------------------------------------------------------------------------
def segment_softmax(src, index, num_nodes=None):
    """torch_geometric.utils.softmax(src, index, dim=0): exp(x - segmax) / (segsum + 1e-16)
    (point_conv.py:91, dgcnn.py:243)."""
    size = _out_size(index, num_nodes)
    idx = _expand_index(index, src)
    smax = torch.full((size,) + tuple(src.shape[1:]), float("-inf"), dtype=src.dtype)
    smax = smax.scatter_reduce(0, idx, src, "amax", include_self=True)
    ex = (src - smax.gather(0, idx)).exp()
    ssum = scatter_add(ex, index, 0, None, size) + 1e-16
    return ex / ssum.gather(0, idx)


def global_add_pool(x, batch, size=None):
    """torch_geometric.nn.glob.global_add_pool: segment sum sized max+1 (point_ops.py:156-157)."""
    return scatter_add(x, batch, 0, None, size)
