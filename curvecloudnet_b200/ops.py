"""Op layer: one python function per C-ABI kernel group (include/ccn_b200.h), torch tensors in/out.

Everything here runs on the GPU through libccn_b200.so; torch is used only to allocate device
buffers and to read back the few data-dependent sizes (marked "host sync").  Neighbour tables and
CSR pointers are int32 on the device; reference-facing index vectors stay int64.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import List, Optional, Tuple

import torch

from . import native
from .native import call

ACT_NONE, ACT_RELU, ACT_LEAKY = 0, 1, 2
AGG = {"max": 0, "mean": 1, "weighted-sum": 2, "attend": 3, "sum": 4, "weighted-sum-norm": 5}
MAX_SAMPLES = 4096
# Fused edge-step kernels (csrc/edge_fused.cu): "auto" = where they are the faster path today (fused_profitable), True = wherever
# the kernel covers the shapes, False = never (the unfused chain of launches).  All three give bit-identical results.
FUSED_EDGE_STEPS = "auto"
TOPO_ERRORS = {1: "batch vector is not sorted", 2: "curve ids are not sorted within a sample",
               4: "more than %d samples in a batch" % MAX_SAMPLES, 8: "batch ids are not 0..B-1"}


def _f32(t: torch.Tensor) -> torch.Tensor:
    if t.dtype != torch.float32:
        raise TypeError("expected float32, got %s" % t.dtype)
    return t if t.is_contiguous() else t.contiguous()


def _i64(t: torch.Tensor) -> torch.Tensor:
    if t.dtype != torch.int64:
        t = t.to(torch.int64)
    return t if t.is_contiguous() else t.contiguous()


def _rows(t: torch.Tensor) -> Tuple[torch.Tensor, int]:
    """2-D fp32 tensor with unit inner stride -> (tensor, leading dimension)."""
    if t.dtype != torch.float32 or t.dim() != 2:
        raise TypeError("expected a 2-D float32 tensor")
    if t.stride(1) != 1 or (t.shape[0] > 1 and t.stride(0) < t.shape[1]):
        t = t.contiguous()
    return t, (t.stride(0) if t.shape[0] > 1 else max(t.shape[1], 1))


def _empty(shape, dtype, like):
    return torch.empty(shape, dtype=dtype, device=like.device)


def _bits_f32(bits: int) -> float:
    import struct
    return struct.unpack("f", struct.pack("i", bits))[0]


def _raise_topo(flags: int):
    msgs = [m for bit, m in TOPO_ERRORS.items() if flags & bit]
    raise AssertionError("; ".join(msgs))   # the reference asserts sortedness (point_ops.py:31,49)


# ----------------------------------------------------------------------------------------------
# topology
# ----------------------------------------------------------------------------------------------
@dataclass
class SamplePtr:
    dev: torch.Tensor          # int32 (B+1) on the device
    host: List[int]            # same on the host

    @property
    def n_samples(self) -> int:
        return len(self.host) - 1

    @property
    def lengths(self) -> List[int]:
        return [self.host[i + 1] - self.host[i] for i in range(self.n_samples)]


@dataclass
class CurveTopology:
    n: int
    n_curves: int
    max_glob: int
    curve_ord: torch.Tensor    # int32 (n)   dense global curve ordinal
    curve_ptr: torch.Tensor    # int32 (n_curves+1)
    curve_glob: torch.Tensor   # int64 (n)   reference curveidx_local2global
    curve_local: torch.Tensor  # int64 (n)   as given
    sptr: SamplePtr
    batch: Optional[torch.Tensor] = None   # int64 (n)

    @property
    def dense(self) -> bool:
        return self.max_glob == self.n_curves


def curve_csr(batch: torch.Tensor, curve: torch.Tensor) -> CurveTopology:
    """curveidx_local2global + batch2ptr in one pass (one host sync for {B, #curves})."""
    batch, curve = _i64(batch), _i64(curve)
    n = batch.shape[0]
    ord_ = _empty((n,), torch.int32, batch)
    cptr = _empty((n + 1,), torch.int32, batch)
    tail = _empty((MAX_SAMPLES + 1 + 4,), torch.int32, batch)   # sample_ptr followed by counts
    soff = _empty((MAX_SAMPLES + 1,), torch.int64, batch)
    glob = _empty((n,), torch.int64, batch)
    counts = tail[MAX_SAMPLES + 1:]
    call("ccn_curve_csr", batch, curve, n, ord_, cptr, tail, soff, glob, counts, MAX_SAMPLES)
    host = tail.tolist()                                         # host sync
    B, n_curves, flags, max_glob = host[MAX_SAMPLES + 1:MAX_SAMPLES + 5]
    if flags:
        _raise_topo(flags)
    sptr = SamplePtr(tail[:B + 1], host[:B + 1])
    return CurveTopology(n, n_curves, max_glob, ord_, cptr[:n_curves + 1], glob, curve, sptr, batch)


def sample_ptr(batch: torch.Tensor) -> SamplePtr:
    """batch2ptr(batch, with_ends=True) (one host sync)."""
    batch = _i64(batch)
    n = batch.shape[0]
    tail = _empty((MAX_SAMPLES + 1 + 4,), torch.int32, batch)
    call("ccn_sample_ptr", batch, n, tail, tail[MAX_SAMPLES + 1:], MAX_SAMPLES)
    host = tail.tolist()
    B, _, flags, _ = host[MAX_SAMPLES + 1:MAX_SAMPLES + 5]
    if flags:
        _raise_topo(flags)
    return SamplePtr(tail[:B + 1], host[:B + 1])


def get_curves(points: torch.Tensor, beam_ids: torch.Tensor, curve_thresh: float = 0.08, n_beams: Optional[int] = None):
    """_get_curves (nuscenes_dataset.py:91-118): -> (reorder i64, inv_reorder i64, curve_idxs i64, sorted xyz f32 (n,3)).
    n_beams: beam ids lie in [0, n_beams) (None: read max(beam)+1 back, one extra host sync)."""
    if points.dtype != torch.float32 or points.dim() != 2 or points.shape[1] < 3:
        raise TypeError("points must be float32 (n, >=3)")
    points, ldp = _rows(points)
    beam_ids = _i64(beam_ids)
    n = points.shape[0]
    if n_beams is None:
        n_beams = int(beam_ids.max().item()) + 1 if n else 1         # host sync
    scratch = _empty((n_beams + max(n, 1),), torch.int32, points)
    reorder = _empty((n,), torch.int64, points)
    inv = _empty((n,), torch.int64, points)
    sorted_pts = _empty((n, 3), torch.float32, points)
    cid = _empty((n,), torch.int64, points)
    counts = _empty((2,), torch.int32, points)
    if n == 0:
        counts.zero_()
        return reorder, inv, cid, sorted_pts, counts
    call("ccn_get_curves", points, ldp, beam_ids, n, int(n_beams), float(curve_thresh), scratch, reorder, inv, sorted_pts,
         cid, counts)
    return reorder, inv, cid, sorted_pts, counts


A2D2_KNN = (4, 4, 4, 3, 4)                       # audi_dataset.py:20
A2D2_CURVE_THRESHES = (0.1, 0.17, 0.1, 0.12, 0.1)   # audi_dataset.py:19


def a2d2_curves(points: torch.Tensor, lidar_id: torch.Tensor, timestamp: torch.Tensor, knn=A2D2_KNN,
                curve_threshes=A2D2_CURVE_THRESHES):
    """SemA2D2._get_curves (audi_dataset.py:156-243): -> (reorder i64 (n), curve_idxs i64 (n), sorted xyz f32 (n,3),
    counts i32 [2] = {number of curves, error flags}); everything in the final (sensor-major) order."""
    import ctypes
    if points.dtype != torch.float32 or points.dim() != 2 or points.shape[1] < 3:
        raise TypeError("points must be float32 (n, >=3)")
    if len(knn) != len(curve_threshes):
        raise ValueError("one k and one threshold per sensor")
    points, ldp = _rows(points)
    lidar_id, timestamp = _i64(lidar_id), _i64(timestamp)
    n, ns = points.shape[0], len(knn)
    nbytes = int(native.load().ccn_a2d2_scratch_bytes(n, ns))
    if nbytes < 0:
        raise native.CcnError("ccn_a2d2_scratch_bytes: invalid argument")
    scratch = _empty((nbytes,), torch.uint8, points)
    reorder = _empty((n,), torch.int64, points)
    sorted_pts = _empty((n, 3), torch.float32, points)
    cid = _empty((n,), torch.int64, points)
    counts = _empty((2,), torch.int32, points)
    if n == 0:
        counts.zero_()
        return reorder, cid, sorted_pts, counts
    k_host = (ctypes.c_int32 * ns)(*[int(k) for k in knn])
    t_host = (ctypes.c_float * ns)(*[float(t) for t in curve_threshes])
    call("ccn_a2d2_curves", points, ldp, lidar_id, timestamp, n, ns, ctypes.addressof(k_host), ctypes.addressof(t_host),
         scratch, reorder, sorted_pts, cid, counts)
    return reorder, cid, sorted_pts, counts


def _finish_subset(parent_batch, idx_buf, count_dev, n_samples) -> Tuple[torch.Tensor, SamplePtr]:
    """Read back {sample_ptr of the selected level, count} with ONE host sync; trims idx."""
    out = _empty((n_samples + 1,), torch.int32, idx_buf)
    call("ccn_subset_sample_ptr", parent_batch, idx_buf, count_dev, idx_buf.shape[0], n_samples, out)
    host = out.tolist()                                          # host sync
    return idx_buf[:host[-1]], SamplePtr(out, host)


# ----------------------------------------------------------------------------------------------
# curve convolution
# ----------------------------------------------------------------------------------------------
def curve_feat_prep(x: Optional[torch.Tensor], pos: torch.Tensor, with_xyz: bool, with_diff: bool,
                    curve_ord: torch.Tensor) -> torch.Tensor:
    pos = _f32(pos)
    n = pos.shape[0]
    cx = 0 if x is None else x.shape[1]
    c_in = cx + (3 if with_xyz else 0)
    out = _empty((n, c_in * (2 if with_diff else 1)), torch.float32, pos)
    if x is not None:
        x, ldx = _rows(x)
    else:
        ldx = 0
    call("ccn_curve_feat_prep", x if x is not None else 0, ldx, cx, pos, int(with_xyz), int(with_diff), curve_ord, n,
         out, out.shape[1])
    return out


def curve_row_map(topo: CurveTopology, pad: int) -> torch.Tensor:
    rows = topo.n + (topo.n_curves + 1) * pad
    row2pt = _empty((max(rows, 1),), torch.int32, topo.curve_ord)
    call("ccn_curve_row_map", topo.curve_ord, topo.n, topo.n_curves, pad, row2pt)
    return row2pt[:rows]


def gather_rows_i32(src: torch.Tensor, idx: torch.Tensor) -> torch.Tensor:
    """out[r] = src[idx[r]] if idx[r] >= 0 else 0   (pad-row insertion of fast_conv1d.py:48-61); idx int32."""
    src, lds = _rows(src)
    out = _empty((idx.shape[0], src.shape[1]), torch.float32, src)
    if idx.shape[0]:
        call("ccn_gather_rows_i32", src, lds, idx, idx.shape[0], src.shape[1], out, src.shape[1])
    return out


def curve_conv_layer(inp: torch.Tensor, w: torch.Tensor, bias: torch.Tensor, row2pt: torch.Tensor, rows: int,
                     out_rows: int, ksize: int, in_mapped: bool, out_mapped: bool, slope: float = 0.01) -> torch.Tensor:
    """w: (ksize, cin, cout) folded, bias: (cout) folded."""
    inp, ld_in = _rows(inp)
    cin, cout = w.shape[1], w.shape[2]
    out = _empty((out_rows, cout), torch.float32, inp)
    call("ccn_curve_conv_layer", inp, ld_in, w, bias, out, cout, row2pt, rows, cin, cout, ksize, int(in_mapped),
         int(out_mapped), float(slope))
    return out


# ----------------------------------------------------------------------------------------------
# sampling / grouping along curves
# ----------------------------------------------------------------------------------------------
def curve_fps(pos: torch.Tensor, topo: CurveTopology, arclen: float, u: torch.Tensor) -> Tuple[torch.Tensor, SamplePtr]:
    """CurveFPS.forward -> (idx int64 ascending, sample ptr of the kept points)."""
    pos = _f32(pos)
    n = pos.shape[0]
    run = _empty((max(n, 1) + (n + 4095) // 4096,), torch.float32, pos)   # running arclength + per-tile counts
    idx = _empty((max(n, 1),), torch.int64, pos)
    cnt = _empty((1,), torch.int32, pos)
    u = u.to(device=pos.device, dtype=torch.float32).reshape(1)
    call("ccn_curve_fps", pos, topo.curve_ord, topo.curve_ptr, n, u, float(arclen), run, idx, cnt)
    return _finish_subset(topo.batch, idx, cnt, topo.sptr.n_samples)


@dataclass
class EdgeCSR:
    row_ptr: torch.Tensor      # int32 (nq+1)
    erow: torch.Tensor         # int32 (E) target / query ordinal of every edge
    ecol: torch.Tensor         # int32 (E) source point index
    n_edges: int
    nq: int
    max_degree: int = 1 << 30  # upper bound of the edges per target (the fused edge kernels tile by it)


def radius_1d_group(pos: torch.Tensor, topo: CurveTopology, idxs: torch.Tensor, radius: float) -> EdgeCSR:
    """radius_1d_group_subset as CSR over the queries (one host sync for E)."""
    if not topo.dense:
        raise NotImplementedError("radius_1d_group: curve ids with gaps inside a sample are not supported")
    pos, idxs = _f32(pos), _i64(idxs)
    nq = idxs.shape[0]
    kc = _empty((max(topo.n_curves, 1),), torch.float32, pos)
    buf = _empty((nq + 4,), torch.int32, pos)                    # row_ptr (nq+1), total, stats (2): one read-back
    row_ptr, total, stats = buf[:nq + 2], buf[nq + 1:nq + 2], buf[nq + 2:]
    call("ccn_radius_1d_group_count", pos, topo.curve_ord, topo.curve_ptr, topo.curve_local, topo.n, topo.n_curves,
         idxs, nq, float(radius), kc, stats, row_ptr, total)
    E, kbits, nbits = buf[nq + 1:].tolist()                      # host sync
    # a query keeps at most k_c candidates, k_c <= max over the curves (stats[0], float bits), and no more than a curve has
    max_deg = max(1, int(min(_bits_f32(kbits), _bits_f32(nbits))))
    col = _empty((max(E, 1),), torch.int32, pos)
    call("ccn_radius_1d_group_fill", topo.curve_ord, topo.curve_local, topo.n, topo.n_curves, idxs, nq, kc, stats,
         row_ptr, col)
    erow = _empty((max(E, 1),), torch.int32, pos)
    call("ccn_csr_rows", row_ptr, nq, erow)
    return EdgeCSR(row_ptr[:nq + 1], erow[:E], col[:E], E, nq, max_deg)


def knn_1d_group(curve_glob: torch.Tensor, idxs: torch.Tensor, knn: int) -> EdgeCSR:
    """knn_1d_group_subset (point_ops.py:114-140) as CSR over the queries (one host sync for E)."""
    curve_glob, idxs = _i64(curve_glob), _i64(idxs)
    nq = idxs.shape[0]
    W = 2 * knn + 1
    idx = _empty((nq, W), torch.int32, curve_glob)
    cnt = _empty((nq + 2,), torch.int32, curve_glob)
    call("ccn_knn_1d_group", curve_glob, curve_glob.shape[0], idxs, nq, knn, idx, cnt)
    return dense_to_csr(idx, cnt)


def group_first_k(values: torch.Tensor, batch: torch.Tensor, n_groups: int, kmax: int):
    """First kmax rows of every group id in order of appearance -> (out (n_groups, kmax, d) zero padded, counts (n_groups,))."""
    values, ldv = _rows(values)
    d = values.shape[1]
    out = _empty((n_groups, kmax, d), torch.float32, values)
    cnt = _empty((max(n_groups, 1),), torch.int32, values)
    call("ccn_group_first_k", values, ldv, _i64(batch), values.shape[0], d, n_groups, kmax, out, cnt)
    return out, cnt[:n_groups]


def knn_1d_superset(pos: torch.Tensor, curve_ord: torch.Tensor, idxs: torch.Tensor, knn: int):
    pos, idxs = _f32(pos), _i64(idxs)
    n = pos.shape[0]
    nbr = _empty((n, knn), torch.int32, pos)
    wgt = _empty((n, knn), torch.float32, pos)
    call("ccn_knn_1d_superset", pos, curve_ord, n, idxs, idxs.shape[0], knn, nbr, wgt)
    return nbr, wgt


# ----------------------------------------------------------------------------------------------
# point down-sampling
# ----------------------------------------------------------------------------------------------
def voxel_fps(pos: torch.Tensor, batch: torch.Tensor, voxel: float, noise: torch.Tensor,
              n_samples: int) -> Tuple[torch.Tensor, SamplePtr]:
    """VoxelFPS.forward -> (idx int64 in voxel-key order, sample ptr of the kept points)."""
    pos, batch = _f32(pos), _i64(batch)
    n = pos.shape[0]
    slots = native.voxel_table_slots(n)
    table = _empty((2 * slots + max(n, 1),), torch.int64, pos)
    ints = _empty((2 * max(n, 1) + 2,), torch.int32, pos)
    idx = _empty((max(n, 1),), torch.int64, pos)
    noise = noise.to(device=pos.device, dtype=torch.float32).contiguous()
    counts = ints[2 * max(n, 1):]
    call("ccn_voxel_fps", pos, batch, n, float(voxel), noise, slots, table, table[slots:], table[2 * slots:],
         ints, ints[max(n, 1):], idx, counts)
    idx, sptr = _finish_subset(batch, idx, counts, n_samples)
    if int(counts[1].item()):
        raise native.CcnError("voxel_fps: voxel coordinates out of the supported range (|ix| < 2^17, B <= 1024)")
    return idx, sptr


def fps_count(lengths: List[int], ratio: float) -> List[int]:
    """K = ceil(len * ratio) evaluated in float32 like the reference (point_ops.py:63)."""
    t = torch.ceil(torch.tensor(lengths, dtype=torch.int64) * ratio).to(torch.int64)
    return [int(v) for v in t.tolist()]


def fps(pos: torch.Tensor, sptr: SamplePtr, ksel: List[int], starts: List[int]) -> torch.Tensor:
    """sample_farthest_points per cloud from explicit start points; returns the selected global
    indices sorted ascending (what fps_pytorch3d returns, point_ops.py:69).  No host sync."""
    pos = _f32(pos)
    n = pos.shape[0]
    B = sptr.n_samples
    lens = sptr.lengths
    total = sum(min(k, l) for k, l in zip(ksel, lens))
    meta = torch.tensor(list(ksel) + list(starts), dtype=torch.int32).to(pos.device, non_blocking=True)
    max_len = max(lens) if lens else 0
    closest = _empty((max(n, 1),), torch.float32, pos) if max_len > 512 * 16 else None
    sel = _empty((max(n, 1),), torch.int32, pos)
    idx = _empty((max(total, 1),), torch.int64, pos)
    call("ccn_fps", pos, sptr.dev, B, n, meta, meta[B:], max_len, closest if closest is not None else 0, sel, idx, 0)
    return idx[:total]


# ----------------------------------------------------------------------------------------------
# neighbour search
# ----------------------------------------------------------------------------------------------
def frnn_knn(pos1, batch1, pos2, batch2, sptr2: SamplePtr, K: int, radius: float, want_d2=False, want_cnt=False,
             cache=None):
    """Fixed-radius kNN -> idx int32 (n1, K) GLOBAL rows of pos2, -1 padded, ascending by (d2, index).
    `cache` (nn.ForwardCache): the hashed grid of (pos2, radius) and, for plain index queries, the whole
    result are memoised for the forward - the reference rebuilds both at every step that searches the
    same level with the same (K, r)."""
    pos1, pos2, batch1, batch2 = _f32(pos1), _f32(pos2), _i64(batch1), _i64(batch2)
    n1, n2 = pos1.shape[0], pos2.shape[0]
    radius = float(radius)
    rkey = (pos1.data_ptr(), pos2.data_ptr(), n1, n2, K, radius)
    if cache is not None and not (want_d2 or want_cnt):
        hit = cache.knn.get(rkey)
        if hit is not None:
            return hit[0], None, None
    gkey = (pos2.data_ptr(), n2, radius)
    grid = cache.grids.get(gkey) if cache is not None else None
    if grid is None:
        cells = native.grid_num_cells(n2)
        cell_off = _empty((cells + 1,), torch.int32, pos1)
        slot = _empty((max(n2, 1),), torch.int32, pos1)
        sorted_ = _empty((max(n2, 1), 4), torch.float32, pos1)
        call("ccn_grid_build", pos2, batch2, n2, radius, cells, cell_off, slot, sorted_)
        grid = (cells, cell_off, sorted_, pos2)
        if cache is not None:
            cache.grids[gkey] = grid
    cells, cell_off, sorted_, _ = grid
    idx = _empty((n1, K), torch.int32, pos1)
    d2 = _empty((n1, K), torch.float32, pos1) if want_d2 else None
    cnt = _empty((n1 + 2,), torch.int32, pos1) if want_cnt else None
    call("ccn_frnn_knn", pos1, batch1, n1, sorted_, cell_off, cells, sptr2.dev, radius, K, idx,
         d2 if want_d2 else 0, cnt if want_cnt else 0)
    if cache is not None and not (want_d2 or want_cnt):
        cache.knn[rkey] = (idx, pos1, pos2)
    return idx, d2, cnt


# exact 3-NN through the hashed grid pays from ~7e8 query x source pairs on (tools/bench_knn.py, profiles/r02o_knn.txt:
# 113 k x 57 k: 4.24 ms brute force -> 0.77 ms; 32 k x 16 k: 0.43 ms brute force, 0.70 ms grid)
KNN_GRID_MIN_PAIRS = 7e8


def knn_exact(pos1, batch1, pos2, sptr2: SamplePtr, K: int, zero_fill=True, want_d2=True, want_cnt=False, batch2=None,
              use_grid=None):
    """pytorch3d.ops.knn_points (point_ops.py:91): exact K nearest by (d2, index).  K <= 4 with `batch2` given: through the
    hashed grid (ccn_knn_grid, same result bit for bit) when the source cloud is large (use_grid None) or forced."""
    pos1, pos2, batch1 = _f32(pos1), _f32(pos2), _i64(batch1)
    n1, n2 = pos1.shape[0], pos2.shape[0]
    idx = _empty((n1, K), torch.int32, pos1)
    d2 = _empty((n1, K), torch.float32, pos1) if want_d2 else None
    cnt = _empty((n1 + 2,), torch.int32, pos1) if want_cnt else None
    if use_grid is None:
        use_grid = float(n1) * float(n2) >= KNN_GRID_MIN_PAIRS
    if use_grid and batch2 is not None and K <= 4 and n1 > 0:
        cells = native.grid_num_cells(n2)
        cell_off = _empty((cells + 1,), torch.int32, pos1)
        slot = _empty((max(n2, 1),), torch.int32, pos1)
        sorted_ = _empty((max(n2, 1), 4), torch.float32, pos1)
        scratch = _empty((16 + n1,), torch.int32, pos1)
        call("ccn_knn_grid", pos1, batch1, n1, pos2, _i64(batch2), n2, sptr2.dev, K, int(zero_fill), 0.0, cells, cell_off, slot,
             sorted_, scratch, idx, d2 if want_d2 else 0, cnt if want_cnt else 0)
        return (idx, d2, cnt) if want_cnt else (idx, d2)
    call("ccn_knn_exact", pos1, batch1, n1, pos2, sptr2.dev, K, int(zero_fill), idx, d2 if want_d2 else 0,
         cnt if want_cnt else 0)
    if want_cnt:
        return idx, d2, cnt
    return idx, d2


def ball_query(pos1, batch1, pos2, sptr2: SamplePtr, radius: float, K: int = 128):
    pos1, pos2, batch1 = _f32(pos1), _f32(pos2), _i64(batch1)
    n1 = pos1.shape[0]
    idx = _empty((n1, K), torch.int32, pos1)
    cnt = _empty((n1 + 2,), torch.int32, pos1)
    call("ccn_ball_query", pos1, batch1, n1, pos2, sptr2.dev, float(radius), K, idx, cnt)
    return idx, cnt


def feat_neighbours(x1, batch1, x2, sptr2: SamplePtr, K: int, radius: float, mode: str, want_d2: bool = False):
    """Neighbours in FEATURE space (DynamicEdgeConv, dgcnn.py:64-89): mode "knn" = fixed-radius kNN by (d2, index)
    (fast_knn on D-dimensional rows, K <= 32), "ball" = first K in index order inside the radius (ball_query)."""
    x1, ld1 = _rows(x1)
    x2, ld2 = _rows(x2)
    if x1.shape[1] != x2.shape[1]:
        raise ValueError("dimension mismatch: points1 of dimension %d while points2 of dimension %d" % (x1.shape[1], x2.shape[1]))
    n1 = x1.shape[0]
    idx = _empty((n1, K), torch.int32, x1)
    cnt = _empty((n1 + 2,), torch.int32, x1)
    d2 = _empty((n1, K), torch.float32, x1) if want_d2 else None
    call("ccn_feat_neighbours", x1, ld1, _i64(batch1), n1, x2, ld2, sptr2.dev, x1.shape[1], float(radius), K,
         {"knn": 0, "ball": 1}[mode], idx, d2 if d2 is not None else 0, cnt)
    return (idx, cnt, d2) if want_d2 else (idx, cnt)


def dense_to_csr(idx: torch.Tensor, cnt: torch.Tensor) -> EdgeCSR:
    """(nq, K) table with leading valid entries + per-row counts (nq+2 buffer) -> CSR (one host sync for E)."""
    nq, K = idx.shape
    call("ccn_excl_scan_i32", cnt, nq, cnt[nq + 1:])
    E = int(cnt[nq + 1].item())                                  # host sync
    erow = _empty((max(E, 1),), torch.int32, idx)
    ecol = _empty((max(E, 1),), torch.int32, idx)
    call("ccn_dense_to_csr", idx, cnt, nq, K, erow, ecol)
    return EdgeCSR(cnt[:nq + 1], erow[:E], ecol[:E], E, nq, K)


# ----------------------------------------------------------------------------------------------
# gathers, aggregation, MLP layer
# ----------------------------------------------------------------------------------------------
def gather_rows(src: torch.Tensor, idx: torch.Tensor) -> torch.Tensor:
    src, lds = _rows(src)
    idx = _i64(idx)
    out = _empty((idx.shape[0], src.shape[1]), torch.float32, src)
    call("ccn_gather_rows_f32", src, lds, idx, idx.shape[0], src.shape[1], out, src.shape[1])
    return out


def gather_i64(src: torch.Tensor, idx: torch.Tensor) -> torch.Tensor:
    src, idx = _i64(src), _i64(idx)
    out = _empty((idx.shape[0],), torch.int64, src)
    call("ccn_gather_i64", src, idx, idx.shape[0], out)
    return out


def subset_gather(pos: torch.Tensor, batch: torch.Tensor, curve: Optional[torch.Tensor], idx: torch.Tensor, npl: int = 0,
                  cache=None):
    """(pos[idx], batch[idx], curve[idx]) in one launch; curve may be None.  npl / cache: the activation planes of the new
    position matrix are written by the same launch and registered in the forward cache."""
    pos, ldp = _rows(_f32(pos))
    batch, idx = _i64(batch), _i64(idx)
    n = idx.shape[0]
    posq = _empty((n, 3), torch.float32, pos)
    batchq = _empty((n,), torch.int64, pos)
    curveq = _empty((n,), torch.int64, pos) if curve is not None else None
    if n:
        pl = alloc_planes(n, 3, npl, pos) if npl and cache is not None else None
        call("ccn_subset_gather", pos, ldp, batch, _i64(curve) if curve is not None else 0, idx, n, posq, batchq,
             curveq if curveq is not None else 0, pl.buf if pl is not None else 0, pl.plane_stride if pl is not None else 0,
             npl)
        if pl is not None:
            cache.put_planes(posq, pl)
    return posq, batchq, curveq


def concat_cols(parts: List[torch.Tensor]) -> torch.Tensor:
    """torch.cat(parts, dim=1) with our copy kernel (column-offset strided writes)."""
    n = parts[0].shape[0]
    c = sum(p.shape[1] for p in parts)
    out = _empty((n, c), torch.float32, parts[0])
    off = 0
    for p in parts:
        p, ld = _rows(p)
        call("ccn_gather_rows_f32", p, ld, 0, n, p.shape[1], out[:, off:], c)
        off += p.shape[1]
    return out


def linear(A: torch.Tensor, W: torch.Tensor, bias: Optional[torch.Tensor], act: int = ACT_NONE,
           slope: float = 0.01) -> torch.Tensor:
    """Y = act(A W^T + bias), fp32."""
    A, lda = _rows(A)
    W, ldw = _rows(W)
    M, K = A.shape
    N = W.shape[0]
    assert W.shape[1] == K, (A.shape, W.shape)
    Y = _empty((M, N), torch.float32, A)
    call("ccn_linear_f32", A, lda, W, ldw, bias if bias is not None else 0, Y, N, M, N, K, act, float(slope))
    return Y


def edge_layer1(H: Optional[torch.Tensor], wp: torch.Tensor, bias: Optional[torch.Tensor], pos_src, pos_dst,
                edges: EdgeCSR, normalize_radius: Optional[float], act: int, slope: float = 0.01, n_planes: int = 0):
    """-> fp32 (E, c), or activation planes when n_planes > 0 (c % 8 == 0)."""
    c = wp.shape[1]
    if H is not None:
        H, ldh = _rows(H)
    else:
        ldh = 0
    if n_planes and c % 16 == 0 and (H is None or ldh % 4 == 0):
        out = alloc_planes(edges.n_edges, c, n_planes, wp)
        call("ccn_edge_layer1_planes", H if H is not None else 0, ldh, wp, bias if bias is not None else 0, _f32(pos_src),
             _f32(pos_dst), edges.erow, edges.ecol, edges.n_edges, c, float(normalize_radius or 0.0), act, float(slope),
             out.buf, out.kpad, out.plane_stride, n_planes)
        return out
    out = _empty((edges.n_edges, c), torch.float32, wp)
    call("ccn_edge_layer1", H if H is not None else 0, ldh, wp, bias if bias is not None else 0, _f32(pos_src),
         _f32(pos_dst), edges.erow, edges.ecol, edges.n_edges, c, float(normalize_radius or 0.0), act, float(slope),
         out, c)
    return out


def edge_pair_layer1(PQ: torch.Tensor, edges: EdgeCSR, act: int, slope: float = 0.01, n_planes: int = 0):
    PQ, ld = _rows(PQ)
    c = PQ.shape[1] // 2
    if n_planes and c % 16 == 0 and ld % 4 == 0:
        out = alloc_planes(edges.n_edges, c, n_planes, PQ)
        call("ccn_edge_pair_layer1_planes", PQ, ld, edges.erow, edges.ecol, edges.n_edges, c, act, float(slope), out.buf,
             out.kpad, out.plane_stride, n_planes)
        return out
    out = _empty((edges.n_edges, c), torch.float32, PQ)
    call("ccn_edge_pair_layer1", PQ, ld, edges.erow, edges.ecol, edges.n_edges, c, act, float(slope), out, c)
    return out


def _both_ok(npl: int, cache, c: int, *mats) -> bool:
    """A producer may write its fp32 result AND the activation planes in one launch (ccn_*_both): tensor-core mode, a
    forward cache to register the planes in, c % 16 == 0 and 16-byte aligned fp32 rows."""
    if not npl or cache is None or c % 16:
        return False
    for t, ld in mats:
        if t is not None and (ld % 4 or t.data_ptr() % 16):
            return False
    return True


def seg_aggregate(msg: torch.Tensor, att: Optional[torch.Tensor], row_ptr: torch.Tensor, nq: int, mode: str,
                  npl: int = 0, cache=None):
    """npl / cache: on the tensor-core modes the result's activation planes are written by the same launch and
    registered in the forward cache (the consumers of a step output are Linear layers)."""
    msg, ldm = _rows(msg)
    c = msg.shape[1]
    if att is not None:
        att, lda = _rows(att)
    else:
        lda = 0
    out = _empty((nq, c), torch.float32, msg)
    if nq and _both_ok(npl, cache, c, (msg, ldm), (att, lda)):
        pl = alloc_planes(nq, c, npl, msg)
        call("ccn_seg_aggregate_both", msg, ldm, att if att is not None else 0, lda, row_ptr, nq, c, AGG[mode], out, c,
             pl.buf, pl.kpad, pl.plane_stride, npl)
        cache.put_planes(out, pl)
        return out
    call("ccn_seg_aggregate", msg, ldm, att if att is not None else 0, lda, row_ptr, nq, c, AGG[mode], out, c)
    return out


def sgcnn_layer1(PQ: torch.Tensor, idx: torch.Tensor, act: int, slope: float = 0.01, n_planes: int = 0):
    PQ, ld = _rows(PQ)
    n, K = idx.shape
    c = PQ.shape[1] // 2
    if n_planes and c % 16 == 0 and ld % 4 == 0:
        out = alloc_planes(n * (K + 1), c, n_planes, PQ)
        call("ccn_sgcnn_layer1_planes", PQ, ld, idx, n, K, c, act, float(slope), out.buf, out.kpad, out.plane_stride,
             n_planes)
        return out
    out = _empty((n * (K + 1), c), torch.float32, PQ)
    call("ccn_sgcnn_layer1", PQ, ld, idx, n, K, c, act, float(slope), out, c)
    return out


def group_max_dense(feat: torch.Tensor, idx: torch.Tensor, npl: int = 0, cache=None) -> torch.Tensor:
    feat, ld = _rows(feat)
    n, K = idx.shape
    c = feat.shape[1]
    out = _empty((n, c), torch.float32, feat)
    if n and _both_ok(npl, cache, c, (feat, ld)):
        pl = alloc_planes(n, c, npl, feat)
        call("ccn_group_max_dense_both", feat, ld, idx, n, K, c, out, c, pl.buf, pl.kpad, pl.plane_stride, npl)
        cache.put_planes(out, pl)
        return out
    call("ccn_group_max_dense", feat, ld, idx, n, K, c, out, c)
    return out


def interp_blend(x: torch.Tensor, nbr: torch.Tensor, wgt: torch.Tensor, from_d2: bool, npl: int = 0, cache=None) -> torch.Tensor:
    x, ldx = _rows(x)
    n, knn = nbr.shape
    c = x.shape[1]
    out = _empty((n, c), torch.float32, x)
    if n and _both_ok(npl, cache, c, (x, ldx)):
        pl = alloc_planes(n, c, npl, x)
        call("ccn_interp_blend_both", x, ldx, nbr, wgt, int(from_d2), n, knn, c, out, c, pl.buf, pl.kpad, pl.plane_stride,
             npl)
        cache.put_planes(out, pl)
        return out
    call("ccn_interp_blend", x, ldx, nbr, wgt, int(from_d2), n, knn, c, out, c)
    return out


def sample_ptr_any(ids: torch.Tensor) -> torch.Tensor:
    """batch2ptr(ids, with_ends=True) for ANY sorted int64 id vector (run starts + [N]) as int64."""
    ids = _i64(ids)
    zeros = torch.zeros_like(ids)
    topo = curve_csr(zeros, ids)
    return topo.curve_ptr.to(torch.int64)


# ----------------------------------------------------------------------------------------------
# tensor-core Linear (tcgen05): weight preparation + launch
# ----------------------------------------------------------------------------------------------
class TcWeight:
    """bf16 hi/lo copies of a folded fp32 weight (N, K), K padded to 64, and their host-side TMA descriptors
    (one pair per N-tile width)."""

    def __init__(self, w: torch.Tensor, seg_widths=None):
        """seg_widths: the K axis is a concatenation of blocks of these widths; every block's columns are placed at
        a multiple of 64 (zero columns in between) so that each block can be its own TMA source (ccn_linear_planes)."""
        if seg_widths is not None and len(seg_widths) > 1:
            assert sum(seg_widths) == w.shape[1]
            offs, tot = [], 0
            for kw in seg_widths:
                offs.append(tot)
                tot += (kw + 63) // 64 * 64
            wide = torch.zeros((w.shape[0], tot), dtype=torch.float32, device=w.device)   # weight preparation
            c0 = 0
            for kw, off in zip(seg_widths, offs):
                wide[:, off:off + kw] = w[:, c0:c0 + kw]
                c0 += kw
            self.k_true, w = w.shape[1], wide
        else:
            self.k_true = w.shape[1]
        w, ldw = _rows(w)
        n, k = w.shape
        self.n, self.k = n, k
        self.kpad = (k + 63) // 64 * 64
        self.max_box = 32 if n <= 32 else 64 if n <= 64 else 128
        buf = torch.empty((2, n, self.kpad), dtype=torch.bfloat16, device=w.device)
        self.hi, self.lo = buf[0], buf[1]
        call("ccn_split_bf16", w, ldw, n, k, self.kpad, self.hi, self.lo)
        self._maps = {}

    def maps(self, box_n: int):
        hit = self._maps.get(box_n)
        if hit is None:
            import ctypes
            lib = native.load()
            bufs, ptrs = [], []
            for t in (self.hi, self.lo):
                b = ctypes.create_string_buffer(128 + 64)
                p = (ctypes.addressof(b) + 63) & ~63                     # CUtensorMap wants 64-byte alignment
                rc = lib.ccn_weight_tensor_map(p, t.data_ptr(), self.n, self.kpad, box_n)
                if rc:
                    raise native.CcnError("ccn_weight_tensor_map failed (code %d)" % rc)
                bufs.append(b)
                ptrs.append(p)
            hit = (ptrs[0], ptrs[1], bufs)
            self._maps[box_n] = hit
        return hit

    def pick_box_planes(self, m: int, n_planes: int) -> int:
        """N-tile width for ccn_linear_planes.  Big M (>= one m-tile per SM): the widest tile whose weight slice stays
        resident in shared memory next to a 3-stage A ring; otherwise the widest tile that still gives >= _BOX_MIN_TILES tiles."""
        cand = [b for b in (256, 128, 64, 32) if b == 32 or b // 2 < self.n]
        tiles_m = (m + 127) // 128
        nk = self.kpad // 64
        if tiles_m >= 148:
            for b in cand:
                if nk * n_planes * b * 128 + 3 * n_planes * 16384 <= SMEM_BUDGET:
                    return b
            return cand[-1]
        for b in cand:
            if tiles_m * ((self.n + b - 1) // b) >= _BOX_MIN_TILES:
                return b
        return 32

    def pick_box(self, m: int) -> int:
        """Widest N tile that still yields >= one wave of CTAs (148 SMs); skinny problems get narrow tiles."""
        box = self.max_box
        mt = (m + 127) // 128
        while box > 32 and mt * ((self.n + box - 1) // box) < 148:
            box //= 2
        return box


SMEM_BUDGET = 227 * 1024 - 2560
# small-M layers: the widest N tile that still gives half a wave of CTAs (measured with tools/bench_linear.py: one wide tile
# per CTA beats two narrow ones that re-read the activation tile, down to ~74 tiles)
_BOX_MIN_TILES = 74


def _tensor_map(base_ptr: int, rows: int, kpad: int, box_rows: int):
    """Host-side CUtensorMap of a bf16 (rows, kpad) row-major matrix, box = 64 x box_rows, SWIZZLE_128B."""
    import ctypes
    b = ctypes.create_string_buffer(128 + 64)
    p = (ctypes.addressof(b) + 63) & ~63
    rc = native.load().ccn_weight_tensor_map(p, base_ptr, rows, kpad, box_rows)
    if rc:
        raise native.CcnError("ccn_weight_tensor_map failed (code %d)" % rc)
    return p, b


class Planes:
    """Split-bf16 activation planes (n_planes, M, kpad): plane 0 = bf16(x), plane 1 = bf16(x - plane 0);
    the inter-layer activation format of the tensor-core MLP path (TMA-loadable, csrc/gemm_planes.cu)."""

    def __init__(self, buf: torch.Tensor, m: int, k: int):
        self.buf, self.m, self.k = buf, m, k
        self.n_planes, self.kpad = buf.shape[0], buf.shape[2]

    @property
    def shape(self):
        return (self.m, self.k)

    @property
    def plane_stride(self):
        return self.buf.stride(0)

    def to_f32(self) -> torch.Tensor:
        """Debug/test helper (torch arithmetic, not on the product path)."""
        x = self.buf[0, :self.m, :self.k].float()
        return x + self.buf[1, :self.m, :self.k].float() if self.n_planes == 2 else x


def alloc_planes(m: int, k: int, n_planes: int, like: torch.Tensor) -> Planes:
    kpad = (k + 15) // 16 * 16        # 32-byte rows: the Linear epilogue writes whole sectors
    return Planes(torch.empty((n_planes, max(m, 1), kpad), dtype=torch.bfloat16, device=like.device), m, k)


def pack_planes(parts, n_planes: int, idxs=None) -> Planes:
    """torch.cat([p[idx] for p, idx in ...], dim=1) written directly as activation planes (up to four blocks)."""
    assert 1 <= len(parts) <= 4
    idxs = list(idxs) if idxs is not None else [None] * len(parts)
    n = parts[0].shape[0] if idxs[0] is None else idxs[0].shape[0]
    args, ctot = [], 0
    for k in range(4):
        if k < len(parts):
            t, ld = _rows(parts[k])
            ix = _i64(idxs[k]) if idxs[k] is not None else 0
            args += [t, ld, t.shape[1], ix]
            ctot += t.shape[1]
        else:
            args += [0, 0, 0, 0]
    out = alloc_planes(n, ctot, n_planes, parts[0])
    call("ccn_pack_planes", *args, n, out.buf, out.kpad, out.plane_stride, n_planes)
    return out


class PlanesCat:
    """[A_0 | A_1 | ...]: a column concatenation of activation planes that is never materialised - every block is
    its own K-segment (TMA source) of ccn_linear_planes."""

    def __init__(self, segs):
        assert 1 <= len(segs) <= 4 and all(s.m == segs[0].m and s.n_planes == segs[0].n_planes for s in segs)
        self.segs = list(segs)
        self.m, self.k, self.n_planes = segs[0].m, sum(s.k for s in segs), segs[0].n_planes

    @property
    def shape(self):
        return (self.m, self.k)

    @property
    def widths(self):
        return tuple(s.k for s in self.segs)

    def to_f32(self):
        return torch.cat([s.to_f32() for s in self.segs], dim=1)


def linear_planes(A, tw: "TcWeight", bias: Optional[torch.Tensor], act: int, slope: float, out: str = "f32"):
    """Y = act(A W^T + bias) with both operands by TMA; A: Planes or PlanesCat (tw prepared for the same widths).
    out: "f32" -> tensor, "planes" -> Planes, "both" -> (tensor, Planes)."""
    segs = A.segs if isinstance(A, PlanesCat) else (A,)
    M, K = A.shape
    assert K == tw.k_true, (A.shape, tw.k_true)
    npl = A.n_planes
    dev = segs[0].buf
    Y = _empty((M, tw.n), torch.float32, dev) if out != "planes" else None
    Yp = alloc_planes(M, tw.n, npl, dev) if out != "f32" else None
    box = tw.pick_box_planes(M, npl)
    w_hi, w_lo, _ = tw.maps(box)
    a = []
    for sg in segs:
        a += [sg.buf, sg.kpad, sg.buf.stride(0), sg.k]
    a += [0, 0, 0, 0] * (4 - len(segs))
    call("ccn_linear_planes", *a, w_hi, w_lo, bias if bias is not None else 0, Y if Y is not None else 0, tw.n,
         Yp.buf if Yp is not None else 0, Yp.kpad if Yp is not None else 0, Yp.buf.stride(0) if Yp is not None else 0,
         M, tw.n, box, 3 if npl == 2 else 1, act, float(slope))
    return Y if out == "f32" else Yp if out == "planes" else (Y, Yp)


def linear_planes_bt(A: "Planes", tw: "TcWeight") -> torch.Tensor:
    """Y (M, K_w) = A (M, N_w) @ W with W = the (N_w, K_w) weight behind `tw`, read as MN-major tiles from its FORWARD
    planes (ccn_linear_planes_bt): the data gradient dX = dY W without a transposed copy of W."""
    M, K = A.shape
    assert K == tw.n and tw.k == tw.k_true, (A.shape, tw.n, tw.k)
    npl = A.n_planes
    n_out = tw.k
    Y = _empty((M, n_out), torch.float32, A.buf)
    # output tile width: the widest of 256 / 128 / 64 whose weight slice stays resident (big M) or that fills the SMs
    tiles_m = (M + 127) // 128
    nk = (tw.n + 63) // 64
    cand = [b for b in (256, 128, 64) if b == 64 or b // 2 < n_out]
    box = cand[-1]
    for b in cand:
        if tiles_m >= 148:
            if nk * npl * b * 128 + 3 * npl * 16384 <= SMEM_BUDGET:
                box = b
                break
        elif tiles_m * ((n_out + b - 1) // b) >= 148:
            box = b
            break
    w_hi, w_lo, _ = tw.maps(64)
    call("ccn_linear_planes_bt", A.buf, A.kpad, A.buf.stride(0), A.k, w_hi, w_lo, Y, n_out, 0, 0, 0, M, n_out, box,
         3 if npl == 2 else 1)
    return Y


def conv_planes(sig: "Planes", w_kio: torch.Tensor, bias: Optional[torch.Tensor], act: int = ACT_NONE, slope: float = 0.01,
                tw: "TcWeight" = None, out: str = "f32", out_map: Optional[torch.Tensor] = None, out_rows: int = 0):
    """'same' zero-padded 1-D convolution along the rows of `sig` (planes of a (rows, cin) signal) as ONE tcgen05 GEMM:
    w_kio (taps, cin, cout); tap t of output row m reads row m + t - taps/2 (ccn_conv_planes).  out = "f32": fp32
    (rows, cout); "planes": activation planes for the next layer of the stack; "both": (fp32, planes).  out_map (int32 (rows), with out_rows):
    output row m is stored at row out_map[m] and dropped when negative (the last layer keeps the curve points only)."""
    taps, cin, cout = w_kio.shape
    assert sig.k == cin
    if tw is None:
        tw = TcWeight(w_kio.permute(2, 0, 1).reshape(cout, taps * cin).contiguous(), [cin] * taps if taps > 1 else None)
    M = sig.m
    rows_out = out_rows if out_map is not None else M
    Y = _empty((rows_out, cout), torch.float32, sig.buf) if out != "planes" else None
    Yp = alloc_planes(rows_out, cout, sig.n_planes, sig.buf) if out != "f32" else None
    box = tw.pick_box_planes(M, sig.n_planes)
    w_hi, w_lo, _ = tw.maps(box)
    call("ccn_conv_planes", sig.buf, sig.kpad, sig.buf.stride(0), cin, taps, w_hi, w_lo, bias if bias is not None else 0,
         Y if Y is not None else 0, cout, Yp.buf if Yp is not None else 0, Yp.kpad if Yp is not None else 0,
         Yp.buf.stride(0) if Yp is not None else 0, M, cout, box, 3 if sig.n_planes == 2 else 1, act, float(slope),
         out_map if out_map is not None else 0)
    return Y if out == "f32" else Yp if out == "planes" else (Y, Yp)


_TC_SCRATCH = {}


def tc_weight_scratch(w: torch.Tensor) -> TcWeight:
    """TcWeight of a weight that changes every step (training): ONE set of bf16 buffers + TMA descriptors per weight
    shape, refilled by ccn_split_bf16.  Safe because every user launches on the same stream right after the refill
    (stream order: split -> GEMM -> next split)."""
    w, ldw = _rows(w)
    key = (w.shape[0], w.shape[1], w.device)
    tw = _TC_SCRATCH.get(key)
    if tw is None:
        tw = _TC_SCRATCH[key] = TcWeight(w)
    else:
        call("ccn_split_bf16", w, ldw, tw.n, tw.k, tw.kpad, tw.hi, tw.lo)
    return tw


def tc_weight_param(w: torch.Tensor) -> TcWeight:
    """TcWeight of a PARAMETER in training: bf16 planes + TMA descriptors kept on the parameter object and refilled only
    when its in-place version changed (once per optimiser step) - the forward GEMM and the data gradient of the same step
    share one split.  Anything that is not a leaf parameter (slices, concatenations) goes through tc_weight_scratch."""
    if not (w.is_leaf and isinstance(w, torch.nn.Parameter) and w.is_contiguous()):
        return tc_weight_scratch(w.detach())
    hit = w.__dict__.get("_ccn_tw")
    if hit is not None and hit[0] == w._version and hit[2] == w.data_ptr():
        return hit[1]
    wd = w.detach()
    if hit is not None and hit[1].n == wd.shape[0] and hit[1].k == wd.shape[1] and hit[1].hi.device == wd.device:
        tw = hit[1]
        call("ccn_split_bf16", wd, wd.stride(0), tw.n, tw.k, tw.kpad, tw.hi, tw.lo)
    else:
        tw = TcWeight(wd)
    w.__dict__["_ccn_tw"] = (w._version, tw, w.data_ptr())
    return tw


def prepare_tc_weight(w: torch.Tensor, seg_widths=None) -> TcWeight:
    return TcWeight(w, seg_widths)


def linear_tc(A: torch.Tensor, tw: TcWeight, bias: Optional[torch.Tensor], act: int, slope: float, mode: int):
    """Y = act(A W^T + bias) on tcgen05; mode 1 = bf16, 3 = split-bf16 x3."""
    A, lda = _rows(A)
    M, K = A.shape
    assert K == tw.k, (A.shape, tw.k)
    Y = _empty((M, tw.n), torch.float32, A)
    box = tw.pick_box(M)
    p_hi, p_lo, _ = tw.maps(box)
    call("ccn_linear_tc", A, lda, p_hi, p_lo, bias if bias is not None else 0, Y, tw.n, M, tw.n, K, box, mode, act,
         float(slope))
    return Y


# ----------------------------------------------------------------------------------------------
# fused edge-MLP steps (csrc/edge_fused.cu): gather -> layer 1 (decomposed) -> layers 2..n -> aggregate in one kernel
# ----------------------------------------------------------------------------------------------
CHAIN_MAX_LAYERS = 6
CHAIN_MAX_WIDTH = 256
STREAM_MAX_WIDTH = 1024


class ChainWeights:
    """The layers AFTER the decomposed first one of an edge MLP (plus, for the attention aggregations, attend_nn's
    layers), prepared for ccn_sgcnn_fused / ccn_pointconv_fused: every layer's folded fp32 weight (N_l, K_l) stacked
    row-wise into ONE split-bf16 matrix (columns zero padded to the widest K), one TMA descriptor per plane
    (64 x 64 boxes) and the host-side ccn_chain_t.  Weight preparation, once per weight version."""

    def __init__(self, layers, n_planes: int, first_seg_widths=None):
        """layers: [(w (N, K) fp32 folded, bias (N) or None, act, slope)].  first_seg_widths (ccn_mlp_chain): the first
        layer's input is a concatenation of blocks of these widths; every block's weight columns start at a multiple of 64."""
        import ctypes
        assert 1 <= len(layers) <= CHAIN_MAX_LAYERS
        if first_seg_widths is not None:
            w0 = layers[0][0]
            assert sum(first_seg_widths) == w0.shape[1]
            tot = sum((kw + 63) // 64 * 64 for kw in first_seg_widths)
            wide0 = torch.zeros((w0.shape[0], tot), dtype=torch.float32, device=w0.device)
            c0 = off = 0
            for kw in first_seg_widths:
                wide0[:, off:off + kw] = w0[:, c0:c0 + kw]
                c0 += kw
                off += (kw + 63) // 64 * 64
            layers = [(wide0,) + tuple(layers[0][1:])] + list(layers[1:])
        kmax = max(w.shape[1] for w, _, _, _ in layers)
        rows = sum(w.shape[0] for w, _, _, _ in layers)
        dev = layers[0][0].device
        wide = torch.zeros((rows, kmax), dtype=torch.float32, device=dev)
        offs, r = [], 0
        for w, _, _, _ in layers:
            wide[r:r + w.shape[0], :w.shape[1]] = w
            offs.append(r)
            r += w.shape[0]
        self.tw = TcWeight(wide)
        self.n_out = layers[-1][0].shape[0]
        self.n_planes = n_planes
        self._keep = [b for _, b, _, _ in layers]
        w_hi, w_lo, self._maps_keep = self.tw.maps(64)

        L = CHAIN_MAX_LAYERS

        class _Chain(ctypes.Structure):
            _fields_ = [("n_layers", ctypes.c_int32), ("mode", ctypes.c_int32), ("k", ctypes.c_int32 * L),
                        ("n", ctypes.c_int32 * L), ("w_row", ctypes.c_int32 * L), ("act", ctypes.c_int32 * L),
                        ("slope", ctypes.c_float * L), ("bias", ctypes.c_void_p * L), ("w_hi_tmap", ctypes.c_void_p),
                        ("w_lo_tmap", ctypes.c_void_p)]

        c = _Chain()
        c.n_layers, c.mode = len(layers), 3 if n_planes == 2 else 1
        for i, (w, b, act, slope) in enumerate(layers):
            c.k[i], c.n[i], c.w_row[i], c.act[i], c.slope[i] = w.shape[1], w.shape[0], offs[i], act, slope
            c.bias[i] = b.data_ptr() if b is not None else None
        c.w_hi_tmap, c.w_lo_tmap = w_hi, w_lo
        self._struct = c
        self.addr = ctypes.addressof(c)

    @staticmethod
    def supported(c1: int, dims, dense: bool = False) -> bool:
        """dims: output widths of the chain layers.  Every width a multiple of 64 and <= 256 (the activations of a tile stay
        resident in shared memory); a dense SGCNN step with ONE chain layer streams its operands instead and takes widths up
        to 1024 (sgcnn_stream_kernel)."""
        wmax = STREAM_MAX_WIDTH if dense and len(dims) == 1 else CHAIN_MAX_WIDTH
        return (len(dims) >= 1 and len(dims) <= CHAIN_MAX_LAYERS and c1 % 64 == 0 and 64 <= c1 <= wmax and
                all(d % 64 == 0 and 64 <= d <= wmax for d in dims))


def fused_profitable(c1: int, n_planes: int, attention: bool, dense_single_layer: bool = False) -> bool:
    """Measured on B200 (tools/bench_fused.py, profiles/r02g_fused.txt): the one-kernel edge step wins for the NARROW chains
    (resident activation tile <= 32 KB: two CTAs per SM overlap each other's latency chains) - the 374 k-row level-1 SGCNN
    steps, 96 us against 116 us in bf16x3 - and loses to the unfused launches on the 128 / 256-wide chains (one CTA per SM,
    the serial form -> MMA -> epilogue chain of a tile is not hidden) and on the attention steps (1.5x slower)."""
    if FUSED_EDGE_STEPS != "auto":
        return bool(FUSED_EDGE_STEPS)
    if dense_single_layer:
        # sgcnn_stream_kernel (one chain layer, widths up to 1024): 76 / 90 / 101 / 130 us against 58 / 48 / 50 / 46 us of the
        # unfused launches at the level-2..5 shapes (profiles/r02l_fused_bf16x3.txt) - forming the A operand inside the GEMM
        # is bound by the gather rate of one SM's few producer warps (per-thread loads: 74 us; TMA row gathers of 256 bytes:
        # 76 us, the copy engine serialises small copies), where the stand-alone layer-1 kernel gathers at full occupancy
        return False
    return (not attention) and c1 * n_planes <= 128


def sgcnn_fused(PQ: torch.Tensor, idx: torch.Tensor, act0: int, slope0: float, chain: ChainWeights, npl: int = 0,
                cache=None) -> torch.Tensor:
    """Dense SGCNN step in one kernel (ccn_sgcnn_fused): max_k chain(act0(P[idx[i][k]] + Q[i])), self loop included,
    missing neighbours at -100.  -> fp32 (n, C_out); its activation planes are registered in `cache`."""
    PQ, ld = _rows(PQ)
    n, K = idx.shape
    c1 = PQ.shape[1] // 2
    out = _empty((n, chain.n_out), torch.float32, PQ)
    pl = alloc_planes(n, chain.n_out, npl, PQ) if npl and cache is not None and n else None
    call("ccn_sgcnn_fused", PQ, ld, idx, n, K, c1, act0, float(slope0), chain.addr, out, chain.n_out,
         pl.buf if pl is not None else 0, pl.kpad if pl is not None else 0, pl.plane_stride if pl is not None else 0,
         npl if pl is not None else 0)
    if pl is not None:
        cache.put_planes(out, pl)
    return out


FUSED_MAX_DEGREE = 64       # CSR tiles hold the targets whose first edge falls into a stride of 129 - max_degree rows


def pointconv_fused(edges: EdgeCSR, chain: ChainWeights, agg: str, msg_layer: int, act0: int, slope0: float, c1: int,
                    H: Optional[torch.Tensor] = None, wp: Optional[torch.Tensor] = None, bias0: Optional[torch.Tensor] = None,
                    pos_src: Optional[torch.Tensor] = None, pos_dst: Optional[torch.Tensor] = None,
                    radius: Optional[float] = None, PQ: Optional[torch.Tensor] = None, npl: int = 0, cache=None):
    """Edge-MLP step over a CSR edge list in one kernel (ccn_pointconv_fused).  First layer either
    act0(H[j] + Wp (pos_src[j] - pos_dst[i]) [/ radius] + bias0) (PointNetConv2.message) or act0(PQ[j][:c1] + PQ[i][c1:])
    (sparse StaticEdgeConv); then the chain; then `agg` over every target's edges.  -> fp32 (nq, C_out)."""
    assert edges.max_degree <= FUSED_MAX_DEGREE
    like = PQ if PQ is not None else wp
    out = _empty((edges.nq, chain.n_out), torch.float32, like)
    pl = alloc_planes(edges.nq, chain.n_out, npl, like) if npl and cache is not None and edges.nq else None
    n_tiles = edges.n_edges // (129 - edges.max_degree) + 1
    scratch = _empty((4 * n_tiles,), torch.int32, like)
    if PQ is not None:
        PQ, ldpq = _rows(PQ)
        a = (0, 0, 0, 0, 0, 0, PQ, ldpq)
    else:
        if H is not None:
            H, ldh = _rows(H)
        else:
            ldh = 0
        a = (H if H is not None else 0, ldh, wp, bias0 if bias0 is not None else 0, _f32(pos_src), _f32(pos_dst), 0, 0)
    call("ccn_pointconv_fused", *a, edges.row_ptr, edges.erow, edges.ecol, edges.nq, edges.n_edges, edges.max_degree, c1,
         float(radius or 0.0), act0, float(slope0), chain.addr, AGG[agg], msg_layer, scratch, out, chain.n_out,
         pl.buf if pl is not None else 0, pl.kpad if pl is not None else 0, pl.plane_stride if pl is not None else 0,
         npl if pl is not None else 0)
    if pl is not None:
        cache.put_planes(out, pl)
    return out


# Per-point MLP chains (ccn_mlp_chain): 2+ consecutive Linear layers of a shared MLP in one launch.  "auto": where measured
# faster than the per-layer launches (mlp_chain_profitable); True / False force it (bit-identical results either way).
CHAIN_MLP_LAYERS = "auto"


def mlp_chain_supported(seg_widths, dims) -> bool:
    """seg_widths: K-blocks of the chain's input; dims: output widths of its layers."""
    k0 = sum((kw + 63) // 64 * 64 for kw in seg_widths)
    return (2 <= len(dims) <= CHAIN_MAX_LAYERS and 1 <= len(seg_widths) <= 4 and 64 <= k0 <= CHAIN_MAX_WIDTH and
            all(d % 64 == 0 and 64 <= d <= CHAIN_MAX_WIDTH for d in dims[:-1]) and 1 <= dims[-1] <= CHAIN_MAX_WIDTH)


def mlp_chain_profitable(seg_widths, dims) -> bool:
    """Measured on B200 (tools/bench_fused.py `mlp` lines, profiles/r02p_mlp_chain.txt; bf16x3, launch to launch with a
    flushed L2): one launch beats the per-layer launches for the 3+-layer chains over a single input block - the out MLP
    [64] -> [64, 64, 17] at 32.5 k rows 21 us against 32 us, [256] -> [128, 128, 64] at 17.8 k rows 28 against 33 us - and for
    the narrow two-layer chains ([64, 64] -> [64, 64]: 17 against 19 us).  It loses (0.8x) where a two-layer chain is 128+
    wide or its input blocks need padding to the 64-column chunks (one CTA per SM, a tile's serial load -> MMA -> epilogue
    chain is not hidden, and the per-layer kernels already run at the launch floor)."""
    if CHAIN_MLP_LAYERS != "auto":
        return bool(CHAIN_MLP_LAYERS)
    if len(dims) >= 3 and len(seg_widths) == 1:
        return True
    k0 = sum(seg_widths)
    return all(kw % 64 == 0 for kw in seg_widths) and k0 <= 128 and max(dims) <= 64


def mlp_chain_plan(layers, seg_widths):
    """Which suffix of an MLP runs as one ccn_mlp_chain launch: -> (start layer, K-blocks of its input) or None.
    layers: [PreparedLinear]; seg_widths: K-blocks of the MLP's input.  Either the whole MLP, or everything after its
    first layer (whose input is too wide / too ragged for a resident tile)."""
    if CHAIN_MLP_LAYERS is False:
        return None
    dims = [l.out_features for l in layers]
    seg_widths = tuple(int(kw) for kw in seg_widths)
    if mlp_chain_supported(seg_widths, dims) and mlp_chain_profitable(seg_widths, dims):
        return 0, seg_widths
    if len(dims) >= 3 and mlp_chain_supported((dims[0],), dims[1:]) and mlp_chain_profitable((dims[0],), dims[1:]):
        return 1, (dims[0],)
    return None


def mlp_chain(A, chain: ChainWeights, out: str = "f32"):
    """A: Planes or PlanesCat (<= 4 blocks); -> fp32 (M, N) | Planes | (fp32, Planes) like linear_planes."""
    segs = A.segs if isinstance(A, PlanesCat) else (A,)
    M = A.shape[0]
    npl = segs[0].n_planes
    dev = segs[0].buf
    Y = _empty((M, chain.n_out), torch.float32, dev) if out != "planes" else None
    Yp = alloc_planes(M, chain.n_out, npl, dev) if out != "f32" else None
    a = []
    for sg in segs:
        a += [sg.buf, sg.kpad, sg.buf.stride(0), sg.k]
    a += [0, 0, 0, 0] * (4 - len(segs))
    call("ccn_mlp_chain", *a, M, chain.addr, Y if Y is not None else 0, chain.n_out, Yp.buf if Yp is not None else 0,
         Yp.kpad if Yp is not None else 0, Yp.buf.stride(0) if Yp is not None else 0)
    return Y if out == "f32" else Yp if out == "planes" else (Y, Yp)
