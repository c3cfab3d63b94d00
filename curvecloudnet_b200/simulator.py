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


class ScanLineGenerator:
    """scanning_simulator/utils/scanlines.py:4-150.  The line end points and the sub-sample offset are drawn on the host from
    numpy's / torch's global generators in the reference's order (seed both identically -> the same lines); everything per
    sample (pixel, first-on-pixel test, mask look-up, the N-point cut, line-id compression) runs on the device."""
    UV_BASIS = np.array([[1, 0], [0, 1]])
    NLINES = 100
    MAX_WHILE_ATTEMPTS = 400

    def __init__(self, resolution, device, style="linear", direction="random", line_density=1.0):
        self.style = style
        self.direction = direction
        self.resolution = resolution
        self.line_density = line_density
        self.device = device
        self.borders = np.array([[0, 0], [0, 0], [0, resolution - 1], [resolution - 1, 0]])
        self.ts = int(2 ** 0.5 * self.resolution * self.line_density)

    def generate_scan_lines(self, B, N, masks, coords=None, normals=None):
        """-> (success, uv (B, N, 3) int64 = (u, v, line id per mesh)).  With coords / normals (B, res, res, 3) the pixel
        look-up of SampleMeshes._sample_curves is done in the same pass: -> (success, uv, points, point_normals)."""
        if self.style == "linear":
            return self.generate_linear_scan_lines(B, N, masks, coords, normals)
        raise NotImplementedError("Have not implemented sine scanning")     # scanlines.py:21-22

    def generate_linear_scan_lines(self, B, N, masks, coords=None, normals=None):
        if not masks.is_cuda:
            raise native.CcnError("ScanLineGenerator: CUDA tensors only (no CPU path exists)")
        dev = masks.device
        res = self.resolution
        m8 = masks.contiguous().view(torch.uint8) if masks.dtype == torch.bool else masks.to(torch.uint8).contiguous()
        uv = torch.zeros((B, N, 3), dtype=torch.int64, device=dev)
        npoints = torch.zeros((B,), dtype=torch.int32, device=dev)
        host = torch.zeros((B,), dtype=torch.int32).pin_memory()
        num_attempts = 0
        while True:
            host.copy_(npoints)                           # the reference's `while torch.any(npoints_per_scan < N).item()`
            if not bool((host < N).any()):
                break
            if self.direction == "random":
                border = self._sample_linear_border_points(self.NLINES, B)
            elif self.direction == "parallel":
                border = self._sample_parallel_border_points(num_attempts, self.NLINES, B)
            elif self.direction == "grid":
                border = self._sample_grid_border_points(num_attempts, self.NLINES, B)
            else:
                raise RuntimeError("Sampling direction must be parallel, grid, or random!")
            f32_math = int(np.issubdtype(border.dtype, np.integer))
            offset = float((torch.rand(1) / self.ts).item())                  # sample_line_points' rand_offset (float32)
            bdev = torch.from_numpy(np.ascontiguousarray(border, dtype=np.float64).reshape(B, self.NLINES, 2, 2)).to(dev)
            call("ccn_sim_scan_lines", bdev, B, self.NLINES, self.ts, offset, f32_math, res, m8, N, num_attempts * self.NLINES,
                 npoints, uv)
            num_attempts += 1
            if num_attempts > self.MAX_WHILE_ATTEMPTS:
                print("Reached maximum line-sampling attempts of %s" % self.MAX_WHILE_ATTEMPTS)
                print("Total number of possible points: %s" % torch.sum(masks.view(B, -1), dim=1))
                return (False, None) if coords is None else (False, None, None, None)
        if coords is None:
            call("ccn_sim_scan_finish", uv, B, N, res, 0, 0, 0, 0)
            return True, uv
        pts = torch.empty((B, N, 3), dtype=torch.float32, device=dev)
        nrm = torch.empty((B, N, 3), dtype=torch.float32, device=dev)
        call("ccn_sim_scan_finish", uv, B, N, res, coords.contiguous(), normals.contiguous(), pts, nrm)
        return True, uv, pts, nrm

    # ---- host-side draws, statement for statement (scanlines.py:91-131); `np.int` of the reference is plain int
    def _sample_linear_border_points(self, N, B):
        border_idxs = np.array([np.random.choice(np.arange(4), size=2, replace=False) for i in range(N * B)])
        border_directions = self.UV_BASIS[border_idxs.flatten() % 2].reshape(-1, 2, 2)
        border_points = self.borders[border_idxs.flatten()].reshape(-1, 2, 2)
        border_offsets = np.random.randint(0, high=self.resolution, size=border_points.shape)
        border_points += border_directions * border_offsets
        return border_points

    def _binary_locs(self, num_attempts, N):
        num_bins = int(np.ceil(np.log2((num_attempts + 1) * N)))
        N_full = 2 ** num_bins
        idxs = np.arange(num_attempts * N_full, (num_attempts + 1) * N_full) + 1
        bin_nums = ((idxs.reshape(-1, 1) & (2 ** np.arange(num_bins)).astype(int)) != 0).astype(int)
        factors = np.ones(num_bins) * 2 ** (-np.arange(1, num_bins + 1).astype(float))
        locs = np.sum(bin_nums * factors.reshape(1, -1), axis=1)
        locs = locs[num_attempts * N: (num_attempts + 1) * N]
        return locs * self.resolution

    def _sample_parallel_border_points(self, num_attempts, N, B):
        locs = self._binary_locs(num_attempts, N)
        border_pnts = np.stack([locs, np.zeros(N), locs, np.ones(N) * self.resolution - 1], axis=1).reshape((N, 2, 2))
        return np.repeat(border_pnts, B, axis=0)

    def _sample_grid_border_points(self, num_attempts, N, B):
        assert (N % 2) == 0
        N = N // 2
        locs = self._binary_locs(num_attempts, N)
        horizontal = np.stack([locs, np.zeros(N), locs, np.ones(N) * self.resolution - 1], axis=1).reshape((N, 2, 2))
        vertical = np.stack([np.zeros(N), locs, np.ones(N) * self.resolution - 1, locs], axis=1).reshape((N, 2, 2))
        border_pnts = np.stack([horizontal, vertical], axis=1).reshape((N * 2, 2, 2))
        return np.repeat(border_pnts, B, axis=0)


class CurveClouds:
    """scanning_simulator/utils/curve_cloud.py:8-190: polylines -> 3-D curves (split at jumps), anti-aliased points / normals,
    per-point curvature vectors and position gradients.  Attributes as the reference's: points / normals (B, N, 3) float64,
    uv, point2uvcurve_idxs, point2curve_idxs (B, N) float32, curvature (B, N, 3) float64, gradients (B, N, 3, 3) float64.
    `with_intersections` (curve_cloud.py:113-165, an O(N^2) segment-pair search; sampling.py:85 asks for it, nothing downstream
    stores it): intersections (2, K) int64 and intersection2batch_idxs (K,), one curve cloud at a time like upstream."""
    CURVE_SPLIT_THRESH = 0.01
    INTERSECTION_DIST_THRESH = 0.01
    ANTI_ALIAS_KERNEL = [1, 2, 3, 3, 3, 2, 2]

    def __init__(self, points, normals, uv, point2uvcurve_idxs, anti_alias=True, with_intersections=False, line_density=1.0,
                 curvature_knn=10):
        if not points.is_cuda:
            raise native.CcnError("CurveClouds: CUDA tensors only (no CPU path exists)")
        B, N = points.shape[0], points.shape[1]
        dev = points.device
        p32 = points.to(torch.float32).contiguous()
        n32 = normals.to(torch.float32).contiguous()
        self.uv = uv
        self.point2uvcurve_idxs = point2uvcurve_idxs
        self.device = dev
        cidx = torch.empty((B, N), dtype=torch.int32, device=dev)
        self._cstart = torch.empty((B, N), dtype=torch.int32, device=dev)
        self._cend = torch.empty((B, N), dtype=torch.int32, device=dev)
        self._ncurves = torch.empty((B,), dtype=torch.int32, device=dev)
        call("ccn_sim_curve_split", p32, B, N, float(self.CURVE_SPLIT_THRESH), cidx, self._cstart, self._cend, self._ncurves)
        self._curve_idx_i32 = cidx
        self.point2curve_idxs = cidx.to(torch.float32)               # the reference's cat([zeros, cumsum]) is float32
        if anti_alias:
            self.points = torch.empty((B, N, 3), dtype=torch.float64, device=dev)
            self.normals = torch.empty((B, N, 3), dtype=torch.float64, device=dev)
            call("ccn_sim_anti_alias", p32, cidx, B, N, self.points)
            call("ccn_sim_anti_alias", n32, cidx, B, N, self.normals)
        else:
            self.points, self.normals = p32.double(), n32.double()
        self.curvature_knn = curvature_knn
        self.curvature, self.gradients = self.compute_curvature_new()
        self.intersections = None
        self.intersection2batch_idxs = None
        if with_intersections:
            self.intersections, self.intersection2batch_idxs = self.compute_intersections()

    def compute_intersections(self):
        """curve_cloud.py:114-166 -> (pairs (2, K) int64 point indices, batch (K,) int64): for every ordered pair of curves
        whose polylines come closer than INTERSECTION_DIST_THRESH, the point of each at the closest approach.  One curve cloud
        only: upstream's second scatter_min (:146) cannot broadcast its index for B > 1 and raises - so does this."""
        B, N = self.points.shape[0], self.points.shape[1]
        dev = self.points.device
        if B != 1:
            raise RuntimeError("CurveClouds.compute_intersections: one curve cloud at a time (B = %d; the reference's "
                               "scatter_min over the columns fails to broadcast for B > 1)" % B)
        if N < 2:
            return torch.zeros((2, 0), dtype=torch.int64, device=dev), torch.zeros((0,), dtype=torch.int64, device=dev)
        G = int(self._ncurves[0].item())                                   # host sync
        E = N - 1
        pts = self.points[0].contiguous()
        cidx = self._curve_idx_i32[0].contiguous()
        i32 = dict(dtype=torch.int32, device=dev)
        cptr = torch.empty((G + 1,), **i32)
        min1 = torch.empty((G, E), dtype=torch.float64, device=dev)
        arg1 = torch.empty((G, E), **i32)
        arg2 = torch.empty((G * G,), **i32)
        flag = torch.empty((G * G,), **i32)
        slot = torch.empty((G * G + 1,), **i32)
        call("ccn_sim_intersections_count", pts, cidx, N, G, float(self.INTERSECTION_DIST_THRESH), cptr, min1, arg1, arg2, flag, slot)
        K = int(slot[G * G].item())                                        # host sync
        pairs = torch.empty((2, K), dtype=torch.int64, device=dev)
        err = torch.empty((1,), **i32)
        call("ccn_sim_intersections_fill", pts, N, G, arg1, arg2, flag, slot, K, pairs if K else 0, err)
        if K and int(err.item()):
            raise IndexError("CurveClouds.compute_intersections: a closest approach at the far end of the last edge "
                             "(upstream indexes past its distance matrix here)")
        return pairs, torch.zeros((K,), dtype=torch.int64, device=dev)

    def compute_curvature_new(self):
        B, N = self.points.shape[0], self.points.shape[1]
        dev = self.points.device
        curvature = torch.empty((B, N, 3), dtype=torch.float64, device=dev)
        gradients = torch.empty((B, N, 3, 3), dtype=torch.float64, device=dev)
        raw = torch.empty((B, N), dtype=torch.float64, device=dev)
        geo = torch.empty((B, N), dtype=torch.float64, device=dev)
        mean_edge = torch.empty((B,), dtype=torch.float64, device=dev)
        call("ccn_sim_curvature", self.points, self._curve_idx_i32, self._cstart, self._cend, B, N, int(self.curvature_knn), 2.5, 1e-4,
             raw, geo, mean_edge, curvature, gradients)
        return curvature, gradients

    def __len__(self):
        return self.points.size(0)

    def to(self, device):
        for name in ("points", "normals", "uv", "point2uvcurve_idxs", "point2curve_idxs", "curvature", "gradients"):
            setattr(self, name, getattr(self, name).to(device))
        if self.intersections is not None:
            self.intersections = self.intersections.to(device)
            self.intersection2batch_idxs = self.intersection2batch_idxs.to(device)


def transfer_labels(points: torch.Tensor, ref_points: torch.Tensor, ref_labels: torch.Tensor):
    """generate_shapenet.py:88-89 (`closest = argmin(cdist(points.double(), pos.double()), dim=1); y[closest]`): labels of a
    sampled curve cloud (N, 3) from the labelled point cloud (M, 3) / (M,) of the same model -> (labels (N,), closest (N,))."""
    if not points.is_cuda:
        raise native.CcnError("transfer_labels: CUDA tensors only (no CPU path exists)")
    q = points.to(torch.float64).contiguous()
    r = ref_points.to(torch.float64).contiguous()
    lab = ref_labels.to(torch.int64).contiguous()
    idx = torch.empty((q.shape[0],), dtype=torch.int64, device=q.device)
    out = torch.empty((q.shape[0],), dtype=torch.int64, device=q.device)
    call("ccn_sim_nearest_label", q, q.shape[0], r, r.shape[0], lab, idx, out)
    return out.to(ref_labels.dtype), idx


class SampleMeshes:
    """scanning_simulator/utils/sampling.py:13-85: batches of meshes -> CurveClouds of `npoints` points each."""

    def __init__(self, npoints, device, resolution=1024, camera_view=(1.0, 0, 0, 0), scan_style="linear", scan_direction="random",
                 line_density=1.0, curvature_knn=10):
        self.npoints = npoints
        self.device = device
        self.resolution = resolution
        self.camera_view = camera_view
        self.scan_style = scan_style
        self.scan_direction = scan_direction
        self.line_density = line_density
        self.curvature_knn = curvature_knn
        R_x = torch.tensor(rodrigues([camera_view[1], 0, 0]), dtype=torch.float32)
        R_y = torch.tensor(rodrigues([0, camera_view[2], 0]), dtype=torch.float32)
        R_z = torch.tensor(rodrigues([0, 0, camera_view[3]]), dtype=torch.float32)
        self.R = (R_x @ R_y @ R_z).to(device)
        self.t = torch.tensor([[0, 0, camera_view[0]]], dtype=torch.float32, device=device)
        self.scan_lines_gen = ScanLineGenerator(resolution, device, scan_style, scan_direction, line_density)

    def meshes2points(self, vertices, faces, labels=None, viz_outdir=None):
        coords, normals, masks = self._render_meshes(vertices, faces, viz_outdir)
        return self._sample_curves(coords, normals, masks)

    def _render_meshes(self, vertices, faces, viz_outdir=None):
        out = [rasterize(vertices[i], faces[i], self.R, self.t, res=self.resolution) for i in range(len(vertices))]
        return torch.stack([o[0] for o in out]), torch.stack([o[1] for o in out]), torch.stack([o[2] for o in out])

    def _sample_curves(self, coords, normals, masks):
        B = coords.size(0)
        success, scans_uv, pts, nrm = self.scan_lines_gen.generate_scan_lines(B, self.npoints, masks, coords, normals)
        if not success:
            return None
        # the reference passes with_intersections=True (sampling.py:83-84), which only works for ONE mesh (every shipped caller's
        # case, generate_shapenet.py:69,83); for several meshes - where upstream raises - the intersections are left out
        return CurveClouds(pts, nrm, scans_uv[:, :, :2], scans_uv[:, :, 2], anti_alias=True, with_intersections=(B == 1),
                           curvature_knn=self.curvature_knn)
