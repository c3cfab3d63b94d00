"""Generate tests/golden/*.pt by running the REAL reference python sources.

TEST INFRASTRUCTURE, build container only (needs /root/reference; the GPU box does not have
it - the committed fixtures travel instead).  Usage:  python -m oracle.make_golden

How the reference is made importable here: its third-party imports (pytorch3d,
torch_geometric, torch_scatter, torch_sparse, torch_cluster, frnn) resolve to oracle/shims/*
(the oracle's restatement of that arithmetic).  Two documented deviations from stock:
  * `point_ops.fast_knn` raises on CPU tensors (point_ops.py:439-440); it is replaced by the
    same call without the `is_cuda` check so the nuScenes/KITTI/A2D2 step lists run on CPU;
  * `SemNuScenes._get_curves` is compiled from its own source text (ast) because importing
    src/data/nuscenes_dataset.py needs nuscenes-devkit/cv2/mitsuba.
So the fixtures pin the reference's OWN code (base.py, modules/*, point_ops.py, fps_ops.py,
_get_curves); the third-party arithmetic under the shims stays "parity unpinned".
"""
from __future__ import annotations

import ast
import os
import sys
import textwrap

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
OUT = os.path.join(ROOT, "tests", "golden")


def _import_reference():
    for p in (ROOT, os.path.join(ROOT, "oracle", "shims"), REF):
        if p not in sys.path:
            sys.path.insert(0, p)
    import frnn
    import src.models.utils.point_ops as rpo

    def fast_knn_no_cuda_check(points1, points2, lengths1, lengths2, K, r, return_nn=False):
        return frnn.frnn_grid_points(points1.contiguous(), points2.contiguous(), lengths1, lengths2, K, r)[1]

    rpo.fast_knn = fast_knn_no_cuda_check
    import src.models.base as rbase
    import src.models.modules.fast_conv1d as rconv
    import src.models.modules.fps_ops as rfps
    return rpo, rbase, rconv, rfps


def _reference_get_curves():
    """Compile SemNuScenes._get_curves from its source text (nuscenes_dataset.py:91-118)."""
    src = open(os.path.join(REF, "src/data/nuscenes_dataset.py")).read()
    tree = ast.parse(src)
    for node in ast.walk(tree):
        if isinstance(node, ast.FunctionDef) and node.name == "_get_curves":
            code = textwrap.dedent(ast.get_source_segment(src, node))
            ns = {"torch": torch}
            exec(compile(code, "nuscenes_dataset.py:_get_curves", "exec"), ns)
            return ns["_get_curves"]
    raise RuntimeError("_get_curves not found")


def _cases():
    from curvecloudnet_b200.synthetic import driving_scan_batch, object_scan_batch
    return {
        "driving_b2": driving_scan_batch([11, 12], subsample=1500),
        "driving_b1": driving_scan_batch([13], subsample=2500),
        "object_b2": object_scan_batch([21, 22], n_points=768),
    }


def make_ops(rpo, rconv, rfps):
    out = {}
    for cname, data in _cases().items():
        pos, batch, curve = data.pos, data.batch, data.curve_idxs
        g = {"pos": pos, "x": data.x, "batch": batch, "curve": curve}
        driving = cname.startswith("driving")
        arclen = 0.007 if driving else 0.04
        r_geo = 0.02 if driving else 0.075
        g["glob"] = rpo.curveidx_local2global(curve.clone(), batch)
        g["ptr"] = rpo.batch2ptr(batch)
        g["ptr_ends"] = rpo.batch2ptr(batch, with_ends=True)
        padded, mask = rpo.to_batch_padded(pos.clone(), batch.clone())
        g["padded_sum"], g["padded_mask"] = padded.sum(dim=1), mask

        torch.manual_seed(100)
        g["curvefps_u"] = torch.rand(1)
        torch.manual_seed(100)
        idx = rfps.CurveFPS(arclen)(pos.clone(), batch.clone(), curve.clone())
        g["curvefps_arclen"], g["curvefps_idx"] = arclen, idx

        vox = 0.03 if driving else 0.08
        torch.manual_seed(101)
        g["voxel_noise"] = torch.rand(pos.shape[0])
        torch.manual_seed(101)
        g["voxel_size"], g["voxel_idx"] = vox, rfps.VoxelFPS(vox)(pos.clone(), batch.clone())

        torch.manual_seed(102)
        lengths = torch.bincount(batch)
        g["fps_starts"] = torch.tensor([int(torch.randint(high=int(l), size=(1,)).item()) for l in lengths])
        torch.manual_seed(102)
        g["fps_ratio"], g["fps_idx"] = 0.25, rpo.fps_pytorch3d(pos, batch, 0.25)

        row, col = rpo.radius_1d_group_subset(pos, idx, curve, batch, r_geo)
        g["r1d_radius"], g["r1d_row"], g["r1d_col"] = r_geo, row, col
        row, col = rpo.knn_1d_group_superset(pos, idx, curve, batch, 3)
        g["s1d_row"], g["s1d_col"] = row, col

        r_knn = 0.04 if driving else 0.2
        g["knn_radius"] = r_knn
        sub = g["voxel_idx"]
        dense, l2, m1 = rpo.knn_ball_group_pytorch3d(pos[sub].clone(), pos, batch[sub].clone(), batch, operation="knn",
                                                     knn=16, radius=r_knn, return_dense=True)
        g["frnn_dense"] = dense
        row, col = rpo.knn_ball_group_pytorch3d(pos[sub].clone(), pos, batch[sub].clone(), batch, operation="knn",
                                                knn=16, radius=r_knn)
        g["frnn_row"], g["frnn_col"] = row, col
        row, col = rpo.knn_ball_group_pytorch3d(pos[sub].clone(), pos, batch[sub].clone(), batch,
                                                operation="ball-group", radius=r_knn)
        g["ball_row"], g["ball_col"] = row, col
        row, col = rpo.knn_ball_group_pytorch3d(pos, pos[sub].clone(), batch, batch[sub].clone(), operation="knn",
                                                knn=3, accel_knn=False)
        g["knn3_row"], g["knn3_col"] = row, col

        gen = torch.Generator().manual_seed(7)
        feat = torch.randn(pos.shape[0], 6, generator=gen)
        g["feat"] = feat
        g["feat_diffs"] = rconv.compute_feature_diffs(feat, curve, batch)
        fsub = torch.randn(sub.shape[0], 8, generator=gen)
        g["feat_sub"] = fsub
        g["interp3"] = rpo.knn_interpolate_pytorch3d(fsub, pos[sub], pos, batch[sub], batch, k=3)
        fidx = torch.randn(idx.shape[0], 8, generator=gen)
        g["feat_idx"] = fidx
        g["interp1d"] = rpo.knn_interpolate_1D_pytorch3d(fidx, idx, pos, batch, curve, 3)

        from curvecloudnet_b200.synthetic import randomize_model_
        for ver, cls, k in (("v2", rconv.SymmetricCurve1DConvV2, 5), ("v1", rconv.SymmetricCurve1DConvFastV1, 7)):
            torch.manual_seed(3)
            mod = cls([9, 16, 12], k, with_xyz=True, with_diff=True)
            randomize_model_(mod, 55)
            mod.eval()
            with torch.no_grad():
                g["conv_%s_out" % ver] = mod(feat, pos, batch, curve)[0]
            g["conv_%s_k" % ver] = k
        out[cname] = g

    # a3: curve grouping of raw sweeps
    from curvecloudnet_b200.synthetic import raycast_driving_scan
    gc = _reference_get_curves()
    raw = raycast_driving_scan(31)
    n = 6000
    pts = torch.from_numpy(raw.points[:n])
    beam = torch.from_numpy(raw.beam_ids[:n])

    class _Self:
        CURVE_THRESH = 0.08
    labels = torch.zeros(n, 1, dtype=torch.uint8)
    refl = torch.from_numpy(raw.reflectance[:n])[:, None]
    p2, cid, _, _, inv = gc(_Self(), pts.clone(), beam.clone(), labels, refl)
    out["get_curves"] = {"points": pts, "beam": beam, "sorted_points": p2, "curve_idxs": cid, "inv_reorder": inv}
    return out


def make_models(rbase):
    from curvecloudnet_b200.configs import build_model
    from curvecloudnet_b200.synthetic import driving_scan_batch, object_scan_batch, randomize_model_
    out = {}
    specs = {
        "nuscenes": lambda: driving_scan_batch([41, 42], subsample=1500),
        "shapenet-class": lambda: object_scan_batch([43, 44], n_points=768),
        "shapenet-seg": lambda: object_scan_batch([45, 46], n_points=768),
    }
    for name, mk in specs.items():
        data = mk()
        if name == "shapenet-class":
            del data.labels          # shapenet_classification.py:24-25 deletes it before the call
        model = build_model(rbase.ModelBase, name)
        randomize_model_(model, 1234)
        model.eval()
        torch.manual_seed(77)
        with torch.no_grad():
            logits = model(data, labels=data.y)
        out[name] = {"data_seeds": {"nuscenes": [41, 42], "shapenet-class": [43, 44], "shapenet-seg": [45, 46]}[name],
                     "pos": data.pos, "x": data.x, "batch": data.batch, "curve": data.curve_idxs,
                     "labels": getattr(data, "labels", None), "weight_seed": 1234, "rng_seed": 77,
                     "logits": logits}
        print(name, tuple(logits.shape), float(logits.abs().max()))
    return out


def grad_projection(name: str, numel: int) -> torch.Tensor:
    """Fixed +-1 vector per parameter name (seeded by a hash of the name): gradients are frozen as their projection
    on it, their L2 norm and - for vectors - in full, which keeps train.pt small."""
    import zlib
    gen = torch.Generator().manual_seed(zlib.crc32(name.encode()))
    return torch.randint(0, 2, (numel,), generator=gen).double() * 2 - 1


def make_train(rbase):
    """One training step of the REAL reference sources (train mode, NLL loss of random targets, loss.backward()):
    with BatchNorm on running statistics (well conditioned: tight comparison) and on batch statistics."""
    import torch.nn.functional as F
    from curvecloudnet_b200.configs import build_model
    from curvecloudnet_b200.synthetic import driving_scan_batch, object_scan_batch, randomize_model_
    out = {}
    specs = {
        "nuscenes": lambda: driving_scan_batch([51, 52], subsample=1500),
        "shapenet-seg": lambda: object_scan_batch([55, 56], n_points=768),
    }
    for name, mk in specs.items():
        data = mk()
        for freeze in (True, False):
            model = build_model(rbase.ModelBase, name)
            randomize_model_(model, 1234)
            model.train()
            if freeze:
                for m in model.modules():
                    if isinstance(m, torch.nn.BatchNorm1d):
                        m.eval()
            torch.manual_seed(78)
            logits = model(data, labels=data.y)
            gen = torch.Generator().manual_seed(6)
            target = torch.randint(0, logits.shape[1], (logits.shape[0],), generator=gen)
            loss = F.nll_loss(F.log_softmax(logits, 1), target)
            loss.backward()
            grads = {}
            for k, p in model.named_parameters():
                if p.grad is None:
                    continue
                gd = p.grad.double().flatten()
                ent = {"norm": float(gd.norm()), "proj": float((gd * grad_projection(k, gd.numel())).sum()),
                       "absmax": float(gd.abs().max())}
                if p.dim() == 1:
                    ent["full"] = p.grad.clone()
                grads[k] = ent
            rec = {"logits": logits.detach(), "loss": float(loss), "grads": grads}
            if not freeze:
                rec["buffers"] = {k: v.clone() for k, v in model.named_buffers() if v.dtype.is_floating_point and v.numel() <= 64}
            out.setdefault(name, {"pos": data.pos, "x": data.x, "batch": data.batch, "curve": data.curve_idxs,
                                  "labels": getattr(data, "labels", None), "weight_seed": 1234, "rng_seed": 78,
                                  "target": target})["frozen_bn" if freeze else "batch_stats"] = rec
            print(name, "frozen" if freeze else "batch-stats", float(loss), len(grads))
    return out


def make_losses():
    """seg_loss (NLL + Lovasz-softmax) of the REAL reference sources on random logits: loss, per-point loss and the
    gradient w.r.t. the logits.  lovasz_losses.py is pure torch and is imported as is; seg_loss / fast_hist are
    taken from the reference files by source extraction (their modules import packages that are absent here)."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("ref_lovasz", os.path.join(REF, "src/models/utils/lovasz_losses.py"))
    lov = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(lov)
    import numpy as np
    import torch.nn.functional as F
    src = open(os.path.join(REF, "src/run/nuscenes_seg.py")).read()
    tree = ast.parse(src)
    ns = {"F": F, "torch": torch, "np": np, "lovasz_softmax_flat": lov.lovasz_softmax_flat}
    usrc = open(os.path.join(REF, "src/utils/utils.py")).read()
    exec(compile(ast.parse(usrc), "utils.py", "exec"), ns)
    for node in tree.body:
        if isinstance(node, ast.FunctionDef) and node.name in ("seg_loss", "fast_hist_crop"):
            exec(compile(ast.Module([node], []), "nuscenes_seg.py", "exec"), ns)
    out = {}
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from golden_utils import grad_projection, loss_case
    for name, n, C, seed in (("small", 700, 17, 1), ("nuscenes", 34000, 17, 2), ("kitti", 9000, 20, 3)):
        logits, gt, w = loss_case(name, n, C, seed)                    # inputs are regenerated from the seed by the tests
        logits.requires_grad_()
        rec = {"n": n, "C": C, "seed": seed}
        for lov_on in (False, True):
            logits.grad = None
            loss, per = ns["seg_loss"](logits, gt, ignore=0, use_lovasz=lov_on, ce_weights=w)
            loss.backward()
            gd = logits.grad.double().flatten()
            ent = {"loss": float(loss), "perpnt_head": per.detach()[:256].clone(), "perpnt_sum": float(per.double().sum()),
                   "dnorm": float(gd.norm()), "dproj": float((gd * grad_projection(name, gd.numel())).sum()),
                   "dhead": logits.grad[:256].clone()}
            if name == "small":
                ent["dlogits"] = logits.grad.clone()
            rec["lovasz" if lov_on else "nll"] = ent
        pred = torch.argmax(logits.detach(), dim=-1)
        ul = np.arange(C - 1)
        rec["hist_crop"] = torch.from_numpy(ns["fast_hist_crop"](pred.numpy(), gt.numpy(), ul))
        rec["iou"] = torch.from_numpy(ns["per_class_iu"](rec["hist_crop"].numpy()))
        out[name] = rec
        print(name, rec["nll"]["loss"], rec["lovasz"]["loss"])
    return out


def make_models_extra(rbase):
    """Eval-mode logits of the REAL reference ModelBase for the remaining shipped step lists: A2D2 (sparse SGCNN with
    softmax attention), Kortx (k=7 convs, knn 30, seg categories) and KITTI (nuScenes step list, voxel 0.025, 20 classes)."""
    from curvecloudnet_b200.configs import build_model
    from curvecloudnet_b200.synthetic import driving_scan_batch, object_scan_batch, randomize_model_
    out = {}
    specs = {
        "a2d2": lambda: driving_scan_batch([61, 62], subsample=1500, n_classes=38),
        "kortx": lambda: object_scan_batch([63, 64], n_points=768),
        "kitti": lambda: driving_scan_batch([65], kind="kitti", n_classes=20, subsample=2500),
    }
    for name, mk in specs.items():
        data = mk()
        model = build_model(rbase.ModelBase, name)
        randomize_model_(model, 1234)
        model.eval()
        torch.manual_seed(79)
        with torch.no_grad():
            logits = model(data, labels=data.y)
        out[name] = {"pos": data.pos, "x": data.x, "batch": data.batch, "curve": data.curve_idxs,
                     "labels": getattr(data, "labels", None), "weight_seed": 1234, "rng_seed": 79, "logits": logits}
        print(name, tuple(logits.shape), float(logits.abs().max()))
    return out


def make_dgcnn(rbase):
    """DynamicEdgeConv (dgcnn.py:16-127) through the REAL reference classes: DGCNNLayer / DGCNNLayerRadius with every
    default aggregation, and a whole ModelBase whose step list uses `dgcnn` twice (no shipped config does)."""
    import src.models.modules.dgcnn as rdg
    from torch_geometric.nn import MLP
    from curvecloudnet_b200.synthetic import object_scan_batch, randomize_model_
    data = object_scan_batch([71, 72], n_points=600)
    gen = torch.Generator().manual_seed(3)
    x = 0.2 * torch.randn(data.pos.shape[0], 5, generator=gen)
    out = {"pos": data.pos, "batch": data.batch, "curve": data.curve_idxs, "x": x, "layers": {}}
    for tag, mk in {
        "knn_max": lambda nn_: rdg.DGCNNLayer(nn_, 8, with_xyz=True),
        "knn_mean": lambda nn_: rdg.DGCNNLayer(nn_, 12, aggr="mean", with_xyz=True),
        "knn_add_posonly": lambda nn_: rdg.DGCNNLayer(nn_, 6, aggr="add", with_xyz=True),
        "ball_max": lambda nn_: rdg.DGCNNLayerRadius(nn_, 0.3, with_xyz=True),
        "ball_mean_noxyz": lambda nn_: rdg.DGCNNLayerRadius(nn_, 0.35, aggr="mean", with_xyz=False),
    }.items():
        feat = None if tag.endswith("posonly") else x
        d = (0 if feat is None else 5) + (0 if tag.endswith("noxyz") else 3)
        mod = mk(MLP([2 * d, 32, 48]))
        randomize_model_(mod, 900 + len(out["layers"]))
        mod.eval()
        with torch.no_grad():
            y = mod(feat, data.pos, data.batch, data.curve_idxs)[0]
        out["layers"][tag] = {"state": {k: v.clone() for k, v in mod.state_dict().items()}, "out": y, "dims": [2 * d, 32, 48]}
        print("dgcnn", tag, tuple(y.shape), float(y.abs().max()))
    model = rbase.ModelBase(in_dim=4, n_out=6, steps=["dgcnn", "dgcnn", "mlp"], feat_dims=[[32, 32], [32, 64], [64]],
                            out_mlp={"dims": [32]}, knn=[10, 6, None], radii=[None, None, None], with_xyz=True)
    randomize_model_(model, 1234)
    model.eval()
    data.x = x[:, :1].contiguous()
    torch.manual_seed(81)
    with torch.no_grad():
        logits = model(data)
    out["model"] = {"kwargs": dict(in_dim=4, n_out=6, steps=["dgcnn", "dgcnn", "mlp"], feat_dims=[[32, 32], [32, 64], [64]],
                                   out_mlp={"dims": [32]}, knn=[10, 6, None], radii=[None, None, None], with_xyz=True),
                    "x": data.x, "weight_seed": 1234, "logits": logits}
    print("dgcnn model", tuple(logits.shape), float(logits.abs().max()))
    return out


def make_helpers(rpo):
    """The two grouping helpers of point_ops.py that no shipped config calls, through the REAL functions."""
    from curvecloudnet_b200.synthetic import object_scan_batch
    data = object_scan_batch([91, 92], n_points=300)
    n = data.pos.shape[0]
    gen = torch.Generator().manual_seed(17)
    out = {"pos": data.pos, "batch": data.batch, "curve": data.curve_idxs, "knn_1d": {}, "first_k": {}}
    idxs = torch.sort(torch.randperm(n, generator=gen)[:97])[0]
    idxs[0], idxs[-1] = 0, n - 1
    out["idxs"] = idxs
    for knn in (1, 3, 8):
        row, col = rpo.knn_1d_group_subset(data.pos, idxs.clone(), data.curve_idxs, data.batch, knn)
        out["knn_1d"][knn] = {"row": row, "col": col}
        print("knn_1d_group_subset", knn, row.shape[0])
    # group_first_k_values: unsorted ids with a gap (id 3 never occurs), 1-D and 2-D values, int and per-batch k
    ids = torch.tensor([2, 0, 0, 5, 2, 2, 1, 0, 5, 2, 4, 4, 0, 2, 1, 2, 2, 0, 5, 4, 2, 0, 0, 2])
    v1 = torch.randn(ids.shape[0], generator=gen)
    v2 = torch.randn(ids.shape[0], 3, generator=gen)
    out["first_k_ids"], out["first_k_v1"], out["first_k_v2"] = ids, v1, v2
    for tag, vals, k in (("v1_k4", v1, 4), ("v2_k3", v2, 3), ("v2_k12", v2, 12),
                         ("v1_ktensor", v1, torch.tensor([3, 1, 6, 2, 2])), ("v2_ktensor", v2, torch.tensor([1, 5, 2, 4, 3]))):
        grouped, mask = rpo.group_first_k_values(vals.clone(), ids.clone(), k.clone() if torch.is_tensor(k) else k)
        out["first_k"][tag] = {"k": k, "grouped": grouped, "mask": mask}
        print("group_first_k_values", tag, tuple(grouped.shape))
    return out


def _reference_class(relpath: str, class_name: str, names, extra=None):
    """A stand-in class whose methods are compiled from the reference's own source text (the module itself cannot be
    imported here: nuscenes-devkit / cv2 / mitsuba / torch_geometric.data are absent).  extra: further module-level names
    the methods use."""
    import numpy as np
    src = open(os.path.join(REF, relpath)).read()
    tree = ast.parse(src)
    ns = {"torch": torch, "np": np, "os": os, "Data": lambda **kw: kw}
    ns.update(extra or {})
    body = {}
    for node in ast.walk(tree):
        if isinstance(node, ast.ClassDef) and node.name == class_name:
            for item in node.body:
                if isinstance(item, ast.FunctionDef) and item.name in names:
                    code = textwrap.dedent(ast.get_source_segment(src, item))
                    exec(compile(code, "%s:%s" % (relpath, item.name), "exec"), ns)
                    static = any(isinstance(d, ast.Name) and d.id == "staticmethod" for d in item.decorator_list)
                    body[item.name] = staticmethod(ns[item.name]) if static else ns[item.name]
                elif isinstance(item, ast.Assign) and isinstance(item.targets[0], ast.Name):
                    body[item.targets[0].id] = eval(compile(ast.Expression(item.value), relpath, "eval"), ns)
    missing = set(names) - set(body)
    if missing:
        raise RuntimeError("not found in %s: %s" % (relpath, sorted(missing)))
    return type(class_name + "FromSource", (), body)


def a2d2_frames():
    """The A2D2-shaped fixtures: a full synthetic frame, a small one, and the small one with sensor 2 absent and a single
    point of sensor 4 (np.unique over the ids that occur, audi_dataset.py:162; the `> 1` guard of :195)."""
    import numpy as np
    from curvecloudnet_b200.synthetic import a2d2_frame
    full = a2d2_frame(1)
    small = a2d2_frame(2, azimuth_steps=150)
    keep = small["lidar_id"] != 2
    one = np.flatnonzero(small["lidar_id"] == 4)[1:]
    keep[one] = False
    ragged = {k: np.ascontiguousarray(v[keep]) for k, v in small.items()}
    return {"full": full, "small": small, "ragged": ragged}


def make_a2d2():
    """SemA2D2.get / _get_curves / _partition_curves_at_discontinuities / _training_augmentation (src/data/audi_dataset.py)
    compiled from source; _load_frame (npz + label image through cv2) is replaced by the synthetic frame.  torch_cluster.knn
    and torch_geometric.utils.to_scipy_sparse_matrix are absent: knn = the oracle's exact (d2, index) kNN, the sparse matrix
    is built with scipy directly; scipy's connected_components is the real one."""
    import numpy as np
    import scipy.sparse as sp
    import scipy.sparse.csgraph  # noqa: F401
    sys.path.insert(0, ROOT)
    from oracle import tp_ops as tp

    def knn(x, y, k, batch_x=None, batch_y=None):
        lx, ly = torch.tensor([x.shape[0]]), torch.tensor([y.shape[0]])
        kk = min(k, x.shape[0])
        idx = tp.knn_points(y[None].float(), x[None].float(), ly, lx, kk)[1][0]          # (ny, kk), ascending (d2, index)
        rows = torch.arange(y.shape[0])[:, None].expand(-1, kk)
        return torch.stack([rows.reshape(-1), idx.reshape(-1)], dim=0)

    def to_scipy_sparse_matrix(edge_index, num_nodes=None):
        r, c = edge_index[0].numpy(), edge_index[1].numpy()
        return sp.coo_matrix((np.ones(r.shape[0]), (r, c)), shape=(num_nodes, num_nodes))

    cls = _reference_class("src/data/audi_dataset.py", "SemA2D2",
                           ["get", "_get_curves", "_partition_curves_at_discontinuities", "_training_augmentation"],
                           extra={"knn": knn, "to_scipy_sparse_matrix": to_scipy_sparse_matrix, "sp": sp})
    out = {}
    for tag, frame in a2d2_frames().items():
        n = frame["points"].shape[0]
        rec = {k: torch.from_numpy(v) for k, v in frame.items()}
        base = int(frame["timestamp"].min())              # the fixture stores int32: timestamps as offsets from the first
        rec["timestamp"] = rec["timestamp"] - base
        rec["timestamp_base"] = base
        ds = cls()
        ds.split = "val"
        ds._load_frame = lambda idx, f=frame, m=n: ({k: v.copy() for k, v in f.items()}, np.arange(m))
        d = ds.get(0)                                   # labels = arange(n): the permutation the reference applied
        rec["val"] = {"pos": d["pos"], "x": d["x"], "reorder": d["y"], "curve_idxs": d["curve_idxs"]}
        ds.split = "train"
        np.random.seed(11)
        d = ds.get(0)
        assert torch.equal(d["curve_idxs"], rec["val"]["curve_idxs"])
        rec["train"] = {"pos": d["pos"], "seed": 11}
        out[tag] = rec
        print("a2d2", tag, n, "points", int(d["curve_idxs"].max()) + 1, "curves")
    return out


def make_objects():
    """Object scans: `pc_normalize` compiled from its source text (src/data/object_datasets.py:61-66) and the body of
    CurvesInMemoryDataset.get (:21-38) around it.  torch_geometric.transforms is absent: Center / NormalizeScale / RandomScale
    are stated from their 2.3.0 definitions (pos - mean; * 0.999999 / max|pos|; * random.uniform) - unpinned third party."""
    import random
    from curvecloudnet_b200.synthetic import object_scan_batch
    src = open(os.path.join(REF, "src/data/object_datasets.py")).read()
    tree = ast.parse(src)
    ns = {"torch": torch}
    for node in tree.body:
        if isinstance(node, ast.FunctionDef) and node.name == "pc_normalize":
            exec(compile(textwrap.dedent(ast.get_source_segment(src, node)), "object_datasets.py:pc_normalize", "exec"), ns)
    pc_normalize = ns["pc_normalize"]
    data = object_scan_batch([31, 32, 33], n_points=768)
    gen = torch.Generator().manual_seed(4)
    raw = data.pos * 3.7 + torch.tensor([0.4, -1.1, 2.0]) + 0.01 * torch.randn(data.pos.shape, generator=gen)   # un-normalised
    out = {"seeds": [31, 32, 33], "raw": raw, "batch": data.batch, "val": [], "train": [], "draws": []}
    random.seed(21)
    torch.manual_seed(21)
    for b in range(3):
        p = raw[data.batch == b].float()
        v = pc_normalize(p)
        out["val"].append(v)
        t = v - v.mean(dim=-2, keepdim=True)                      # T.NormalizeScale = Center + scale
        t = t * ((1 / t.abs().max()) * 0.999999)
        scale = random.uniform(0.95, 1.05)                        # T.RandomScale((0.95, 1.05))
        t = t * scale
        shift = (torch.rand((1, 3)) - 0.5) * 0.05                 # object_datasets.py:36
        t = t + shift
        out["train"].append(t)
        out["draws"].append([scale] + shift[0].tolist())
    out["val"], out["train"] = torch.cat(out["val"]), torch.cat(out["train"])
    out["draws"] = torch.tensor(out["draws"], dtype=torch.float64)
    print("objects", tuple(out["val"].shape), float(out["val"].norm(dim=1).max()), float(out["train"].abs().max()))
    return out


def kortx_scan(seed: int, n: int = 9000):
    """A Kortx-shaped raw scan: points along ~70 laser lines of a height field (f64, 0.5 m extent), a raw curve counter that
    grows at every line change - with spurious single-point increments the way the real files have them - a time stamp and a
    part label per point."""
    import numpy as np
    rng = np.random.default_rng(500 + seed)
    lines = 70
    per = n // lines
    pts, raw, cur = [], [], 0
    for li in range(lines):
        m = per + (n - per * lines if li == lines - 1 else 0)
        s = np.linspace(-0.25, 0.25, m)
        th = rng.uniform(0, np.pi)
        off = rng.uniform(-0.2, 0.2)
        xy = off * np.array([-np.sin(th), np.cos(th)])[None] + s[:, None] * np.array([np.cos(th), np.sin(th)])[None]
        z = 0.08 * np.sin(9 * xy[:, 0]) * np.cos(7 * xy[:, 1])
        hole = rng.integers(0, m - 40)
        z[hole:hole + rng.integers(0, 30)] += 0.35                      # a jump: cut_curves splits here ('table' threshold 0.2)
        pts.append(np.concatenate([xy, z[:, None]], axis=1))
        ids = np.full(m, cur)
        for k in rng.integers(1, m - 2, size=3):                       # spurious +1, +1 on consecutive points
            ids[k:] += 1
            ids[k + 1:] += 1
        raw.append(ids)
        cur = ids[-1] + 1
    pts = np.concatenate(pts).astype(np.float64)
    raw = np.concatenate(raw).astype(np.int64)
    t = (np.arange(pts.shape[0]) * 13).astype(np.int64)
    seg = rng.integers(12, 16, size=pts.shape[0]).astype(np.int64)
    return pts, t, raw, seg


def make_kortx():
    """scanning_simulator/kortx/generate_kortx.py run AS A SCRIPT MODULE (its main(), fix_curve_idxs and cut_curves; pandas is
    here) over a temp directory holding one synthetic instance with two views; its visualisation import is stubbed."""
    import json
    import tempfile
    import types
    import numpy as np
    import pandas as pd
    for p in (ROOT, os.path.join(ROOT, "oracle", "shims"), REF):
        if p not in sys.path:
            sys.path.insert(0, p)
    viz = types.ModuleType("scanning_simulator.utils.visualization")
    viz.visualize_kortx_test_pc = viz.visualize_kortx_test_pc_mitsuba = lambda *a, **k: None
    sys.modules["scanning_simulator.utils.visualization"] = viz
    import importlib.util
    spec = importlib.util.spec_from_file_location("ref_generate_kortx", os.path.join(REF, "scanning_simulator/kortx/generate_kortx.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    out = {}
    for tag, setting, seed, n in (("table", "table", 1, 9000), ("handheld", "handheld", 2, 12345)):
        tmp = tempfile.mkdtemp(prefix="ccn_kortx_")
        os.makedirs(os.path.join(tmp, "inst"))
        with open(os.path.join(tmp, "inst", "info.json"), "w") as f:
            json.dump({"object-class": "chair", "capture-setting": setting}, f)
        pts, t, raw, seg = kortx_scan(seed, n)
        pd.DataFrame(np.concatenate([pts, t[:, None], raw[:, None], seg[:, None]], axis=1)).to_csv(
            os.path.join(tmp, "inst", "view0.csv"), header=False, index=False, float_format="%.17g")
        # what the script will read: pandas' default float parser is not round-trip exact, so the fixture keeps ITS values
        pts = pd.read_csv(os.path.join(tmp, "inst", "view0.csv"), header=None).iloc[:, :3].to_numpy().copy()
        args = types.SimpleNamespace(data_path=tmp, npoints=2048, samples_per_scan=5, viz=False,
                                     outfile=os.path.join(tmp, "out.pth"))
        mod.main(args, "cpu")
        batch = torch.load(args.outfile, weights_only=False)
        out[tag] = {"points": torch.from_numpy(pts), "raw": torch.from_numpy(raw), "seg": torch.from_numpy(seg),
                    "setting": setting, "pos": batch.pos.view(5, 2048, 3), "curve_idxs": batch.curve_idxs.view(5, 2048),
                    "y": batch.y.view(5, 2048), "label": int(batch.labels[0])}
        print("kortx", tag, n, "points ->", tuple(out[tag]["pos"].shape), "curves per window",
              [int(c.max()) + 1 for c in out[tag]["curve_idxs"]])
    return out


def make_ingest():
    """Scan ingest: SemNuScenes.get / SemKITTI.get with their own _load_frame, _get_curves and
    _training_augmentation, compiled from source and run on synthetic raw files written to a temp directory."""
    import tempfile
    import numpy as np
    from curvecloudnet_b200.synthetic import raycast_driving_scan
    out = {}
    tmp = tempfile.mkdtemp(prefix="ccn_ingest_")

    # ---- nuScenes: (n,5) f32 records + u8 lidarseg labels, learning map 0..31 -> 0..16
    raw = raycast_driving_scan(41)
    n = 5000
    rec = np.concatenate([raw.points[:n], raw.reflectance[:n, None], raw.beam_ids[:n, None]], axis=1).astype(np.float32)
    g = np.random.RandomState(5)
    lmap = {k: int(g.randint(0, 17)) for k in range(32)}
    lab = g.randint(0, 32, size=n).astype(np.uint8)
    os.makedirs(os.path.join(tmp, "samples", "LIDAR_TOP"), exist_ok=True)
    os.makedirs(os.path.join(tmp, "lidarseg"), exist_ok=True)
    rec.tofile(os.path.join(tmp, "samples", "LIDAR_TOP", "a.pcd.bin"))
    lab.tofile(os.path.join(tmp, "lidarseg", "a_lidarseg.bin"))

    class _Nusc:
        dataroot = tmp

        def get(self, table, token):
            return {"data": {"LIDAR_TOP": "tok"}} if table == "sample" else {"filename": "lidarseg/a_lidarseg.bin"}

    cls = _reference_class("src/data/nuscenes_dataset.py", "SemNuScenes",
                           ["get", "_load_frame", "_get_curves", "_training_augmentation"])
    cls.len = lambda self: 1
    g_nusc = {"raw": torch.from_numpy(rec), "labels": torch.from_numpy(lab),
              "learning_map": torch.tensor([lmap[k] for k in range(32)], dtype=torch.int32)}
    for split, seed in (("val", None), ("train", 7), ("test", None)):
        ds = cls()
        ds.split, ds.polarmix_aug, ds.learning_map, ds.nusc, ds.data_path = split, False, lmap, _Nusc(), tmp
        ds.nusc_infos = [{"lidar_path": "x/samples/LIDAR_TOP/a.pcd.bin", "token": "t"}]
        if seed is not None:
            np.random.seed(seed)
        d = ds.get(0)
        full = {"pos": d["pos"], "x": d["x"], "y": d["y"].to(torch.uint8), "curve_idxs": d["curve_idxs"], "reorder": d["reorder"]}
        if split != "val":          # only what differs from the validation split is stored
            for k in ("x", "curve_idxs", "reorder"):
                assert torch.equal(full[k], g_nusc["val"][k])
            full = {"pos": full["pos"]} if split == "train" else {"y": full["y"]}
        g_nusc[split] = full
        g_nusc[split + "_seed"] = -1 if seed is None else seed
    out["nuscenes"] = g_nusc

    # ---- KITTI: (n,4) f32 records + i32 labels (instance id in the high 16 bits), sparse learning map
    raw = raycast_driving_scan(42, beams=64, azimuth_steps=300, elev_deg=(-24.8, 2.0))
    n = min(6000, raw.points.shape[0])
    rec = np.concatenate([raw.points[:n], raw.reflectance[:n, None]], axis=1).astype(np.float32)
    keys = [0, 1, 10, 11, 13, 15, 16, 18, 20, 30, 31, 32, 40, 44, 48, 49, 50, 51, 52, 60, 70, 71, 72, 80, 81, 99, 252,
            253, 254, 255, 256, 257, 258, 259]
    kmap = {k: int(g.randint(0, 20)) for k in keys}
    sem = np.asarray(keys, dtype=np.int32)[g.randint(0, len(keys), size=n)]
    klab = (sem | (g.randint(0, 500, size=n).astype(np.int32) << 16)).astype(np.int32)
    os.makedirs(os.path.join(tmp, "seq", "velodyne"), exist_ok=True)
    os.makedirs(os.path.join(tmp, "seq", "labels"), exist_ok=True)
    fbin = os.path.join(tmp, "seq", "velodyne", "000000.bin")
    rec.tofile(fbin)
    klab.tofile(os.path.join(tmp, "seq", "labels", "000000.label"))
    kcls = _reference_class("src/data/kitti_dataset.py", "SemKITTI",
                            ["get", "_load_frame", "_get_curves", "_training_augmentation"])
    lut = np.full(260, -1, dtype=np.int32)
    for k, v in kmap.items():
        lut[k] = v
    g_kitti = {"raw": torch.from_numpy(rec), "labels": torch.from_numpy(klab), "learning_map": torch.from_numpy(lut)}
    for split, seed in (("val", None), ("train", 9), ("test", None)):
        ds = kcls()
        ds.split, ds.polarmix_aug, ds.learning_map, ds.fpaths = split, False, kmap, [fbin]
        if seed is not None:
            np.random.seed(seed)
        d = ds.get(0)
        full = {"pos": d["pos"], "x": d["x"], "y": d["y"].to(torch.uint8), "curve_idxs": d["curve_idxs"]}
        if split != "val":
            for k in ("x", "curve_idxs"):
                assert torch.equal(full[k], g_kitti["val"][k])
            full = {"pos": full["pos"]} if split == "train" else {"y": full["y"]}
        g_kitti[split] = full
        g_kitti[split + "_seed"] = -1 if seed is None else seed
    out["kitti"] = g_kitti
    import shutil
    shutil.rmtree(tmp, ignore_errors=True)
    return out


def polarmix_inputs(n=1500):
    """Two synthetic nuScenes sweeps (records + class labels) for the PolarMix fixture; tests rebuild them from here."""
    import numpy as np
    from curvecloudnet_b200.synthetic import raycast_driving_scan
    out = []
    for seed in (71, 72):
        raw = raycast_driving_scan(seed)
        sl = slice(seed * 10, seed * 10 + n)
        rec = np.concatenate([raw.points[sl], raw.reflectance[sl, None], raw.beam_ids[sl, None]], axis=1).astype(np.float32)
        lab = np.random.RandomState(seed).randint(0, 17, size=n).astype(np.uint8)
        out.append((rec, lab))
    return out


def make_polarmix():
    """PolarMix: the REAL `polarmix` of src/data/data_utils.py (imported from its file), driven
    like `_apply_polarmix` (nuscenes_dataset.py:151-178) drives it: sweep-2 ids += max(sweep-1 ids), rows
    [xyz, reflectance, curve id] concatenated in f32, polarmix(...), torch.unique(..., return_inverse=True)."""
    import importlib.util
    import numpy as np
    from oracle import ccn_ingest as oi
    spec = importlib.util.spec_from_file_location("ref_data_utils", os.path.join(REF, "src/data/data_utils.py"))
    du = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(du)
    ident = {k: k for k in range(32)}
    (rec1, lab1), (rec2, lab2) = polarmix_inputs()
    d1 = oi.ingest_scan(rec1, lab1, ident, "nuscenes", None, pos_div=1.0)
    d2 = oi.ingest_scan(rec2, lab2, ident, "nuscenes", None, pos_div=1.0)
    classes = [2, 3, 4, 5, 6, 7, 9, 10]                                  # nuscenes_dataset.py:12-13
    out = {"labels1": torch.from_numpy(lab1), "labels2": torch.from_numpy(lab2), "instance_classes": classes, "cases": []}
    want = {(True, True), (True, False), (False, True), (False, False)}
    seed = 0
    while want:
        seed += 1
        g = np.random.RandomState(seed)
        gates = (bool(g.random_sample() < 0.5), bool(g.random_sample() < 0.5))
        if gates not in want:
            continue
        want.discard(gates)
        alpha = (np.random.RandomState(100 + seed).random_sample() - 1) * np.pi
        beta = alpha + np.pi
        omega = [0.37 * np.pi * 2 / 3, 1.61 * np.pi * 2 / 3]
        cid2 = d2["curve_idxs"] + torch.max(d1["curve_idxs"])
        in1 = torch.cat([d1["pos"], d1["x"], d1["curve_idxs"].reshape(-1, 1)], dim=-1)
        in2 = torch.cat([d2["pos"], d2["x"], cid2.reshape(-1, 1)], dim=-1)
        np.random.seed(seed)                                             # the two gates are drawn inside polarmix
        aug, laug = du.polarmix(in1.numpy(), d1["y"].numpy().flatten().astype(np.uint8), in2.numpy(),
                                d2["y"].numpy().flatten().astype(np.uint8), alpha=alpha, beta=beta,
                                instance_classes=classes, Omega=omega)
        dense = torch.unique(torch.from_numpy(aug[:, 4].copy()), return_inverse=True)[1]
        out["cases"].append({"seed": seed, "do_swap": gates[0], "do_paste": gates[1], "alpha": float(alpha),
                             "beta": float(beta), "omega": [float(o) for o in omega],
                             "pos": torch.from_numpy(aug[:, :3].copy()), "x": torch.from_numpy(aug[:, 3:4].copy()),
                             "y": torch.from_numpy(laug.astype(np.uint8)), "curve_idxs": dense})
    return out


def _compact(obj):
    """int64 -> int32 (all values are small indices) to keep the committed fixtures small;
    tests/golden_utils.load() widens them again."""
    if torch.is_tensor(obj):
        return obj.to(torch.int32) if obj.dtype == torch.int64 else obj
    if isinstance(obj, dict):
        return {k: _compact(v) for k, v in obj.items()}
    return obj


def main():
    os.makedirs(OUT, exist_ok=True)
    if len(sys.argv) > 1 and sys.argv[1] == "losses":     # only the loss / metrics fixture (needs no shims)
        torch.save(make_losses(), os.path.join(OUT, "losses.pt"))
        print("losses.pt", os.path.getsize(os.path.join(OUT, "losses.pt")) // 1024, "KiB")
        return
    if len(sys.argv) > 1 and sys.argv[1] == "ingest":     # only the scan-ingest fixture (source-compiled dataset methods)
        sys.path.insert(0, ROOT)
        torch.save(_compact(make_ingest()), os.path.join(OUT, "ingest.pt"))
        print("ingest.pt", os.path.getsize(os.path.join(OUT, "ingest.pt")) // 1024, "KiB")
        return
    if len(sys.argv) > 1 and sys.argv[1] == "objects":    # only the object-scan normalisation fixture
        sys.path.insert(0, ROOT)
        torch.save(_compact(make_objects()), os.path.join(OUT, "objects.pt"))
        print("objects.pt", os.path.getsize(os.path.join(OUT, "objects.pt")) // 1024, "KiB")
        return
    if len(sys.argv) > 1 and sys.argv[1] == "kortx":      # only the Kortx raw-scan fixture (the real generate_kortx.py)
        torch.save(_compact(make_kortx()), os.path.join(OUT, "kortx.pt"))
        print("kortx.pt", os.path.getsize(os.path.join(OUT, "kortx.pt")) // 1024, "KiB")
        return
    if len(sys.argv) > 1 and sys.argv[1] == "a2d2":       # only the A2D2 curve-builder fixture
        sys.path.insert(0, ROOT)
        torch.save(_compact(make_a2d2()), os.path.join(OUT, "a2d2.pt"))
        print("a2d2.pt", os.path.getsize(os.path.join(OUT, "a2d2.pt")) // 1024, "KiB")
        return
    if len(sys.argv) > 1 and sys.argv[1] == "polarmix":   # only the PolarMix fixture (the real data_utils.polarmix)
        sys.path.insert(0, ROOT)
        torch.save(_compact(make_polarmix()), os.path.join(OUT, "polarmix.pt"))
        print("polarmix.pt", os.path.getsize(os.path.join(OUT, "polarmix.pt")) // 1024, "KiB")
        return
    rpo, rbase, rconv, rfps = _import_reference()
    if len(sys.argv) > 1 and sys.argv[1] == "helpers":    # only the unused grouping helpers of point_ops.py
        torch.save(_compact(make_helpers(rpo)), os.path.join(OUT, "helpers.pt"))
        print("helpers.pt", os.path.getsize(os.path.join(OUT, "helpers.pt")) // 1024, "KiB")
        return
    if len(sys.argv) > 1 and sys.argv[1] == "dgcnn":      # only the DynamicEdgeConv fixture
        torch.save(_compact(make_dgcnn(rbase)), os.path.join(OUT, "dgcnn.pt"))
        print("dgcnn.pt", os.path.getsize(os.path.join(OUT, "dgcnn.pt")) // 1024, "KiB")
        return
    if len(sys.argv) > 1 and sys.argv[1] == "models_extra":
        torch.save(_compact(make_models_extra(rbase)), os.path.join(OUT, "models_extra.pt"))
        print("models_extra.pt", os.path.getsize(os.path.join(OUT, "models_extra.pt")) // 1024, "KiB")
        return
    if len(sys.argv) > 1 and sys.argv[1] == "train":      # only the training-step fixture
        torch.save(_compact(make_train(rbase)), os.path.join(OUT, "train.pt"))
        print("train.pt", os.path.getsize(os.path.join(OUT, "train.pt")) // 1024, "KiB")
        return
    torch.save(_compact(make_ops(rpo, rconv, rfps)), os.path.join(OUT, "ops.pt"))
    torch.save(_compact(make_models(rbase)), os.path.join(OUT, "models.pt"))
    for f in ("ops.pt", "models.pt"):
        print(f, os.path.getsize(os.path.join(OUT, f)) // 1024, "KiB")


if __name__ == "__main__":
    main()
