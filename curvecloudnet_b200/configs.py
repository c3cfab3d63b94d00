"""The shipped CurveCloudNet step lists as python tables.

`ModelBase` consumes the reference's YAML `model:` section unchanged
(/root/reference/configs/curvecloudnet-{eval,train}/*.yaml).  Those files do not travel to the
GPU box, so the same dictionaries are rebuilt here from one row per step
(step, feat_dims, knn, ratio, radius, num_skips, kernel_size); tests/test_configs.py checks them
against the YAMLs whenever /root/reference is present.
"""
from __future__ import annotations

import copy


def _s(name, **kw):
    return dict(step_name=name, **kw) if kw else name


def _assemble(rows, **top):
    steps, fd, knn, ratios, radii, skips, ks = [], [], [], [], [], [], []
    for step, dims, k, ratio, r, ns, kern in rows:
        steps.append(copy.deepcopy(step)); fd.append(list(dims)); knn.append(k); ratios.append(ratio)
        radii.append(r); skips.append(ns); ks.append(kern)
    cfg = dict(type="generic", steps=steps, feat_dims=fd, knn=knn, ratios=ratios, radii=radii,
               num_skips=skips, kernel_sizes=ks)
    cfg.update(top)
    return cfg


def _driving(voxel1: float):
    conv = _s("conv1d-fast-v2", with_diff=True, with_xyz=True)
    sg = _s("sgcnn", with_xyz=True, aggr_type="max")
    fp = _s("fp", with_xyz=True)
    sk = "skip-connect"

    def sa(aggr, down, **kw):
        return _s("sa", aggr_type=aggr, downsample_type=down, **kw, normalize_radius=True, use_fast_knn=True)

    rows = [
        (conv, (32, 32, 32), None, None, None, None, 5),
        (_s("sa-geo", curve_fps_arclen=0.007, use_curve_fps=True, use_curve_knn=True, with_xyz=True,
            aggr_type="attend", normalize_radius=True), (64, 128, 192, 256), None, None, 0.02, None, None),
        (_s("mlp", plain_last=False, with_xyz=True), (256, 128, 128, 64), None, None, None, None, None),
        (sg, (64, 64, 64), 20, None, 0.04, None, 3),
        (sk, (128, 128, 128), None, None, None, 1, None),
        (sa("attend", "voxel", voxel_size=voxel1), (128, 128, 128), 32, 0.3, 0.04, None, None),
        (sg, (128, 128), 20, None, 0.08, None, None),
        (sk, (256, 256), None, None, None, 1, None),
        (sa("attend", "voxel", voxel_size=0.07), (256, 256, 256), 32, 0.3, 0.1, None, None),
        (sg, (256, 256), 20, None, 0.3, None, None),
        (sk, (512, 512), None, None, None, 1, None),
        (sa("attend", "fps"), (512, 512, 512), 32, 0.3, 0.3, None, None),
        (sg, (512, 512), 20, None, 0.3, None, None),
        (sk, (1024, 1024), None, None, None, 1, None),
        (sa("max", "fps"), (1024, 1024, 1024), 32, 0.3, 0.5, None, None),
        (sg, (1024, 1024), 20, None, 0.8, None, None),
        (_s("sgcnn", aggr_type="max"), (1024, 1024), 20, None, 0.8, None, None),
        (sk, (3072, 2048, 1024), 20, None, None, 2, None),
        (fp, (2051, 1024, 512), 3, None, None, None, None),
        ("sgcnn", (512, 512), 20, None, 0.3, None, None),
        (sk, (1024, 1024, 512), None, None, None, 1, None),
        (fp, (1027, 512, 256), 3, None, None, None, None),
        ("sgcnn", (256, 256), 20, None, 0.3, None, None),
        (sk, (512, 512, 256), None, None, None, 1, None),
        (fp, (515, 256, 128), 3, None, None, None, None),
        ("sgcnn", (128, 128), 20, None, 0.08, None, None),
        (sk, (256, 256, 128), None, None, None, 1, None),
        (fp, (259, 128, 64), 3, None, None, None, None),
        (sg, (64, 64, 64), 20, None, 0.04, None, 3),
        (sk, (128, 64, 64), None, None, None, 1, None),
        (_s("fp-geo", with_xyz=True), (99, 128, 128), 3, None, None, None, None),
        (conv, (32, 32, 32), 8, None, 0.02, None, 5),
        (sk, (160, 128, 64), None, None, None, 1, None),
    ]
    return _assemble(rows, use_bias=False, version=2.0, out_mlp=dict(dims=[64, 64], dropout=0.0),
                     skip_connect_state_store=["conv1d-fast-v2", "sgcnn"])


def nuscenes_model_config():
    """configs/curvecloudnet-eval/nuscenes-curvecloudnet.yaml `model:` (33 steps, in_dim 4, 17 classes)."""
    return _driving(0.03)


def kitti_model_config():
    """configs/curvecloudnet-eval/kitti-curvecloudnet.yaml `model:` (voxel 0.025; in_dim 4, 20 classes)."""
    return _driving(0.025)


def _object(arclen, r_geo, geo_dims, mlp_dims, conv_k, sa_r, sa_ratio, knn, tail, out_mlp, a2d2=False):
    """A2D2 / ShapeNet-seg / Kortx / ShapeNet-class family."""
    conv = _s("conv1d-fast-v1", with_diff=True, with_xyz=True)
    sk = "skip-connect"
    fp = _s("fp", with_xyz=True)
    if a2d2:
        sg_xyz = _s("sgcnn", with_xyz=True, aggr_type="attend", use_sparse_feat_agg=True)
        sg = _s("sgcnn", aggr_type="attend", use_sparse_feat_agg=True)
        sg_r = [0.25, 0.5, 1.5]
        conv_r = 0.1
    else:
        sg_xyz = _s("sgcnn", with_xyz=True, use_fast_knn=False, use_sparse_feat_agg=True)
        sg = _s("sgcnn", use_fast_knn=False, use_sparse_feat_agg=True)
        sg_r = [None, None, None]
        conv_r = None
    sa = _s("sa", aggr_type="attend", normalize_radius=True, use_fast_knn=False, downsample_type="fps")
    c1 = mlp_dims[-1]
    rows = [
        (_s("sa-geo", curve_fps_arclen=arclen, use_curve_fps=True, use_curve_knn=True, with_xyz=True,
            aggr_type="attend", normalize_radius=True), geo_dims, None, None, r_geo, None, None),
        (_s("mlp", plain_last=False, with_xyz=True), mlp_dims, None, None, None, None, None),
        (conv, (c1, c1), knn, None, conv_r, None, conv_k),
        (sk, tail["skip3"], None, None, None, 1, None),
        (sa, (128, 128, 128), None, sa_ratio[0], sa_r[0], None, None),
        (sg_xyz, (128, 128), knn, None, sg_r[0], None, None),
        (sk, (256, 256), None, None, None, 1, None),
        (sa, (256, 256, 256), None, sa_ratio[1], sa_r[1], None, None),
        (sg_xyz, (256, 256), knn, None, sg_r[1], None, None),
        (sk, (512, 512), None, None, None, 1, None),
        (sa, (512, 512, 512), None, sa_ratio[2], sa_r[2], None, None),
        (sg_xyz, (512, 512), knn, None, sg_r[2], None, None),
        (sg, (512, 512), knn, None, sg_r[2], None, None),
        (sk, tail["skip13"], knn, None, None, 2, None),
    ]
    if tail.get("classification"):
        rows.append(("sa-global", (1024, 1024), None, None, None, None, None))
    else:
        rows += [
            (fp, (1027, 512, 256), 3, None, None, None, None),
            (sg, (256, 256), knn, None, sg_r[1], None, None),
            (sk, (512, 512, 256), None, None, None, 1, None),
            (fp, (515, 256, 128), 3, None, None, None, None),
            (sg, (128, 128), knn, None, sg_r[0], None, None),
            (sk, (256, 256, 128), None, None, None, 1, None),
            (fp, tail["fp20"], 3, None, None, None, None),
            (conv, tail["conv21"], knn, None, conv_r, None, conv_k),
            (sk, tail["skip22"], None, None, None, 1, None),
            (_s("fp-geo", with_xyz=True), tail["fpgeo"], 3, None, None, None, None),
            (conv, tail["conv24"], None, None, None, None, conv_k),
            (sk, tail["skip25"], None, None, None, 1, None),
        ]
    return _assemble(rows, use_bias=True, version=1.0, out_mlp=out_mlp,
                     skip_connect_state_store=["conv1d-fast-v1", "sgcnn"])


def shapenet_class_model_config():
    """configs/curvecloudnet-eval/shapenet-class-curvecloudnet.yaml (15 steps, in_dim 3, 16 classes)."""
    return _object(0.04, 0.075, (64, 128, 256, 512), (256, 128, 64), 7, (0.2, 0.4, 0.8), (0.25, 0.25, 0.25), 30,
                   dict(skip3=(128, 128), skip13=(1536, 1024, 1024), classification=True),
                   dict(dims=[512, 256, 128], dropout=0.0, with_seg_category=False))


def shapenet_seg_model_config():
    """configs/curvecloudnet-eval/shapenet-seg-curvecloudnet.yaml (26 steps, in_dim 3, 50 part classes)."""
    return _object(0.03, 0.04, (64, 128, 256, 512, 1024), (512, 256, 128), 5, (0.18, 0.35, 0.7), (0.25, 0.25, 0.5), 20,
                   dict(skip3=(256, 128), skip13=(1536, 1024, 512), fp20=(259, 128, 128), conv21=(128, 128),
                        skip22=(256, 128, 128), fpgeo=(134, 128, 128), conv24=(64, 64), skip25=(192, 128, 64)),
                   dict(dims=[64, 64], dropout=0.0, with_seg_category=True))


def kortx_model_config():
    """configs/curvecloudnet-eval/kortx-testsplit-curvecloudnet.yaml (26 steps, conv k=7, knn 30)."""
    return _object(0.04, 0.075, (64, 128, 256, 512), (256, 128, 64), 7, (0.2, 0.4, 0.8), (0.25, 0.25, 0.25), 30,
                   dict(skip3=(128, 128), skip13=(1536, 1024, 512), fp20=(259, 128, 64), conv21=(64, 48),
                        skip22=(112, 64, 64), fpgeo=(70, 64, 64), conv24=(64, 64), skip25=(128, 64, 64)),
                   dict(dims=[64, 64], dropout=0.0, with_seg_category=True))


def a2d2_model_config():
    """configs/curvecloudnet-eval/audi-curvecloudnet.yaml (26 steps, FRNN sgcnn with attention)."""
    return _object(0.01, 0.015, (64, 128, 256, 512), (256, 128, 64), 5, (0.03, 0.06, 0.15), (0.35, 0.25, 0.25), 30,
                   dict(skip3=(128, 128), skip13=(1536, 1024, 512), fp20=(259, 128, 64), conv21=(64, 64),
                        skip22=(128, 64, 64), fpgeo=(68, 128, 128), conv24=(128, 128), skip25=(256, 128, 64)),
                   dict(dims=[64, 64], dropout=0.0), a2d2=True)


MODEL_CONFIGS = {
    "nuscenes": (nuscenes_model_config, 4, 17),
    "kitti": (kitti_model_config, 4, 20),
    "shapenet-class": (shapenet_class_model_config, 3, 16),
    "shapenet-seg": (shapenet_seg_model_config, 3, 50),
    "kortx": (kortx_model_config, 3, 2),
    "a2d2": (a2d2_model_config, 4, 38),
}


def build_model(model_cls, name: str, **extra):
    """Instantiate `model_cls` (ModelBase of this package, of the oracle, or the reference's) the way
    src/utils/load_utils.py:17-27 does."""
    fn, in_dim, n_out = MODEL_CONFIGS[name]
    cfg = fn()
    kw = {k: v for k, v in cfg.items() if k not in ("steps", "feat_dims", "out_mlp", "type")}
    kw.update(extra)
    return model_cls(in_dim, n_out, steps=cfg["steps"], feat_dims=cfg["feat_dims"], out_mlp=cfg["out_mlp"],
                     keypoint_visualization="none", dataset_source=name, **kw)
