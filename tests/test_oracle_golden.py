"""The oracle's restatement of the reference's own python against the golden vectors that
oracle/make_golden.py produced by running the REAL reference sources (CPU, not gpu)."""
import pytest
import torch

from golden_utils import FixedDraws, load, rel_err
from oracle import ccn_modules as M
from oracle import ccn_point_ops as po

OPS = load("ops.pt")
CASES = [c for c in OPS if c != "get_curves"]


@pytest.mark.parametrize("case", CASES)
def test_csr_helpers(case):
    g = OPS[case]
    assert torch.equal(po.curve_ids_global(g["curve"], g["batch"]), g["glob"])
    assert torch.equal(po.batch_to_ptr(g["batch"]), g["ptr"])
    assert torch.equal(po.batch_to_ptr(g["batch"], True), g["ptr_ends"])
    padded, mask = po.pad_by_batch(g["pos"], g["batch"])
    assert torch.equal(mask, g["padded_mask"])
    assert torch.allclose(padded.sum(dim=1), g["padded_sum"])


@pytest.mark.parametrize("case", CASES)
def test_samplers(case):
    g = OPS[case]
    assert torch.equal(po.curve_fps(g["pos"], g["batch"], g["curve"], g["curvefps_arclen"], g["curvefps_u"]),
                       g["curvefps_idx"])
    assert torch.equal(po.voxel_fps(g["pos"], g["batch"], g["voxel_size"], g["voxel_noise"]), g["voxel_idx"])
    d = FixedDraws(starts=g["fps_starts"].tolist())
    assert torch.equal(po.fps_sorted(g["pos"], g["batch"], g["fps_ratio"], d), g["fps_idx"])


@pytest.mark.parametrize("case", CASES)
def test_groupings(case):
    g = OPS[case]
    pos, batch, curve, idx, sub = g["pos"], g["batch"], g["curve"], g["curvefps_idx"], g["voxel_idx"]
    row, col = po.radius_1d_group_subset(pos, idx, curve, batch, g["r1d_radius"])
    assert torch.equal(row, g["r1d_row"]) and torch.equal(col, g["r1d_col"])
    row, col = po.knn_1d_group_superset(pos, idx, curve, batch, 3)
    assert torch.equal(row, g["s1d_row"]) and torch.equal(col, g["s1d_col"])
    r = g["knn_radius"]
    dense, _, _ = po.knn_ball_group(pos[sub], pos, batch[sub], batch, "knn", radius=r, knn=16, return_dense=True)
    assert torch.equal(dense, g["frnn_dense"])
    row, col = po.knn_ball_group(pos[sub], pos, batch[sub], batch, "knn", radius=r, knn=16)
    assert torch.equal(row, g["frnn_row"]) and torch.equal(col, g["frnn_col"])
    row, col = po.knn_ball_group(pos[sub], pos, batch[sub], batch, "ball-group", radius=r)
    assert torch.equal(row, g["ball_row"]) and torch.equal(col, g["ball_col"])
    row, col = po.knn_ball_group(pos, pos[sub], batch, batch[sub], "knn", knn=3, accel_knn=False)
    assert torch.equal(row, g["knn3_row"]) and torch.equal(col, g["knn3_col"])


@pytest.mark.parametrize("case", CASES)
def test_float_ops(case):
    from curvecloudnet_b200.synthetic import randomize_model_
    g = OPS[case]
    pos, batch, curve, idx, sub = g["pos"], g["batch"], g["curve"], g["curvefps_idx"], g["voxel_idx"]
    assert torch.equal(M.feature_diffs(g["feat"], curve, batch), g["feat_diffs"])
    y = po.knn_interpolate(g["feat_sub"], pos[sub], pos, batch[sub], batch, k=3)
    assert rel_err(y, g["interp3"]) < 1e-6
    y = po.knn_interpolate_1d(g["feat_idx"], idx, pos, batch, curve, 3)
    assert rel_err(y, g["interp1d"]) < 1e-6
    for ver, cls in (("v2", M.SymmetricCurve1DConvV2), ("v1", M.SymmetricCurve1DConvFastV1)):
        mod = cls([9, 16, 12], g["conv_%s_k" % ver], with_xyz=True, with_diff=True)
        randomize_model_(mod, 55)
        mod.eval()
        with torch.no_grad():
            out = mod(g["feat"], pos, batch, curve)[0]
        assert rel_err(out, g["conv_%s_out" % ver]) < 1e-6


def test_get_curves():
    g = OPS["get_curves"]
    reorder, inv, cid = po.get_curves(g["points"], g["beam"], 0.08)
    assert torch.equal(inv, g["inv_reorder"])
    assert torch.equal(cid, g["curve_idxs"])
    assert torch.equal(g["points"][reorder], g["sorted_points"])


@pytest.mark.parametrize("name", ["nuscenes", "shapenet-class", "shapenet-seg", "a2d2", "kortx", "kitti"])
def test_model_logits(name):
    """End to end: oracle ModelBase == the reference's ModelBase (run over the shims) on the same
    seeded weights, inputs and RNG draws."""
    from curvecloudnet_b200.configs import build_model
    from curvecloudnet_b200.synthetic import ScanBatch, randomize_model_
    from oracle.ccn_model import ModelBase
    g = load("models.pt" if name in ("nuscenes", "shapenet-class", "shapenet-seg") else "models_extra.pt")[name]
    data = ScanBatch(g["x"], g["pos"], g["batch"], g["curve"], None, g["labels"])
    model = build_model(ModelBase, name)
    randomize_model_(model, g["weight_seed"])
    model.eval()
    torch.manual_seed(g["rng_seed"])
    with torch.no_grad():
        logits = model(data)
    assert logits.shape == g["logits"].shape
    assert rel_err(logits, g["logits"]) < 2e-6


# ---------------------------------------------------------------------------------------------------------------
# training step: the oracle's autograd against the step frozen from the REAL reference sources (train.pt)
# ---------------------------------------------------------------------------------------------------------------
import pytest  # noqa: E402

from golden_utils import check_train_golden  # noqa: E402


@pytest.mark.parametrize("mode", ["frozen_bn", "batch_stats"])
@pytest.mark.parametrize("name", ["shapenet-seg", "nuscenes"])
def test_oracle_training_step_matches_reference(name, mode):
    import torch
    import torch.nn.functional as F
    from curvecloudnet_b200.configs import build_model
    from curvecloudnet_b200.synthetic import randomize_model_
    from oracle.ccn_model import ModelBase
    from oracle.tp_nn import Batch
    g = load("train.pt")[name]
    model = build_model(ModelBase, name)
    randomize_model_(model, g["weight_seed"])
    model.train()
    if mode == "frozen_bn":
        for m in model.modules():
            if isinstance(m, torch.nn.BatchNorm1d):
                m.eval()
    kw = dict(x=g["x"], pos=g["pos"], batch=g["batch"], curve_idxs=g["curve"])
    if g["labels"] is not None:
        kw["labels"] = g["labels"]
    torch.manual_seed(g["rng_seed"])
    logits = model(Batch(**kw))
    loss = F.nll_loss(F.log_softmax(logits, 1), g["target"])
    loss.backward()
    # the restatement runs the same CPU torch kernels as the reference did: agreement to rounding
    check_train_golden(model, logits, loss, g[mode], 1e-5, 1e-4 if mode == "frozen_bn" else 1e-3, "%s %s oracle" % (name, mode))


# ---------------------------------------------------------------------------------------------------------------
# DynamicEdgeConv (dgcnn.py:16-127; reached by no shipped config) against the REAL reference classes (dgcnn.pt)
# ---------------------------------------------------------------------------------------------------------------
DGCNN_LAYERS = {
    "knn_max": lambda Mo, nn_: Mo.DGCNNLayer(nn_, 8, with_xyz=True),
    "knn_mean": lambda Mo, nn_: Mo.DGCNNLayer(nn_, 12, aggr="mean", with_xyz=True),
    "knn_add_posonly": lambda Mo, nn_: Mo.DGCNNLayer(nn_, 6, aggr="add", with_xyz=True),
    "ball_max": lambda Mo, nn_: Mo.DGCNNLayerRadius(nn_, 0.3, with_xyz=True),
    "ball_mean_noxyz": lambda Mo, nn_: Mo.DGCNNLayerRadius(nn_, 0.35, aggr="mean", with_xyz=False),
}


@pytest.mark.parametrize("tag", list(DGCNN_LAYERS))
def test_dgcnn_layers(tag):
    from oracle.tp_nn import MLP
    g = load("dgcnn.pt")
    rec = g["layers"][tag]
    mod = DGCNN_LAYERS[tag](M, MLP(rec["dims"]))
    mod.load_state_dict(rec["state"], strict=True)
    mod.eval()
    x = None if tag.endswith("posonly") else g["x"]
    with torch.no_grad():
        out = mod(x, g["pos"], g["batch"], g["curve"])[0]
    assert out.shape == rec["out"].shape
    assert rel_err(out, rec["out"]) < 2e-6


def test_dgcnn_model_logits():
    from curvecloudnet_b200.synthetic import ScanBatch, randomize_model_
    from oracle.ccn_model import ModelBase
    g = load("dgcnn.pt")
    rec = g["model"]
    model = ModelBase(**rec["kwargs"])
    randomize_model_(model, rec["weight_seed"])
    model.eval()
    with torch.no_grad():
        logits = model(ScanBatch(rec["x"], g["pos"], g["batch"], g["curve"], None, None))
    assert rel_err(logits, rec["logits"]) < 2e-6


# ---------------------------------------------------------------------------------------------------------------
# the two grouping helpers of point_ops.py that no shipped config calls, against the REAL functions (helpers.pt)
# ---------------------------------------------------------------------------------------------------------------
def test_unused_grouping_helpers():
    g = load("helpers.pt")
    for knn, rec in g["knn_1d"].items():
        row, col = po.knn_1d_group_subset(g["pos"], g["idxs"], g["curve"], g["batch"], int(knn))
        assert torch.equal(row, rec["row"]) and torch.equal(col, rec["col"]), knn
    for tag, rec in g["first_k"].items():
        vals = g["first_k_v1"] if tag.startswith("v1") else g["first_k_v2"]
        grouped, mask = po.group_first_k_values(vals, g["first_k_ids"], rec["k"])
        assert torch.equal(mask, rec["mask"]) and torch.equal(grouped, rec["grouped"]), tag


# ---------------------------------------------------------------------------------------------------------------
# A2D2 curve builder (audi_dataset.py:120-243) against SemA2D2's own methods (a2d2.pt)
# ---------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("tag", ["full", "small", "ragged"])
def test_a2d2_curve_builder(tag):
    import numpy as np
    from oracle import ccn_ingest as I
    g = load("a2d2.pt")[tag]
    frame = {"points": g["points"].numpy(), "reflectance": g["reflectance"].numpy(), "lidar_id": g["lidar_id"].numpy(),
             "timestamp": g["timestamp"].numpy() + int(g["timestamp_base"])}
    got = I.a2d2_ingest_frame(frame)
    assert torch.equal(got["reorder"], g["val"]["reorder"])
    assert torch.equal(got["curve_idxs"], g["val"]["curve_idxs"])
    assert torch.equal(got["pos"], g["val"]["pos"]) and torch.equal(got["x"], g["val"]["x"])
    np.random.seed(int(g["train"]["seed"]))
    aug = I.a2d2_draw_augmentation()
    assert torch.equal(I.a2d2_ingest_frame(frame, aug)["pos"], g["train"]["pos"])
