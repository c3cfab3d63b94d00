"""Scan ingest: raw sweep records -> Data fields.

tests/golden/ingest.pt holds the outputs of the reference's OWN `get()` / `_load_frame` / `_get_curves` /
`_training_augmentation` (SemNuScenes, SemKITTI; compiled from their source text by oracle/make_golden.py `ingest`) on
synthetic raw files.  CPU tests pin the oracle restatement (oracle/ccn_ingest.py) against it; GPU tests compare
`ccn_ingest_scan` (through curvecloudnet_b200.ingest) with the fixture and, on full-size sweeps, with the oracle.
Tolerances: everything integer / index / label / reflectance is bit-exact; positions are bit-exact without augmentation
and within 1 ulp (f32) with it (the reference's rotation is a BLAS f64 product whose summation order is not specified).
"""
import numpy as np
import pytest
import torch

from golden_utils import load

G = load("ingest.pt")


def _lmap_dict(lut):
    return {k: int(v) for k, v in enumerate(lut.tolist()) if v >= 0}


def _labels_np(ds):
    lab = G[ds]["labels"]
    return lab.numpy().astype(np.uint8) if ds == "nuscenes" else lab.numpy().astype(np.int32)


def _ulp_close(a, b, ulps=1, floor=1e-30):
    """|a - b| <= ulps * 2^-23 * max(|b|, floor); `floor` > 0 where a translation may cancel the magnitude."""
    a, b = a.double(), b.double()
    tol = ulps * torch.maximum(b.abs(), torch.full_like(b, floor)).float().double() * 2.0 ** -23
    return bool(torch.all((a - b).abs() <= tol))


# ------------------------------------------------------------------------------------------------ CPU: the oracle
@pytest.mark.parametrize("ds", ["nuscenes", "kitti"])
def test_oracle_ingest_matches_reference(ds):
    from oracle import ccn_ingest as oi
    g = G[ds]
    raw = g["raw"].numpy()
    lm = _lmap_dict(g["learning_map"])
    val = oi.ingest_scan(raw, _labels_np(ds), lm, ds, None)
    for k in ("pos", "x", "curve_idxs"):
        assert torch.equal(val[k], g["val"][k]), k
    assert torch.equal(val["y"], g["val"]["y"].long())
    if ds == "nuscenes":
        assert torch.equal(val["inv_reorder"], g["val"]["reorder"])
    np.random.seed(int(g["train_seed"]))
    aug = oi.draw_augmentation(ds)
    tr = oi.ingest_scan(raw, _labels_np(ds), lm, ds, aug)
    assert _ulp_close(tr["pos"], g["train"]["pos"])
    assert (tr["pos"] != g["train"]["pos"]).float().mean() < 1e-3      # the 1-ulp allowance is for rare ties only
    te = oi.ingest_scan(raw, None, lm, ds, None)
    assert torch.equal(te["y"], g["test"]["y"].long())


@pytest.mark.parametrize("ds", ["nuscenes", "kitti"])
def test_augmentation_draws_replay_the_reference_order(ds):
    from curvecloudnet_b200.ingest import ScanAugmentation
    from oracle import ccn_ingest as oi
    np.random.seed(123)
    a = oi.draw_augmentation(ds)
    np.random.seed(123)
    b = ScanAugmentation.draw(ds)
    assert (a["rot_c"], a["rot_s"], a["flip"], a["scale"]) == (b.rot_c, b.rot_s, b.flip, b.scale)
    assert tuple(a["translate"]) == tuple(b.translate)
    if ds == "kitti":
        assert b.rot_c >= 0.0                                          # +-90 degrees


# ------------------------------------------------------------------------------------------------ GPU: the CUDA path
@pytest.mark.gpu
@pytest.mark.parametrize("ds", ["nuscenes", "kitti"])
def test_ingest_matches_reference_golden(ds):
    from curvecloudnet_b200.ingest import LearningMap, ScanAugmentation, ingest_scan
    g = G[ds]
    lm = LearningMap(_lmap_dict(g["learning_map"]))
    raw_bytes = g["raw"].numpy().tobytes()                              # the .bin file content
    lab_bytes = _labels_np(ds).tobytes()
    d = ingest_scan(raw_bytes, lab_bytes, lm, ds)
    for k in ("pos", "x", "curve_idxs"):
        assert torch.equal(getattr(d, k).cpu(), g["val"][k]), k
    assert torch.equal(d.y.cpu(), g["val"]["y"].long())
    if ds == "nuscenes":
        assert torch.equal(d.reorder.cpu(), g["val"]["reorder"])
    assert int(d.counts[0]) == int(g["val"]["curve_idxs"].max()) + 1
    np.random.seed(int(g["train_seed"]))
    t = ingest_scan(g["raw"], _labels_np(ds), lm, ds, ScanAugmentation.draw(ds))
    assert _ulp_close(t.pos.cpu(), g["train"]["pos"])
    assert (t.pos.cpu() != g["train"]["pos"]).float().mean() < 1e-3
    assert torch.equal(t.y.cpu(), g["val"]["y"].long()) and torch.equal(t.curve_idxs.cpu(), g["val"]["curve_idxs"])
    te = ingest_scan(g["raw"], None, None, ds)
    assert torch.equal(te.y.cpu(), g["test"]["y"].long())


@pytest.mark.gpu
@pytest.mark.parametrize("ds,seed,beams,az,elev", [("nuscenes", 51, 32, 1085, (-30.67, 10.67)),
                                                   ("kitti", 52, 64, 1875, (-24.8, 2.0))])
def test_ingest_full_sweep_matches_oracle(ds, seed, beams, az, elev):
    from curvecloudnet_b200.ingest import LearningMap, ScanAugmentation, ingest_scan
    from curvecloudnet_b200.synthetic import raycast_driving_scan
    from oracle import ccn_ingest as oi
    raw = raycast_driving_scan(seed, beams=beams, azimuth_steps=az, elev_deg=elev)
    cols = [raw.points, raw.reflectance[:, None]] + ([raw.beam_ids[:, None]] if ds == "nuscenes" else [])
    rec = np.concatenate(cols, axis=1).astype(np.float32)
    n = rec.shape[0]
    gen = np.random.RandomState(seed)
    mapping = {k: int(gen.randint(0, 17)) for k in range(0, 64, 2)}
    keys = np.asarray(sorted(mapping))
    lab = keys[gen.randint(0, len(keys), size=n)]
    lab = lab.astype(np.uint8) if ds == "nuscenes" else (lab | (gen.randint(0, 99, size=n) << 16)).astype(np.int32)
    np.random.seed(seed)
    aug_o = oi.draw_augmentation(ds)
    np.random.seed(seed)
    aug = ScanAugmentation.draw(ds)
    for a_o, a in ((None, None), (aug_o, aug)):
        want = oi.ingest_scan(rec, lab, mapping, ds, a_o)
        got = ingest_scan(rec, lab, LearningMap(mapping), ds, a)
        assert torch.equal(got.x.cpu(), want["x"]) and torch.equal(got.y.cpu(), want["y"])
        assert torch.equal(got.curve_idxs.cpu(), want["curve_idxs"]) and torch.equal(got.reorder.cpu(), want["inv_reorder"])
        if a is None:
            assert torch.equal(got.pos.cpu(), want["pos"])
        else:
            assert _ulp_close(got.pos.cpu(), want["pos"])
    # size-independent properties: x / y are a permutation of the inputs, curve ids are non-decreasing and dense
    assert torch.equal(torch.sort(got.x.cpu().flatten())[0], torch.sort(torch.from_numpy(rec[:, 3].copy()))[0])
    cid = got.curve_idxs.cpu()
    assert int(cid[0]) == 0 and bool(torch.all((cid[1:] - cid[:-1]) >= 0)) and bool(torch.all((cid[1:] - cid[:-1]) <= 1))


@pytest.mark.gpu
def test_ingest_edge_cases_and_errors():
    from curvecloudnet_b200.ingest import LearningMap, ingest_scan
    lm = LearningMap({0: 0, 1: 3})
    d = ingest_scan(np.zeros((0, 5), np.float32), np.zeros((0,), np.uint8), lm, "nuscenes")
    assert d.pos.shape == (0, 3) and d.y.shape == (0,) and d.counts.tolist() == [0, 0]
    one = np.array([[1.0, 2.0, 3.0, 7.0, 5.0]], np.float32)
    d = ingest_scan(one, np.array([1], np.uint8), lm, "nuscenes")
    assert d.y.tolist() == [3] and d.curve_idxs.tolist() == [0] and d.x.tolist() == [[7.0]]
    assert torch.equal(d.pos.cpu(), torch.tensor([[1.0, 2.0, 3.0]]) / 20)
    with pytest.raises(KeyError):                                       # learning_map.__getitem__ KeyError upstream
        ingest_scan(one, np.array([2], np.uint8), lm, "nuscenes")
    with pytest.raises(ValueError):                                     # ring id outside [0, n_beams)
        ingest_scan(np.array([[1.0, 2.0, 3.0, 7.0, 40.0]], np.float32), np.array([1], np.uint8), lm, "nuscenes")
    with pytest.raises(ValueError):                                     # truncated file
        ingest_scan(np.zeros(7, np.float32), None, None, "nuscenes")
    with pytest.raises(ValueError):
        ingest_scan(one, np.array([1, 1], np.uint8), lm, "nuscenes")


@pytest.mark.gpu
def test_ingest_batch_feeds_the_model():
    """raw records -> ingest_batch -> forward: the ingested Batch is what ModelBase.forward consumes."""
    from curvecloudnet_b200.configs import build_model
    from curvecloudnet_b200.ingest import LearningMap, ingest_batch
    from curvecloudnet_b200.synthetic import randomize_model_, raycast_driving_scan
    sweeps = []
    for seed in (61, 62):
        raw = raycast_driving_scan(seed, azimuth_steps=200)
        rec = np.concatenate([raw.points, raw.reflectance[:, None], raw.beam_ids[:, None]], axis=1).astype(np.float32)
        sweeps.append({"raw": rec, "labels": np.zeros(rec.shape[0], np.uint8)})
    batch = ingest_batch(sweeps, LearningMap({0: 0}), "nuscenes", augment=True, rng=np.random.RandomState(3))
    n = sum(s["raw"].shape[0] for s in sweeps)
    assert batch.pos.shape == (n, 3) and batch.batch.shape == (n,) and int(batch.batch[-1]) == 1
    from curvecloudnet_b200.model import ModelBase
    model = build_model(ModelBase, "nuscenes")
    randomize_model_(model, 1234)
    model = model.cuda().eval()
    with torch.no_grad():
        out = model(batch)
    assert out.shape[0] == n and bool(torch.isfinite(out).all())


# ================================================================================================ PolarMix
# tests/golden/polarmix.pt: outputs of the REAL `polarmix` (src/data/data_utils.py:69-85, with swap / rotate_copy) driven
# like `_apply_polarmix` drives it, for the four gate combinations.  Selection, order, labels, reflectance and dense
# curve ids are bit-exact; the rotated copies are within 1 ulp (f64 BLAS product upstream).
PM = load("polarmix.pt")


def _pm_sweeps():
    from oracle.make_golden import polarmix_inputs
    return polarmix_inputs()


def _pm_check(got, case):
    n = case["pos"].shape[0]
    assert got["pos"].shape[0] == n
    assert torch.equal(got["y"].cpu(), case["y"].long())
    assert torch.equal(got["x"].cpu().reshape(-1), case["x"].reshape(-1))
    assert torch.equal(got["curve_idxs"].cpu(), case["curve_idxs"])
    assert _ulp_close(got["pos"].cpu(), case["pos"])
    assert (got["pos"].cpu() != case["pos"]).float().mean() < 1e-3


@pytest.mark.parametrize("ci", range(4))
def test_oracle_polarmix_matches_reference(ci):
    from oracle import ccn_ingest as oi
    case = PM["cases"][ci]
    ident = {k: k for k in range(32)}
    (rec1, lab1), (rec2, lab2) = _pm_sweeps()
    assert torch.equal(torch.from_numpy(lab1), PM["labels1"].to(torch.uint8))
    d1 = oi.ingest_scan(rec1, lab1, ident, "nuscenes", None, pos_div=1.0)
    d2 = oi.ingest_scan(rec2, lab2, ident, "nuscenes", None, pos_div=1.0)
    mix = {"alpha": case["alpha"], "beta": case["beta"], "omega": case["omega"], "do_swap": case["do_swap"],
           "do_paste": case["do_paste"]}
    _pm_check(oi.polarmix(d1, d2, mix, PM["instance_classes"]), case)


def test_polarmix_draws_replay_the_reference_order():
    from curvecloudnet_b200.ingest import PolarMixDraw
    from oracle import ccn_ingest as oi
    np.random.seed(77)
    a = oi.draw_polarmix(1000)
    np.random.seed(77)
    b = PolarMixDraw.draw(1000)
    assert (a["idx2"], a["alpha"], a["beta"], a["omega"][0], a["omega"][1], a["do_swap"], a["do_paste"]) == \
        (b.idx2, b.alpha, b.beta, b.omega0, b.omega1, b.do_swap, b.do_paste)


@pytest.mark.gpu
@pytest.mark.parametrize("ci", range(4))
def test_polarmix_matches_reference_golden(ci):
    from curvecloudnet_b200.ingest import LearningMap, PolarMixDraw, ingest_scan, polarmix
    case = PM["cases"][ci]
    lm = LearningMap({k: k for k in range(32)})
    (rec1, lab1), (rec2, lab2) = _pm_sweeps()
    d1 = ingest_scan(rec1, lab1, lm, "nuscenes", pos_div=1.0)
    d2 = ingest_scan(rec2, lab2, lm, "nuscenes", pos_div=1.0)
    mix = PolarMixDraw(0, case["alpha"], case["beta"], case["omega"][0], case["omega"][1], case["do_swap"], case["do_paste"])
    out = polarmix(d1, d2, mix, PM["instance_classes"])
    _pm_check({"pos": out.pos, "x": out.x, "y": out.y, "curve_idxs": out.curve_idxs}, case)
    assert out.counts.tolist() == [case["pos"].shape[0], int(case["curve_idxs"].max()) + 1]


@pytest.mark.gpu
@pytest.mark.parametrize("ds,beams,az,elev", [("nuscenes", 32, 1085, (-30.67, 10.67)), ("kitti", 64, 1875, (-24.8, 2.0))])
def test_polarmix_full_sweeps_match_oracle(ds, beams, az, elev):
    """Full-size sweeps through ingest -> PolarMix -> augmentation, every gate combination, against the oracle."""
    from curvecloudnet_b200.ingest import (INSTANCE_CLASSES, LearningMap, PolarMixDraw, ScanAugmentation,
                                           ingest_scan_polarmix)
    from curvecloudnet_b200.synthetic import raycast_driving_scan
    from oracle import ccn_ingest as oi
    sweeps = []
    for seed in (81, 82):
        raw = raycast_driving_scan(seed, beams=beams, azimuth_steps=az, elev_deg=elev)
        cols = [raw.points, raw.reflectance[:, None]] + ([raw.beam_ids[:, None]] if ds == "nuscenes" else [])
        rec = np.concatenate(cols, axis=1).astype(np.float32)
        lab = np.random.RandomState(seed).randint(0, 17, size=rec.shape[0])
        sweeps.append({"raw": rec, "labels": lab.astype(np.uint8 if ds == "nuscenes" else np.int32)})
    ident = {k: k for k in range(32)}
    lm = LearningMap(ident)
    np.random.seed(5)
    aug = ScanAugmentation.draw(ds)
    np.random.seed(5)
    aug_o = oi.draw_augmentation(ds)
    d1 = oi.ingest_scan(sweeps[0]["raw"], sweeps[0]["labels"], ident, ds, None, pos_div=1.0)
    d2 = oi.ingest_scan(sweeps[1]["raw"], sweeps[1]["labels"], ident, ds, None, pos_div=1.0)
    for do_swap, do_paste in ((True, True), (False, True), (True, False), (False, False)):
        mix = PolarMixDraw(1, -2.3, -2.3 + np.pi, 0.9, 3.1, do_swap, do_paste)
        got = ingest_scan_polarmix(sweeps[0], sweeps[1], mix, lm, ds, aug)
        want = oi.polarmix(d1, d2, {"alpha": mix.alpha, "beta": mix.beta, "omega": [mix.omega0, mix.omega1],
                                    "do_swap": do_swap, "do_paste": do_paste}, INSTANCE_CLASSES[ds])
        wpos = torch.from_numpy(oi.training_augmentation(want["pos"].numpy(), aug_o)) / oi.POS_NORMALIZE
        assert got.pos.shape[0] == wpos.shape[0]
        assert torch.equal(got.y.cpu(), want["y"]) and torch.equal(got.x.cpu(), want["x"])
        assert torch.equal(got.curve_idxs.cpu(), want["curve_idxs"])
        assert _ulp_close(got.pos.cpu(), wpos, ulps=2, floor=0.05)      # two roundings (copy rotation, augmentation)
        cid = got.curve_idxs.cpu()
        assert int(cid[0]) == 0 and bool(torch.all((cid[1:] - cid[:-1]) >= 0)) and bool(torch.all((cid[1:] - cid[:-1]) <= 1))


@pytest.mark.gpu
def test_polarmix_degenerate_cases():
    """No instance points in the partner sweep, an empty yaw sector, and a sector holding everything."""
    from curvecloudnet_b200.ingest import LearningMap, PolarMixDraw, ingest_scan, polarmix
    from oracle import ccn_ingest as oi
    ident = {k: k for k in range(32)}
    lm = LearningMap(ident)
    (rec1, lab1), (rec2, lab2) = _pm_sweeps()
    none2 = np.zeros_like(lab2)                                         # class 0 is not an instance class
    for labels2, alpha, beta in ((none2, -2.0, 1.0), (lab2, 0.5, 0.5), (lab2, -4.0, 4.0)):
        d1 = ingest_scan(rec1, lab1, lm, "nuscenes", pos_div=1.0)
        d2 = ingest_scan(rec2, labels2, lm, "nuscenes", pos_div=1.0)
        o1 = oi.ingest_scan(rec1, lab1, ident, "nuscenes", None, pos_div=1.0)
        o2 = oi.ingest_scan(rec2, labels2, ident, "nuscenes", None, pos_div=1.0)
        mix = PolarMixDraw(0, alpha, beta, 0.4, 2.9, True, True)
        got = polarmix(d1, d2, mix, PM["instance_classes"])
        want = oi.polarmix(o1, o2, {"alpha": alpha, "beta": beta, "omega": [0.4, 2.9], "do_swap": True, "do_paste": True},
                           PM["instance_classes"])
        assert got.pos.shape[0] == want["pos"].shape[0]
        assert torch.equal(got.y.cpu(), want["y"]) and torch.equal(got.curve_idxs.cpu(), want["curve_idxs"])
        assert torch.equal(got.x.cpu(), want["x"]) and _ulp_close(got.pos.cpu(), want["pos"])


# ---------------------------------------------------------------------------------------------------------------
# A2D2 (audi_dataset.py:120-243): curves from a kNN graph per lidar.  tests/golden/a2d2.pt = SemA2D2's own get /
# _get_curves / _partition_curves_at_discontinuities / _training_augmentation on synthetic frames (CPU pin of the
# oracle: tests/test_oracle_golden.py::test_a2d2_curve_builder)
# ---------------------------------------------------------------------------------------------------------------
def _a2d2_frame(g):
    return {"points": g["points"].numpy(), "reflectance": g["reflectance"].numpy(), "lidar_id": g["lidar_id"].numpy(),
            "timestamp": g["timestamp"].numpy() + int(g["timestamp_base"])}


def test_a2d2_augmentation_draws_replay_the_reference_order():
    from curvecloudnet_b200.ingest import A2D2Augmentation
    from oracle import ccn_ingest as I
    np.random.seed(5)
    a = A2D2Augmentation.draw()
    np.random.seed(5)
    b = I.a2d2_draw_augmentation()
    assert (a.flip_y, a.scale) == (b["flip_y"], b["scale"])


@pytest.mark.gpu
@pytest.mark.parametrize("tag", ["full", "small", "ragged"])
def test_a2d2_ingest_matches_reference_golden(tag):
    """ccn_a2d2_curves + ccn_a2d2_finish + ccn_augment_points against the reference's own `get()`: the permutation, the
    curve ids, reflectance and positions bit-exact (validation and training split; no rotation in this augmentation, so
    nothing is left to a BLAS summation order)."""
    from curvecloudnet_b200.ingest import A2D2Augmentation, ingest_a2d2_frame
    g = load("a2d2.pt")[tag]
    frame = _a2d2_frame(g)
    n = frame["points"].shape[0]
    d = ingest_a2d2_frame(frame, labels=np.arange(n, dtype=np.int64))
    assert torch.equal(d.reorder.cpu(), g["val"]["reorder"]) and torch.equal(d.y.cpu(), g["val"]["reorder"])
    assert torch.equal(d.curve_idxs.cpu(), g["val"]["curve_idxs"])
    assert int(d.counts[0]) == int(g["val"]["curve_idxs"].max()) + 1 and int(d.counts[1]) == 0
    assert torch.equal(d.x.cpu(), g["val"]["x"]) and torch.equal(d.pos.cpu(), g["val"]["pos"])
    np.random.seed(int(g["train"]["seed"]))
    t = ingest_a2d2_frame(frame, augmentation=A2D2Augmentation.draw())
    assert torch.equal(t.pos.cpu(), g["train"]["pos"])
    assert torch.equal(t.curve_idxs, d.curve_idxs)


@pytest.mark.gpu
def test_a2d2_ingest_fresh_frames_match_oracle_and_errors():
    """Frames the fixture does not hold (other seeds, points in acquisition order, one sensor only, duplicate points, a
    handful of points) against the CPU oracle; a lidar id out of range is reported."""
    from curvecloudnet_b200.ingest import ingest_a2d2_frame
    from curvecloudnet_b200.synthetic import a2d2_frame
    from oracle import ccn_ingest as I
    frames = [a2d2_frame(7, azimuth_steps=400), a2d2_frame(8, azimuth_steps=300, shuffle=False)]
    one = a2d2_frame(9, azimuth_steps=200)
    keep = one["lidar_id"] == 3
    frames.append({k: np.ascontiguousarray(v[keep]) for k, v in one.items()})
    dup = a2d2_frame(10, azimuth_steps=120)
    dup["points"][5:40] = dup["points"][4]                       # coincident points: distance ties in the kNN graph
    dup["lidar_id"][5:40] = dup["lidar_id"][4]
    frames.append(dup)
    frames.append({k: np.ascontiguousarray(v[:3]) for k, v in one.items()})
    for f in frames:
        want = I.a2d2_ingest_frame(f)
        got = ingest_a2d2_frame(f)
        assert torch.equal(got.reorder.cpu(), want["reorder"]), f["points"].shape
        assert torch.equal(got.curve_idxs.cpu(), want["curve_idxs"])
        assert torch.equal(got.pos.cpu(), want["pos"]) and torch.equal(got.x.cpu(), want["x"])
    bad = {k: v.copy() for k, v in frames[0].items()}
    bad["lidar_id"][3] = 9
    with pytest.raises(ValueError, match="lidar id"):
        ingest_a2d2_frame(bad)
    empty = {k: v[:0] for k, v in frames[0].items()}
    e = ingest_a2d2_frame(empty)
    assert e.pos.shape == (0, 3) and e.curve_idxs.numel() == 0 and int(e.counts[0]) == 0


@pytest.mark.gpu
def test_a2d2_frame_feeds_the_model():
    """An ingested A2D2-shaped frame runs through the A2D2 step list (configs 'a2d2') and matches the oracle model fed with
    the oracle's ingest of the same frame."""
    from curvecloudnet_b200.configs import build_model
    from curvecloudnet_b200.ingest import ingest_a2d2_frame
    from curvecloudnet_b200.model import Batch, ModelBase
    from curvecloudnet_b200.synthetic import ScanBatch, a2d2_frame, randomize_model_
    from golden_utils import rel_err
    from oracle import ccn_ingest as I
    from oracle.ccn_model import ModelBase as OracleModel
    f = a2d2_frame(12, azimuth_steps=250)
    want_in = I.a2d2_ingest_frame(f)
    n = want_in["pos"].shape[0]
    ref = randomize_model_(build_model(OracleModel, "a2d2"), 1234).eval()
    torch.manual_seed(3)
    with torch.no_grad():
        want = ref(ScanBatch(want_in["x"], want_in["pos"], torch.zeros(n, dtype=torch.int64), want_in["curve_idxs"], None, None))
    d = ingest_a2d2_frame(f)
    model = build_model(ModelBase, "a2d2", precision="bf16x3")
    model.load_state_dict(ref.state_dict(), strict=True)
    model = model.cuda().eval()
    torch.manual_seed(3)
    with torch.no_grad():
        got = model(Batch.from_data_list([d]))
    assert rel_err(got.cpu(), want) <= 1e-4


# ---------------------------------------------------------------------------------------------------------------
# object scans: CurvesInMemoryDataset.get / pc_normalize (src/data/object_datasets.py:21-66); tests/golden/objects.pt holds
# pc_normalize's own output (compiled from source) and the shapenet-seg training transform around it
# ---------------------------------------------------------------------------------------------------------------
def test_object_augmentation_draws_replay_the_reference_order():
    import random
    from curvecloudnet_b200.ingest import ObjectAugmentation
    g = load("objects.pt")
    random.seed(21)
    torch.manual_seed(21)
    for b in range(3):
        a = ObjectAugmentation.draw()
        assert abs(a.scale - float(g["draws"][b, 0])) < 1e-15
        assert torch.allclose(torch.tensor(a.shift, dtype=torch.float64), g["draws"][b, 1:], atol=1e-9)


@pytest.mark.gpu
def test_object_scans_match_reference_golden():
    """ccn_object_normalize against pc_normalize (and the training transform): 1e-6 relative to the unit scale of the result -
    the means are f64 sums here, f32 pairwise sums upstream."""
    from curvecloudnet_b200.ingest import ObjectAugmentation, ingest_object_scans
    g = load("objects.pt")
    raw, batch = g["raw"].cuda(), g["batch"].cuda()
    val = ingest_object_scans(raw, batch)
    assert float((val.cpu() - g["val"]).abs().max()) <= 1e-6
    augs = [ObjectAugmentation(float(d[0]), tuple(float(v) for v in d[1:])) for d in g["draws"]]
    tr = ingest_object_scans(raw, batch, augs)
    assert float((tr.cpu() - g["train"]).abs().max()) <= 1e-6
    for b in range(3):                                       # what pc_normalize promises: centred, the farthest point at 1
        v = val[batch == b]
        assert float(v.mean(0).abs().max()) < 1e-6 and abs(float(v.norm(dim=1).max()) - 1.0) < 1e-6


# ---------------------------------------------------------------------------------------------------------------
# Kortx raw scans: tests/golden/kortx.pt = the output file of the REAL scanning_simulator/kortx/generate_kortx.py (main,
# fix_curve_idxs, cut_curves) run over a synthetic instance directory (oracle/make_golden.py kortx)
# ---------------------------------------------------------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("tag", ["table", "handheld"])
def test_kortx_scan_matches_the_reference_script(tag):
    from curvecloudnet_b200.ingest import ingest_kortx_scan
    g = load("kortx.pt")[tag]
    d = ingest_kortx_scan(g["points"].numpy(), g["raw"].numpy(), g["seg"].numpy(), npoints=2048, samples_per_scan=5,
                          capture_setting=g["setting"])
    assert d["pos"].dtype == torch.float64 and torch.equal(d["pos"].cpu(), g["pos"])          # f64 division: bit-exact
    assert torch.equal(d["curve_idxs"].cpu(), g["curve_idxs"]) and torch.equal(d["y"].cpu(), g["y"])
    c = d["all_curve_idxs"]
    assert int(c[0]) == 0 and bool((c[1:] - c[:-1] >= 0).all()) and bool((c[1:] - c[:-1] <= 1).all())
    with pytest.raises(ValueError, match="fewer than one window"):
        ingest_kortx_scan(g["points"].numpy()[:100], g["raw"].numpy()[:100], None)
