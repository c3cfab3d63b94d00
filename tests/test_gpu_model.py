"""End-to-end GPU parity: ModelBase forward on the sm_100a kernels vs the logits the REAL reference sources
produced (tests/golden/models.pt), same weights, same inputs, same random draws (torch seed 77)."""
import pytest
import torch

from golden_utils import load, rel_err

pytestmark = pytest.mark.gpu
MODELS = load("models.pt")
MODELS.update(load("models_extra.pt"))        # a2d2, kortx, kitti - same generator, the REAL reference sources


PRECISION_TOL = {"fp32": 1e-4, "bf16x3": 1e-4, "bf16": 2e-2}     # BASELINE.json north_star tolerances


@pytest.mark.parametrize("precision", list(PRECISION_TOL))
@pytest.mark.parametrize("name", list(MODELS))
def test_model_logits_match_reference(name, precision):
    from curvecloudnet_b200.configs import build_model
    from curvecloudnet_b200.model import Batch, ModelBase
    from curvecloudnet_b200.synthetic import randomize_model_
    g = MODELS[name]
    model = build_model(ModelBase, name, precision=precision)
    randomize_model_(model, g["weight_seed"])
    model = model.cuda().eval()
    kw = dict(x=g["x"], pos=g["pos"], batch=g["batch"], curve_idxs=g["curve"])
    if g["labels"] is not None:
        kw["labels"] = g["labels"]
    data = Batch(**kw).to("cuda")
    torch.manual_seed(g["rng_seed"])
    with torch.no_grad():
        logits = model(data)
    assert logits.shape == g["logits"].shape
    err = rel_err(logits.cpu(), g["logits"])
    print("%s %s rel err %.3e" % (name, precision, err))
    assert err <= PRECISION_TOL[precision], "%s logits off by %.3e relative" % (precision, err)


def test_state_dict_roundtrip_with_oracle():
    """Checkpoint compatibility: the oracle's (= reference's) state_dict loads with strict=True."""
    from curvecloudnet_b200.configs import build_model
    from curvecloudnet_b200.model import ModelBase
    from oracle.ccn_model import ModelBase as OracleModel
    src = build_model(OracleModel, "nuscenes")
    dst = build_model(ModelBase, "nuscenes")
    dst.load_state_dict(src.state_dict(), strict=True)


@pytest.mark.parametrize("native_executor", [True, False])
@pytest.mark.parametrize("n_points", [400, 120, 40])
@pytest.mark.parametrize("name", ["nuscenes", "kitti"])
def test_tiny_scans_match_oracle(name, n_points, native_executor):
    """Scans far below the sizes the levels are tuned for (400 / 120 / 40 points: neighbourhoods larger than the cloud,
    sampling levels that shrink to a handful of points, curves of one or two points): the reference's code path still
    answers there (the oracle runs it on the CPU), so ours has to as well - same logits inside the north-star tolerance,
    through both executors."""
    from curvecloudnet_b200.configs import build_model
    from curvecloudnet_b200.model import ModelBase
    from curvecloudnet_b200.synthetic import driving_scan_batch, randomize_model_
    from oracle.ccn_model import ModelBase as OracleModel
    data = driving_scan_batch([3], kind=name, subsample=n_points)
    ref = randomize_model_(build_model(OracleModel, name), 1234).eval()
    torch.manual_seed(5)
    with torch.no_grad():
        want = ref(data)
    model = build_model(ModelBase, name, precision="bf16x3", native_executor=native_executor)
    model.load_state_dict(ref.state_dict(), strict=True)
    model = model.cuda().eval()
    torch.manual_seed(5)
    with torch.no_grad():
        got = model(data.to("cuda"))
    assert got.shape == want.shape and bool(torch.isfinite(got).all())
    err = rel_err(got.cpu(), want)
    print("%s %d points native=%s rel err %.3e" % (name, n_points, native_executor, err))
    assert err <= 1e-4


@pytest.mark.parametrize("name", ["nuscenes", "shapenet-seg", "a2d2"])
def test_inference_against_f64_oracle(name):
    """How far each precision mode is from an (almost) exact answer: the oracle with its feature arithmetic in f64 (same
    weights, positions and index decisions).  The f32 oracle - i.e. the reference's own arithmetic - is ~1e-6 away from it;
    fp32 must sit at the same distance (3x), bf16x3 inside the 1e-4 north-star budget, bf16 inside 2e-2 - measured against the
    exact answer, not against another f32 run."""
    from curvecloudnet_b200.configs import build_model
    from curvecloudnet_b200.model import ModelBase
    from curvecloudnet_b200.synthetic import driving_scan_batch, object_scan_batch, randomize_model_
    from oracle.ccn_model import ModelBase as OracleModel
    data = driving_scan_batch([4], subsample=2500) if name in ("nuscenes", "a2d2") else object_scan_batch([5, 6])
    ref = randomize_model_(build_model(OracleModel, name), 1234).eval()
    state = {k: v.clone() for k, v in ref.state_dict().items()}
    outs = {}
    for dt in (torch.float32, torch.float64):
        torch.manual_seed(9)
        with torch.no_grad():
            outs[dt] = ref.to(dt)(data).float()
    truth = outs[torch.float64]
    base = rel_err(outs[torch.float32], truth)
    for precision, bound in (("fp32", max(3 * base, 2e-6)), ("bf16x3", 1e-4), ("bf16", 2e-2)):
        model = build_model(ModelBase, name, precision=precision)
        model.load_state_dict(state, strict=True)
        model = model.cuda().eval()
        torch.manual_seed(9)
        with torch.no_grad():
            got = model(data.to("cuda"))
        err = rel_err(got.cpu(), truth)
        print("%s %s vs f64 oracle: %.3e (f32 oracle %.3e)" % (name, precision, err, base))
        assert err <= bound, (precision, err, bound)


@pytest.mark.parametrize("native_executor", [True, False])
def test_mixed_size_batch_matches_oracle(native_executor):
    """One batch holding a 6000-point scan, a 40-point scan and a 900-point scan (levels of the small scans shrink to a few
    points while the large one is still dense; per-sample pointers, FPS counts and neighbour tables all become ragged):
    logits against the oracle through both executors."""
    from curvecloudnet_b200.configs import build_model
    from curvecloudnet_b200.model import ModelBase
    from curvecloudnet_b200.synthetic import _collate, driving_scan_sample, randomize_model_
    from oracle.ccn_model import ModelBase as OracleModel
    data = _collate([driving_scan_sample(5, subsample=6000), driving_scan_sample(6, subsample=40),
                     driving_scan_sample(7, subsample=900)], with_x=True)
    ref = randomize_model_(build_model(OracleModel, "nuscenes"), 1234).eval()
    torch.manual_seed(6)
    with torch.no_grad():
        want = ref(data)
    model = build_model(ModelBase, "nuscenes", precision="bf16x3", native_executor=native_executor)
    model.load_state_dict(ref.state_dict(), strict=True)
    model = model.cuda().eval()
    torch.manual_seed(6)
    with torch.no_grad():
        got = model(data.to("cuda"))
    err = rel_err(got.cpu(), want)
    print("mixed batch native=%s rel err %.3e" % (native_executor, err))
    assert got.shape == want.shape and err <= 1e-4


@pytest.mark.parametrize("native_executor", [True, False])
@pytest.mark.parametrize("structure", ["singletons", "one_curve"])
def test_degenerate_curve_structures_match_oracle(structure, native_executor):
    """Curve ids at the two extremes - every point its own curve (no neighbour along any curve: the convolution sees pads only,
    CurveFPS keeps every point, the 1-D groupings are self loops) and one single curve through the whole scan - against the
    oracle, both executors."""
    from curvecloudnet_b200.configs import build_model
    from curvecloudnet_b200.model import ModelBase
    from curvecloudnet_b200.synthetic import driving_scan_batch, randomize_model_
    from oracle.ccn_model import ModelBase as OracleModel
    data = driving_scan_batch([8], subsample=800)
    n = data.pos.shape[0]
    data.curve_idxs = torch.arange(n) if structure == "singletons" else torch.zeros(n, dtype=torch.int64)
    ref = randomize_model_(build_model(OracleModel, "nuscenes"), 1234).eval()
    torch.manual_seed(1)
    with torch.no_grad():
        want = ref(data)
    model = build_model(ModelBase, "nuscenes", precision="bf16x3", native_executor=native_executor)
    model.load_state_dict(ref.state_dict(), strict=True)
    model = model.cuda().eval()
    torch.manual_seed(1)
    with torch.no_grad():
        got = model(data.to("cuda"))
    err = rel_err(got.cpu(), want)
    print("%s native=%s rel err %.3e" % (structure, native_executor, err))
    assert err <= 1e-4
