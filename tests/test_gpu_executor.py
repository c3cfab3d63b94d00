"""The native step executor (csrc/executor.cu, ccn_forward_run: ONE C-ABI call per forward) against the python executor of
curvecloudnet_b200/model.py (reference src/models/base.py:133-175): it issues the same kernel launches in the same order,
so the logits must be BIT-IDENTICAL under the same torch seed (same random draws); parity of the python executor against
the oracle / the reference-generated goldens is established by tests/test_gpu_model.py and test_gpu_fullsize_oracle.py."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _models(name, precision):
    from curvecloudnet_b200.configs import build_model
    from curvecloudnet_b200.model import ModelBase
    from curvecloudnet_b200.synthetic import randomize_model_
    a = randomize_model_(build_model(ModelBase, name, precision=precision), 1234).cuda().eval()
    b = randomize_model_(build_model(ModelBase, name, precision=precision, native_executor=False), 1234).cuda().eval()
    return a, b


@pytest.mark.parametrize("precision", ["bf16x3", "bf16"])
@pytest.mark.parametrize("name,seeds,sub", [("nuscenes", [3], 1500), ("nuscenes", [0, 1], 3000), ("nuscenes", [0], None),
                                            ("kitti", [2], 2500), ("kitti", [2, 3], None)])
def test_native_executor_equals_python_executor(name, seeds, sub, precision):
    from curvecloudnet_b200.synthetic import driving_scan_batch
    data = driving_scan_batch(seeds, kind=name, subsample=sub).to("cuda")
    native_model, py_model = _models(name, precision)
    with torch.no_grad():
        for rep in range(2):                       # the second pass reuses the executor's arena
            torch.manual_seed(7 + rep)
            got = native_model(data)
            torch.manual_seed(7 + rep)
            want = py_model(data)
            assert native_model.__dict__["_native"][1] is not None, "the native executor did not take this model"
            assert got.shape == want.shape and torch.equal(got, want), "native and python executors differ (pass %d)" % rep


def test_native_executor_grows_its_arena_and_replays_the_draws():
    from curvecloudnet_b200.synthetic import driving_scan_batch
    data = driving_scan_batch([4], subsample=4000).to("cuda")
    native_model, py_model = _models("nuscenes", "bf16x3")
    with torch.no_grad():
        torch.manual_seed(1)
        native_model(data)                          # builds the executor
        ex = native_model.__dict__["_native"][1]
        ex.arena = torch.empty(1 << 20, dtype=torch.uint8, device="cuda")      # far too small: CCN_ENOMEM -> grow -> run again
        torch.manual_seed(2)
        got = native_model(data)
        torch.manual_seed(2)
        want = py_model(data)
    assert ex.arena.numel() > (1 << 20)
    assert torch.equal(got, want)


def test_native_executor_reports_unsorted_input_like_the_reference():
    """point_ops.py:31,49 assert sorted ids; the native executor raises the same AssertionError."""
    from curvecloudnet_b200.synthetic import driving_scan_batch
    data = driving_scan_batch([5], subsample=1200).to("cuda")
    native_model, _ = _models("nuscenes", "bf16x3")
    data.batch = data.batch.clone()
    data.batch[10] = 1
    with torch.no_grad(), pytest.raises(AssertionError):
        native_model(data)


def test_other_configs_fall_back_to_the_python_executor():
    from curvecloudnet_b200.synthetic import object_scan_batch
    data = object_scan_batch([0, 1]).to("cuda")
    native_model, py_model = _models("shapenet-class", "bf16x3")
    with torch.no_grad():
        torch.manual_seed(3)
        got = native_model(data)
        torch.manual_seed(3)
        want = py_model(data)
    assert native_model.__dict__["_native"][1] is None
    assert torch.equal(got, want)


@pytest.mark.parametrize("native", [True, False])
def test_fused_edge_step_setting_does_not_change_the_logits(native):
    """ops.FUSED_EDGE_STEPS: "auto" (fused kernels where they are the faster path), True (wherever they cover the shapes),
    False (unfused launches) - the same arithmetic in the same order, bit-identical logits on both executors."""
    from curvecloudnet_b200 import ops
    from curvecloudnet_b200.configs import build_model
    from curvecloudnet_b200.model import ModelBase
    from curvecloudnet_b200.synthetic import driving_scan_batch, randomize_model_
    data = driving_scan_batch([6], subsample=5000).to("cuda")
    outs = {}
    prev = ops.FUSED_EDGE_STEPS
    try:
        for setting in ("auto", True, False):
            ops.FUSED_EDGE_STEPS = setting
            model = randomize_model_(build_model(ModelBase, "nuscenes", precision="bf16x3", native_executor=native), 1234).cuda().eval()
            with torch.no_grad():
                torch.manual_seed(9)
                outs[setting] = model(data)
    finally:
        ops.FUSED_EDGE_STEPS = prev
    assert torch.equal(outs["auto"], outs[True]) and torch.equal(outs["auto"], outs[False])


@pytest.mark.parametrize("precision", ["bf16x3", "bf16"])
@pytest.mark.parametrize("native", [True, False])
def test_mlp_chain_setting_does_not_change_the_logits(native, precision):
    """ops.CHAIN_MLP_LAYERS: "auto" (ccn_mlp_chain where ops.mlp_chain_profitable says so), True (every per-point MLP the
    kernel covers), False (one launch per Linear) - bit-identical logits on both executors, and the chain launches happen."""
    from curvecloudnet_b200 import native as nat, ops
    from curvecloudnet_b200.configs import build_model
    from curvecloudnet_b200.model import ModelBase
    from curvecloudnet_b200.synthetic import driving_scan_batch, randomize_model_
    data = driving_scan_batch([6], subsample=5000).to("cuda")
    outs, chained = {}, {}
    prev, real_call = ops.CHAIN_MLP_LAYERS, ops.call
    try:
        for setting in ("auto", True, False):
            ops.CHAIN_MLP_LAYERS = setting
            model = randomize_model_(build_model(ModelBase, "nuscenes", precision=precision, native_executor=native), 1234).cuda().eval()
            count = [0]

            def counting(name, *a):
                count[0] += name == "ccn_mlp_chain"
                return real_call(name, *a)
            ops.call = counting
            with torch.no_grad():
                torch.manual_seed(9)
                outs[setting] = model(data)
            ops.call = real_call
            chained[setting] = count[0]
    finally:
        ops.CHAIN_MLP_LAYERS, ops.call = prev, real_call
    assert torch.equal(outs["auto"], outs[True]) and torch.equal(outs["auto"], outs[False])
    if not native:           # the python executor's launches go through ops.call
        assert chained[False] == 0 and chained["auto"] >= 2 and chained[True] > chained["auto"]


@pytest.mark.parametrize("native", [True, False])
def test_dependent_launch_setting_does_not_change_the_logits(native):
    """ccn_set_dependent_launch: every kernel launched with the programmatic-stream-serialization attribute (and waiting on
    its predecessor before its first global access) against plain stream-ordered launches - bit-identical logits, several
    forwards in a row so that a missing wait would race with the previous kernel's tail."""
    from curvecloudnet_b200 import native as nat
    from curvecloudnet_b200.configs import build_model
    from curvecloudnet_b200.model import ModelBase
    from curvecloudnet_b200.synthetic import driving_scan_batch, randomize_model_
    data = driving_scan_batch([4]).to("cuda")
    model = randomize_model_(build_model(ModelBase, "nuscenes", precision="bf16x3", native_executor=native), 1234).cuda().eval()
    prev = nat.set_dependent_launch(-1)
    outs = {}
    try:
        for on in (1, 0, 1):
            nat.set_dependent_launch(on)
            assert nat.set_dependent_launch(-1) == on
            for rep in range(3):
                with torch.no_grad():
                    torch.manual_seed(5)
                    y = model(data)
                outs.setdefault(on, y)
                assert torch.equal(y, outs[on]) and torch.equal(y, outs[1])
    finally:
        nat.set_dependent_launch(prev)


@pytest.mark.parametrize("npl", [2, 1])
@pytest.mark.parametrize("cx,with_xyz,with_diff,pad", [(128, True, True, 6), (1, True, True, 2), (0, True, False, 0), (21, False, True, 4)])
def test_curve_feat_planes_equals_prep_gather_pack(cx, with_xyz, with_diff, pad, npl):
    """ccn_curve_feat_planes (the native executor's one-pass signal planes of a curve-conv step) against the three calls it
    replaces - ccn_curve_feat_prep -> ccn_gather_rows_i32 (padded rows) -> ccn_pack_planes: bit-identical planes."""
    from curvecloudnet_b200 import ops
    from curvecloudnet_b200.native import call
    from curvecloudnet_b200.synthetic import driving_scan_batch
    data = driving_scan_batch([5, 6], subsample=4000).to("cuda")
    topo = ops.curve_csr(data.batch, data.curve_idxs)
    n = data.pos.shape[0]
    gen = torch.Generator().manual_seed(cx)
    x = torch.randn(n, cx, generator=gen).cuda() if cx else None
    feats = ops.curve_feat_prep(x, data.pos, with_xyz, with_diff, topo.curve_ord)
    row2pt = ops.curve_row_map(topo, pad)
    want = ops.pack_planes([ops.gather_rows_i32(feats, row2pt)], npl)
    rows = row2pt.shape[0]
    got = ops.alloc_planes(rows, feats.shape[1], npl, data.pos)
    got.buf.fill_(7.0)
    call("ccn_curve_feat_planes", x if x is not None else 0, x.stride(0) if x is not None else 0, cx, data.pos.contiguous(), int(with_xyz),
         int(with_diff), topo.curve_ord, n, row2pt, rows, got.buf, got.kpad, got.plane_stride, npl)
    assert rows > n or pad == 0
    assert torch.equal(got.buf[:, :rows], want.buf[:, :rows])


@pytest.mark.parametrize("name,seeds", [("nuscenes", [0]), ("kitti", [2, 3])])
def test_lookahead_levels_do_not_change_the_logits(name, seeds, monkeypatch):
    """CCN_EXEC_LOOKAHEAD = 0 (one stream, step order) / 1 (side + geometry streams) / 2 (default: also voxel pre-samples
    collected at step boundaries, the first draw taken at entry, curve statistics ahead): the same kernels with the same
    arguments and the same draws in the same order, whatever runs beside what - bit-identical logits, three forwards each
    (whether a voxel read-back has landed at a step boundary is a matter of timing: both branches occur)."""
    from curvecloudnet_b200.configs import build_model
    from curvecloudnet_b200.model import ModelBase
    from curvecloudnet_b200.synthetic import driving_scan_batch, randomize_model_
    data = driving_scan_batch(seeds, kind=name).to("cuda")
    outs = {}
    for level in ("0", "1", "2"):
        monkeypatch.setenv("CCN_EXEC_LOOKAHEAD", level)
        model = randomize_model_(build_model(ModelBase, name, precision="bf16x3"), 1234).cuda().eval()
        for rep in range(3):
            with torch.no_grad():
                torch.manual_seed(21)
                y = model(data)
            assert model.__dict__["_native"][1] is not None
            outs.setdefault(level, y)
            assert torch.equal(y, outs[level]) and torch.equal(y, outs["0"]), "look-ahead level %s, forward %d" % (level, rep)
