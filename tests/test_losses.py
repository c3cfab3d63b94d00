"""Loss / metrics epilogue (seg_loss = weighted NLL + Lovasz-softmax, fast_hist, per_class_iu, the TTA loop):
the oracle and the CUDA kernels against tests/golden/losses.pt, which oracle/make_golden.py froze from the REAL
reference sources (lovasz_losses.py imported as is; seg_loss / fast_hist_crop / per_class_iu executed from their files)."""
import pytest
import torch

from golden_utils import grad_projection, load, loss_case, rel_err

GOLD = load("losses.pt")


def _check(name, use_lovasz, loss, per, dlogits, tol):
    rec = GOLD[name]["lovasz" if use_lovasz else "nll"]
    assert abs(float(loss) - rec["loss"]) <= tol * abs(rec["loss"]), (float(loss), rec["loss"])
    assert rel_err(per[:256].cpu(), rec["perpnt_head"]) <= tol
    assert abs(float(per.double().sum()) - rec["perpnt_sum"]) <= tol * abs(rec["perpnt_sum"])
    gd = dlogits.detach().cpu().double().flatten()
    assert abs(float(gd.norm()) - rec["dnorm"]) <= 10 * tol * rec["dnorm"]
    assert abs(float((gd * grad_projection(name, gd.numel())).sum()) - rec["dproj"]) <= 10 * tol * rec["dnorm"]
    assert rel_err(dlogits[:256].cpu(), rec["dhead"]) <= 10 * tol
    if "dlogits" in rec:
        assert rel_err(dlogits.cpu(), rec["dlogits"]) <= 10 * tol


@pytest.mark.parametrize("use_lovasz", [False, True])
@pytest.mark.parametrize("name", list(GOLD))
def test_oracle_seg_loss_matches_reference(name, use_lovasz):
    from oracle import ccn_losses as L
    r = GOLD[name]
    logits, gt, w = loss_case(name, r["n"], r["C"], r["seed"])
    logits.requires_grad_()
    loss, per = L.seg_loss(logits, gt, ignore=0, use_lovasz=use_lovasz, ce_weights=w)
    loss.backward()
    _check(name, use_lovasz, loss, per.detach(), logits.grad, 1e-6)
    pred = torch.argmax(logits.detach(), dim=-1)
    hist = L.fast_hist_crop(pred, gt, list(range(r["C"] - 1)))
    assert torch.equal(hist, r["hist_crop"].to(hist.dtype))
    assert torch.allclose(L.per_class_iu(hist), r["iou"].double(), equal_nan=True)


@pytest.mark.gpu
@pytest.mark.parametrize("use_lovasz", [False, True])
@pytest.mark.parametrize("name", list(GOLD))
def test_cuda_seg_loss_matches_reference(name, use_lovasz):
    from curvecloudnet_b200 import losses as L
    r = GOLD[name]
    logits, gt, w = loss_case(name, r["n"], r["C"], r["seed"])
    logits = logits.cuda().requires_grad_()
    loss, per = L.seg_loss(logits, gt.cuda(), ignore=0, use_lovasz=use_lovasz, ce_weights=w.cuda())
    (loss * 1.0).backward()
    _check(name, use_lovasz, loss, per, logits.grad, 2e-5)
    hist = L.fast_hist_crop(torch.argmax(logits.detach(), dim=-1), gt.cuda(), list(range(r["C"] - 1)))
    assert torch.equal(hist.cpu(), r["hist_crop"].to(torch.int64))              # integer work: bit-exact
    assert torch.allclose(L.per_class_iu(hist).cpu(), r["iou"].double(), equal_nan=True)


@pytest.mark.gpu
def test_cuda_lovasz_edge_cases():
    from curvecloudnet_b200 import losses as L
    from oracle import ccn_losses as O
    for n, C, seed in ((1, 3, 0), (2, 2, 1), (4097, 5, 2), (300, 40, 3)):
        gen = torch.Generator().manual_seed(seed)
        logits = torch.randn(n, C, generator=gen)
        gt = torch.randint(0, C, (n,), generator=gen)
        for ignore in (0, -1):
            a = logits.clone().requires_grad_()
            keep = gt != ignore
            ref = O.lovasz_softmax_flat(torch.softmax(a, -1)[keep], gt[keep])
            ref = ref.sum() if ref.dim() else ref
            if a.requires_grad and ref.requires_grad:
                ref.backward()
            b = logits.clone().cuda().requires_grad_()
            got = L.lovasz_softmax_flat(b, gt.cuda(), ignore)
            got.backward()
            assert abs(float(got) - float(ref)) <= 2e-5 * max(abs(float(ref)), 1e-3), (n, C, ignore)
            want = a.grad if a.grad is not None else torch.zeros_like(a)
            assert float((b.grad.cpu() - want).abs().max()) <= 2e-5 * max(float(want.abs().max()), 1e-3)
    # everything ignored: zero loss, zero gradient
    b = torch.randn(50, 4).cuda().requires_grad_()
    got = L.lovasz_softmax_flat(b, torch.zeros(50, dtype=torch.int64).cuda(), 0)
    got.backward()
    assert float(got) == 0.0 and float(b.grad.abs().max()) == 0.0


@pytest.mark.gpu
def test_cuda_tta_forward_matches_oracle():
    """val()'s augmentation loop: 12 forwards (own random draws each), averaged, log_softmax, argmax."""
    from curvecloudnet_b200 import losses as L
    from curvecloudnet_b200.configs import build_model
    from curvecloudnet_b200.model import ModelBase
    from curvecloudnet_b200.synthetic import driving_scan_batch, randomize_model_
    from oracle import ccn_losses as O
    from oracle.ccn_model import ModelBase as OracleModel
    data = driving_scan_batch([7], subsample=1200)
    ref = build_model(OracleModel, "nuscenes")
    randomize_model_(ref, 1234)
    ref.eval()
    torch.manual_seed(3)
    want, want_pred = O.tta_forward(ref, data)
    model = build_model(ModelBase, "nuscenes", precision="fp32")
    model.load_state_dict(ref.state_dict(), strict=True)
    model = model.cuda().eval()
    torch.manual_seed(3)
    got, pred = L.tta_forward(model, data.to("cuda"))
    assert rel_err(got.cpu(), want) <= 1e-4
    agree = float((pred.cpu() == want_pred).float().mean())
    assert agree >= 0.999, agree


@pytest.mark.gpu
def test_cuda_tta_forward_stacked_matches_oracle():
    """tta_forward(group=4): four augmented copies per forward as one batch; same seed -> same draws as the oracle's
    stacked statement; and the stacked result agrees with the loop up to the sampling randomness."""
    from curvecloudnet_b200 import losses as L
    from curvecloudnet_b200.configs import build_model
    from curvecloudnet_b200.model import ModelBase
    from curvecloudnet_b200.synthetic import driving_scan_batch, randomize_model_
    from oracle import ccn_losses as O
    from oracle.ccn_model import ModelBase as OracleModel
    data = driving_scan_batch([8], subsample=900)
    ref = build_model(OracleModel, "nuscenes")
    randomize_model_(ref, 1234)
    ref.eval()
    torch.manual_seed(4)
    want, want_pred = O.tta_forward_stacked(ref, data, 4)
    model = build_model(ModelBase, "nuscenes", precision="fp32")
    model.load_state_dict(ref.state_dict(), strict=True)
    model = model.cuda().eval()
    dev = data.to("cuda")
    torch.manual_seed(4)
    got, pred = L.tta_forward(model, dev, group=4)
    assert rel_err(got.cpu(), want) <= 1e-4
    assert float((pred.cpu() == want_pred).float().mean()) >= 0.999
    assert torch.equal(dev.pos.cpu(), data.pos)                        # the input batch is left untouched
    # two-sample batch: copy g of sample b lands in sample g*B + b
    data2 = driving_scan_batch([8, 9], subsample=700)
    st = L.tta_stack(data2.to("cuda"), [(1.0, 1.0), (-1.05, 1.05)])
    n = data2.pos.shape[0]
    assert st.batch.cpu().tolist() == data2.batch.tolist() + (data2.batch + 2).tolist()
    assert torch.equal(st.pos[:n].cpu(), data2.pos)
    assert torch.equal(st.pos[n:, 2].cpu(), data2.pos[:, 2]) and torch.equal(st.curve_idxs[n:].cpu(), data2.curve_idxs)


@pytest.mark.gpu
def test_part_iou_and_object_epilogues():
    """Object-run epilogues: ccn_part_iou against the reference's per-object loop (shapenet_seg.py:88-99, oracle restatement;
    parts absent from an object -> absent_score 1, an object whose every point is wrong, first-maximum ties), the loss of
    :182-186, the category summary of :103-111, and the classification metrics of shapenet_classification.py:77-93 against
    sklearn's accuracy_score / balanced_accuracy_score."""
    import torch.nn.functional as F
    from sklearn import metrics
    from curvecloudnet_b200 import losses as L
    from oracle import ccn_losses as OL
    gen = torch.Generator().manual_seed(12)
    cats = [0, 10, 1, 15, 4, 10, 7, 3]
    sizes = [700, 650, 300, 2048, 512, 37, 1, 900]
    ptr = torch.tensor([0] + sizes).cumsum(0)
    n = int(ptr[-1])
    out = torch.randn(n, 50, generator=gen)
    y = torch.empty(n, dtype=torch.int64)
    for b, c in enumerate(cats):
        part = L.SHAPENET_SEG_CLASSES[L.SHAPENET_CATEGORY_NAMES[c]]
        lo, hi = int(ptr[b]), int(ptr[b + 1])
        y[lo:hi] = torch.tensor(part)[torch.randint(0, len(part) - (1 if b == 1 else 0), (hi - lo,), generator=gen)]
        if b % 2 == 0:                                   # make the prediction mostly right on every second object
            out[lo:hi].scatter_(1, y[lo:hi, None], 4.0 + torch.rand(hi - lo, 1, generator=gen))
    out[int(ptr[3]):int(ptr[3]) + 50, 47:50] = 0.25                      # equal scores: the first part wins
    lp = F.log_softmax(out, dim=-1)
    want = OL.part_iou_loop(lp, y, ptr, cats, L.SHAPENET_SEG_CLASSES, L.SHAPENET_CATEGORY_NAMES)
    got = L.part_iou(lp.cuda(), y.cuda(), ptr.cuda(), cats)
    assert got.dtype == torch.float32 and torch.allclose(got.cpu(), torch.tensor(want), rtol=0, atol=1e-7), (got.tolist(), want)
    loss, per = L.object_seg_loss(out.cuda(), y.cuda())
    ref = F.nll_loss(lp, y, reduction="none")
    assert abs(float(loss) - float(ref.mean())) <= 1e-5 * float(ref.mean()) and torch.allclose(per.cpu(), ref, rtol=1e-5, atol=1e-6)
    summ = L.category_iou_summary(got, torch.tensor(cats).cuda())
    ious, cat = torch.tensor(want), torch.tensor(cats)
    per_cat = torch.zeros(16).scatter_add_(0, cat, ious) / torch.zeros(16).scatter_add_(0, cat, torch.ones(8)).clamp(min=1)
    assert torch.allclose(summ["per_category"].cpu(), per_cat, atol=1e-6) and abs(summ["inst_iou"] - float(ious.mean())) < 1e-6
    assert abs(summ["mean_iou"] - float(per_cat.mean())) < 1e-6 and abs(summ["median_iou"] - float(per_cat.median())) < 1e-6
    # classification
    logits = torch.randn(500, 16, generator=gen)
    labels = torch.randint(0, 13, (500,), generator=gen)                 # three classes never occur
    logits.scatter_(1, labels[:, None], 1.5, reduce="add")
    clp = F.log_softmax(logits, dim=-1)
    m = L.classification_metrics(clp.cuda(), labels.cuda(), 16)
    pred = clp.argmax(dim=-1)
    assert abs(m["accuracy"] - metrics.accuracy_score(labels.numpy(), pred.numpy())) < 1e-12
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        assert abs(m["class_accuracy"] - metrics.balanced_accuracy_score(labels.numpy(), pred.numpy())) < 1e-12
    assert int(m["hist"].sum()) == 500
