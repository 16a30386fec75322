"""GPU parity: K5 (multiclass curves) — the conditional row softmax must return torch's bits."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16])
def test_softmax_rows_bits_equal_torch(dtype):
    """Every lane layout of torch's persistent warp softmax: C from 1 to 1024 (all powers of two, their
    neighbours, and a spread in between), ragged row counts, large-magnitude logits, -inf entries."""
    from torchmetrics_forked_b200 import kernels

    g = torch.Generator(device="cuda").manual_seed(5)
    widths = sorted({1, 2, 3, 4, 5, 7, 8, 9, 15, 16, 17, 31, 32, 33, 63, 64, 65, 100, 127, 128, 129, 150, 255, 256,
                     257, 500, 511, 512, 513, 777, 1000, 1023, 1024})
    for C in widths:
        rows = 1237 if C <= 128 else 301
        x = (torch.randn(rows, C, generator=g, device="cuda") * 6).to(dtype)
        x[0, 0] = 40.0
        if C > 2:
            x[1, C // 2] = float("-inf")
        got = kernels.softmax_normalize(x)
        assert torch.equal(got, x.softmax(1)), (C, dtype)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_softmax_rows_wide_dynamic_range(dtype):
    """Logits spread over +-150: the exp values run through the whole range from 1 down to denormals and zero, i.e.
    through both branches of the shared-divisor division (three FFMAs above 2^-80, the full IEEE division below)."""
    from torchmetrics_forked_b200 import kernels

    g = torch.Generator(device="cuda").manual_seed(11)
    for C in (2, 3, 10, 31, 33, 100, 257, 1000):
        x = (torch.randn(20011, C, generator=g, device="cuda") * 35).to(dtype)
        x[::7] *= 0.02  # rows with close scores in between
        got = kernels.softmax_normalize(x)
        want = x.softmax(1)
        assert torch.equal(got, want), (C, dtype, int((got != want).sum()))
        if dtype == torch.float32:
            tiny = (want > 0) & (want < 2.0 ** -80)
            assert C == 2 or bool(tiny.any())  # the slow branch was really taken


def test_softmax_normalize_decision_and_special_rows():
    from torchmetrics_forked_b200 import kernels

    g = torch.Generator(device="cuda").manual_seed(6)
    probs = torch.rand(999, 10, generator=g, device="cuda")
    out = kernels.softmax_normalize(probs)
    assert torch.equal(out, probs) and out.data_ptr() != probs.data_ptr()  # all inside [0,1]: a copy
    spiked = probs.clone()
    spiked[500, 3] = 1.0001
    assert torch.equal(kernels.softmax_normalize(spiked), spiked.softmax(1))  # one value flips every row (H1)
    # ... unless its row is ignored
    target = torch.zeros(999, dtype=torch.long, device="cuda")
    target[500] = -1
    assert torch.equal(kernels.softmax_normalize(spiked, target, -1), spiked)
    assert torch.equal(kernels.softmax_normalize(spiked, target, -2), spiked.softmax(1))
    # NaN and +inf rows come out all-NaN exactly like torch's
    weird = torch.randn(64, 33, generator=g, device="cuda")
    weird[3, 5] = float("nan")
    weird[7, 0] = float("inf")
    got, want = kernels.softmax_normalize(weird), weird.softmax(1)
    assert torch.equal(got.isnan(), want.isnan()) and torch.equal(got.nan_to_num(), want.nan_to_num())
    # wide rows keep torch's kernel; empty input
    wide = torch.randn(17, 3000, generator=g, device="cuda")
    assert torch.equal(kernels.softmax_normalize(wide), wide.softmax(1))
    assert kernels.softmax_normalize(probs[:0]).shape == (0, 10)


# ------------------------------------------------------------------------------------------ binned states
from conftest import golden_names, ignore_of, load_golden  # noqa: E402

import hotpath as O  # noqa: E402


def cuda_softmax(x: np.ndarray) -> np.ndarray:
    return torch.from_numpy(np.ascontiguousarray(x, dtype=np.float32)).cuda().softmax(1).cpu().numpy()


def _cuda(x):
    return torch.from_numpy(np.ascontiguousarray(x)).cuda()


@pytest.mark.parametrize("name", golden_names("f2_mccurve_"))
def test_multiclass_curve_golden(name):
    """Binned [T, C, 2, 2] / micro [T, 2, 2] states bit for bit against the oracle fed with the DEVICE softmax (H5:
    CPU and CUDA softmax differ in low bits), and against the reference's own states where no softmax is involved;
    exact-mode state, curves and the scalar metrics against the reference's golden values."""
    from torchmetrics_forked_b200 import classification as C
    from torchmetrics_forked_b200 import functional as F
    from torchmetrics_forked_b200 import kernels

    c = load_golden(name)
    nC, ign = int(c["num_classes"]), ignore_of(c)
    thr = torch.from_numpy(c["thresholds"])
    preds, target = _cuda(c["preds"]), _cuda(c["target"])
    softmaxed = "logits" in name or name.endswith(("multidim", "wide"))
    for micro in (False, True):
        m = C.MulticlassPrecisionRecallCurve(nC, thresholds=thr, average="micro" if micro else None,
                                             ignore_index=ign).cuda()
        h = preds.shape[0] // 3
        m.update(preds[:h], target[:h])
        m.update(preds[h:], target[h:])
        # two updates decide the softmax independently, like two reference updates: build the oracle the same way
        want = sum(O.multiclass_curve_binned(c["preds"][s], c["target"][s], nC, c["thresholds"], ign, cuda_softmax, micro)
                   for s in (slice(0, h), slice(h, None)))
        assert np.array_equal(m.confmat.cpu().numpy(), want), (name, micro)
    got = F.multiclass_precision_recall_curve(preds, target, nC, thresholds=thr, ignore_index=ign)
    state = O.multiclass_curve_binned(c["preds"], c["target"], nC, c["thresholds"], ign, cuda_softmax)
    if not softmaxed:
        assert np.array_equal(state, c["state"])
    tps, fps, fns = state[:, :, 1, 1], state[:, :, 0, 1], state[:, :, 1, 0]
    with np.errstate(invalid="ignore", divide="ignore"):
        prec = np.where(tps + fps == 0, 0.0, tps / np.maximum(tps + fps, 1))
    np.testing.assert_allclose(got[0].cpu().numpy()[:, :-1], prec.T, rtol=1e-6)
    # exact mode
    e = C.MulticlassPrecisionRecallCurve(nC, thresholds=None, ignore_index=ign).cuda()
    e.update(preds, target)
    want_probs, want_t = O.multiclass_rows(c["preds"], c["target"], nC, ign, cuda_softmax)
    assert np.array_equal(e.preds[0].cpu().numpy(), want_probs) and np.array_equal(e.target[0].cpu().numpy(), want_t)
    fps_l, tps_l, thr_l = kernels.clf_curves(e.preds[0], e.target[0], None, class_targets=True)
    for k, (wf, wt, wh) in enumerate(O.multiclass_clf_curves(want_probs, want_t)):
        assert np.array_equal(fps_l[k].cpu().numpy(), wf) and np.array_equal(tps_l[k].cpu().numpy(), wt)
        assert np.array_equal(thr_l[k].cpu().numpy(), wh)
        if not softmaxed:
            assert np.array_equal(wf, c[f"fps_{k}"]) and np.array_equal(wh, c[f"thr_{k}"])
    tol = dict(rtol=2e-5, atol=1e-6) if softmaxed else dict(rtol=1e-6, atol=1e-7)
    for avg in ("macro", "weighted", "none"):
        for T_, tag in ((None, "exact"), (len(thr), "binned")):
            a = F.multiclass_auroc(preds, target, nC, average=avg, thresholds=T_, ignore_index=ign)
            np.testing.assert_allclose(a.cpu().numpy(), c[f"auroc_{avg}_{tag}"], **tol)
            a = F.multiclass_average_precision(preds, target, nC, average=avg, thresholds=T_, ignore_index=ign)
            np.testing.assert_allclose(a.cpu().numpy(), c[f"ap_{avg}_{tag}"], **tol)
    au = C.AUROC(task="multiclass", num_classes=nC, average="macro", thresholds=None, ignore_index=ign).cuda()
    ap = C.MulticlassAveragePrecision(nC, average="weighted", thresholds=None, ignore_index=ign).cuda()
    for s in (slice(0, 100), slice(100, None)):
        au.update(preds[s], target[s])
        ap.update(preds[s], target[s])
    if not softmaxed:  # per-update softmax decisions coincide only when no update needs one
        np.testing.assert_allclose(float(au.compute()), c["auroc_macro_exact"], **tol)
        np.testing.assert_allclose(float(ap.compute()), c["ap_weighted_exact"], **tol)
    if not softmaxed:
        for stem, kw in (("recall_at_fixed_precision", "min_precision"), ("precision_at_fixed_recall", "min_recall"),
                         ("specificity_at_sensitivity", "min_sensitivity")):
            for T_, tag in ((None, "exact"), (len(thr), "binned")):
                a, b = getattr(F, f"multiclass_{stem}")(preds, target, nC, **{kw: 0.4}, thresholds=T_, ignore_index=ign)
                np.testing.assert_allclose(a.cpu().numpy(), c[f"{stem}_{tag}"][0], rtol=1e-6, atol=1e-7)
                np.testing.assert_allclose(b.cpu().numpy(), c[f"{stem}_{tag}"][1], rtol=1e-6, atol=1e-7)
        for avg in ("macro", "micro"):
            pr = F.multiclass_precision_recall_curve(preds, target, nC, thresholds=len(thr), average=avg, ignore_index=ign)
            roc = C.MulticlassROC(nC, thresholds=len(thr), average=avg, ignore_index=ign).cuda()
            roc.update(preds, target)
            for k, v in zip(("p", "r", "t"), pr):
                np.testing.assert_allclose(v.cpu().numpy(), c[f"prc_{avg}_{k}"], rtol=1e-6, atol=1e-7)
            for k, v in zip(("f", "t", "h"), roc.compute()):
                np.testing.assert_allclose(v.cpu().numpy(), c[f"roc_{avg}_{k}"], rtol=1e-6, atol=1e-7)


@pytest.mark.parametrize("C", [2, 3, 10, 16, 17, 31, 32, 33, 100, 1000])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_multiclass_curve_fused_kernel_random(C, dtype):
    """k2_mc_rows over every lane layout, ragged rows, ignore_index, uneven thresholds and 16-bit scores, against
    the oracle with the device softmax; C = 1000, T = 200 does not fit one SM's shared memory (global atomics)."""
    from torchmetrics_forked_b200 import classification as M

    g = torch.Generator().manual_seed(400 + C)
    n = 4099 if C <= 100 else 517
    for thr in (torch.linspace(0, 1, 200), torch.sort(torch.rand(23, generator=g)).values):
        preds = (torch.randn(n, C, generator=g) * 3).to(dtype)
        target = torch.randint(C, (n,), generator=g)
        target[torch.rand(n, generator=g) < 0.1] = -1
        m = M.MulticlassPrecisionRecallCurve(C, thresholds=thr, ignore_index=-1).cuda()
        m.update(preds.cuda(), target.cuda())
        sm = lambda a: torch.from_numpy(a).to(dtype).cuda().softmax(1).float().cpu().numpy()  # noqa: E731
        # 16-bit scores: above 1e6 comparisons the reference loops over thresholds and `preds >= thresholds[i]`
        # (0-d threshold) is evaluated in the dtype of preds, i.e. against the ROUNDED threshold (:472, :518)
        # The reference switches on the number of rows it KEPT after the ignore filter, and so does this
        # implementation (round 2).  C = 16 with 4099 rows sits between the kept and the unfiltered count
        # (3689 * 256 < 10^6 < 4099 * 256).
        kept = int((target != -1).sum())
        thr_seen = thr.to(dtype).float() if dtype != torch.float32 and kept * C * C > 1_000_000 else thr
        want = O.multiclass_curve_binned(preds.float().numpy(), target.numpy(), C, thr_seen.numpy(), -1, sm)
        assert np.array_equal(m.confmat.cpu().numpy(), want), (C, dtype, len(thr))
    # probabilities: pass A only; out-of-range target is reported
    probs = torch.rand(300, C, generator=g).softmax(1)
    m = M.MulticlassPrecisionRecallCurve(C, thresholds=50).cuda()
    tgt = torch.randint(C, (300,), generator=g)
    m.update(probs.cuda(), tgt.cuda())
    want = O.multiclass_curve_binned(probs.numpy(), tgt.numpy(), C, torch.linspace(0, 1, 50).numpy())
    assert np.array_equal(m.confmat.cpu().numpy(), want)
    bad = tgt.clone()
    bad[5] = C
    with pytest.raises(RuntimeError, match="target"):
        M.MulticlassPrecisionRecallCurve(C, thresholds=50).cuda().update(probs.cuda(), bad.cuda())


def test_multiclass_curve_wide_rows_fallback():
    """C > 1024: torch's softmax behind the device-side decision, one-hot binning through the multilabel kernels."""
    from torchmetrics_forked_b200 import classification as M

    g = torch.Generator().manual_seed(9)
    C, n = 1500, 200
    preds, target = torch.randn(n, C, generator=g), torch.randint(C, (n,), generator=g)
    target[::7] = -1
    for micro in (False, True):
        m = M.MulticlassPrecisionRecallCurve(C, thresholds=9, average="micro" if micro else None, ignore_index=-1).cuda()
        m.update(preds.cuda(), target.cuda())
        want = O.multiclass_curve_binned(preds.numpy(), target.numpy(), C, torch.linspace(0, 1, 9).numpy(), -1,
                                         cuda_softmax, micro)
        assert np.array_equal(m.confmat.cpu().numpy(), want)
