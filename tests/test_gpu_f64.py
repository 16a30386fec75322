"""float64 scores: the front end of torchmetrics_forked_b200/_f64.py against the UNMODIFIED reference on the same GPU.

The reference evaluates float64 `preds` in float64.  The inputs below are built so that rounding the scores to float32
would change the answer: scores a few 1e-12 above / below a float32 threshold, rows whose two largest entries agree in
float32, exact-curve scores that tie in float32 but not in float64.  Integer states must be equal bit for bit; derived
floats within 1e-6 relative (1e-12 for the float64 thresholds of the exact curves, which are looked up, not computed).
"""
import pytest
import torch

import ref_loader

pytestmark = pytest.mark.gpu

if not ref_loader.staged_available() and not ref_loader.reference_available():
    pytest.skip("oracle/_ref is not staged (run __graft_entry__.build() where /root/reference exists)",
                allow_module_level=True)


@pytest.fixture(scope="module")
def R():
    return ref_loader.load_staged_reference()


@pytest.fixture(scope="module")
def M():
    import torchmetrics_forked_b200 as ours
    import torchmetrics_forked_b200.classification  # noqa: F401
    import torchmetrics_forked_b200.functional.classification  # noqa: F401

    return ours


def _pair(R, M, name, *args, **kw):
    return getattr(M.classification, name)(*args, **kw).cuda(), getattr(R.classification, name)(*args, **kw).cuda()


def _states_equal(a, b):
    for name in b._defaults:
        x, y = getattr(a, name), getattr(b, name)
        if isinstance(y, list):
            x = torch.cat([c.reshape(-1) for c in x]) if isinstance(x, list) else x.reshape(-1)
            y = torch.cat([c.reshape(-1) for c in y])
            assert x.dtype == y.dtype or not y.is_floating_point(), name
            x = x.to(y.dtype)
            if y.is_floating_point():  # NaN scores are stored as they came: equal where both hold one
                assert torch.equal(x.isnan(), y.isnan()), name
                x, y = x.nan_to_num(nan=-7.0), y.nan_to_num(nan=-7.0)
            assert torch.equal(x, y), name
        else:
            assert torch.equal(x, y), name


def _close(a, b, rtol=1e-6):
    if isinstance(b, (tuple, list)):
        assert len(a) == len(b)
        for x, y in zip(a, b):
            _close(x, y, rtol)
        return
    torch.testing.assert_close(a.double(), b.double(), rtol=rtol, atol=1e-7, equal_nan=True)


def _near_grid(n, gen, thresholds):
    """float64 probabilities sitting 1e-12 around float32 threshold values (and some ordinary ones)."""
    thr = thresholds.double()
    pick = thr[torch.randint(thr.numel(), (n,), generator=gen, device="cuda")]
    nudge = (torch.randint(3, (n,), generator=gen, device="cuda") - 1).double() * 1e-12
    p = (pick + nudge).clamp(0, 1)
    plain = torch.rand(n, generator=gen, device="cuda", dtype=torch.float64)
    return torch.where(torch.rand(n, generator=gen, device="cuda") < 0.7, p, plain)


@pytest.mark.parametrize("ignore", [None, -1])
@pytest.mark.parametrize("logits", [False, True])
def test_binary_label_metrics(R, M, ignore, logits):
    g = torch.Generator(device="cuda").manual_seed(11)
    thr32 = torch.tensor([0.3], device="cuda")  # 0.3 is no float32: the float64 compare sees the difference
    for name in ("BinaryStatScores", "BinaryConfusionMatrix", "BinaryAccuracy", "BinaryF1Score"):
        a, b = _pair(R, M, name, threshold=0.3, ignore_index=ignore)
        for _ in range(2):
            p = _near_grid(5000, g, thr32)
            p = torch.where(torch.rand(5000, generator=g, device="cuda") < 0.3, torch.full_like(p, 0.3), p)
            if logits:
                p = torch.logit(p.clamp(1e-9, 1 - 1e-9))
            t = torch.randint(2, (5000,), generator=g, device="cuda")
            if ignore is not None:
                t[::9] = ignore
            a.update(p, t)
            b.update(p, t)
        _states_equal(a, b)
        _close(a.compute(), b.compute())


@pytest.mark.parametrize("ignore", [None, 1])
def test_multilabel_label_metrics(R, M, ignore):
    g = torch.Generator(device="cuda").manual_seed(12)
    for name, kw in (("MultilabelStatScores", {"average": None}), ("MultilabelConfusionMatrix", {}),
                     ("MultilabelF1Score", {"average": "macro"})):
        a, b = _pair(R, M, name, 6, threshold=0.7, ignore_index=ignore, **kw)
        p = _near_grid(6 * 3000, g, torch.tensor([0.7], device="cuda")).reshape(3000, 6)
        t = torch.randint(2, (3000, 6), generator=g, device="cuda")
        a.update(p, t)
        b.update(p, t)
        a.update(p * 8 - 4, t)  # logits: the range test fires
        b.update(p * 8 - 4, t)
        _states_equal(a, b)
        _close(a.compute(), b.compute())


@pytest.mark.parametrize("top_k", [1, 3])
@pytest.mark.parametrize("ignore", [None, 2])
def test_multiclass_label_metrics(R, M, top_k, ignore):
    g = torch.Generator(device="cuda").manual_seed(13)
    C, n = 7, 4000
    cases = [("MulticlassStatScores", {"average": None, "top_k": top_k}), ("MulticlassAccuracy", {"top_k": top_k})]
    if top_k == 1:
        cases.append(("MulticlassConfusionMatrix", {}))
    for name, kw in cases:
        a, b = _pair(R, M, name, C, ignore_index=ignore, **kw)
        p = torch.randn(n, C, generator=g, device="cuda", dtype=torch.float64)
        # the largest entry and a runner-up that differ by 1e-13: one float32 value, two float64 values
        top = p.argmax(1)
        other = (top + 1 + torch.randint(C - 1, (n,), generator=g, device="cuda")) % C
        rows = torch.arange(n, device="cuda")
        p[rows, other] = p[rows, top] - 1e-13
        t = torch.randint(C, (n,), generator=g, device="cuda")
        a.update(p, t)
        b.update(p, t)
        _states_equal(a, b)
        _close(a.compute(), b.compute())


def test_multiclass_samplewise_and_multidim(R, M):
    g = torch.Generator(device="cuda").manual_seed(14)
    p = torch.randn(6, 5, 40, generator=g, device="cuda", dtype=torch.float64)
    t = torch.randint(5, (6, 40), generator=g, device="cuda")
    for kw in ({"multidim_average": "samplewise", "average": None}, {"multidim_average": "global", "average": "macro"},
               {"multidim_average": "samplewise", "average": None, "top_k": 2}):
        a, b = _pair(R, M, "MulticlassStatScores", 5, **kw)
        a.update(p, t)
        b.update(p, t)
        _states_equal(a, b)
        _close(a.compute(), b.compute())


@pytest.mark.parametrize("thresholds", [17, [0.1, 0.35, 0.5, 0.9]])
@pytest.mark.parametrize("ignore", [None, -1])
def test_binary_binned_curves(R, M, thresholds, ignore):
    g = torch.Generator(device="cuda").manual_seed(15)
    grid = torch.linspace(0, 1, thresholds, device="cuda") if isinstance(thresholds, int) else torch.tensor(
        thresholds, device="cuda")
    for name in ("BinaryPrecisionRecallCurve", "BinaryAUROC", "BinaryAveragePrecision", "BinaryROC"):
        a, b = _pair(R, M, name, thresholds=thresholds, ignore_index=ignore)
        for logits in (False, True):
            p = _near_grid(6000, g, grid)
            if logits:
                p = p * 6 - 3
            t = torch.randint(2, (6000,), generator=g, device="cuda")
            if ignore is not None:
                t[::5] = ignore
            a.update(p, t)
            b.update(p, t)
        _states_equal(a, b)
        _close(a.compute(), b.compute())


@pytest.mark.parametrize("ignore", [None, 0])
def test_multilabel_and_multiclass_binned_curves(R, M, ignore):
    g = torch.Generator(device="cuda").manual_seed(16)
    grid = torch.linspace(0, 1, 21, device="cuda")
    a, b = _pair(R, M, "MultilabelPrecisionRecallCurve", 4, thresholds=21, ignore_index=ignore)
    p = _near_grid(4 * 2500, g, grid).reshape(2500, 4)
    t = torch.randint(2, (2500, 4), generator=g, device="cuda")
    a.update(p, t)
    b.update(p, t)
    _states_equal(a, b)
    _close(a.compute(), b.compute())
    for average in (None,):  # (the reference's class cannot hold the micro state: its functional is checked below)
        a, b = _pair(R, M, "MulticlassPrecisionRecallCurve", 5, thresholds=21, ignore_index=ignore, average=average)
        probs = _near_grid(5 * 2500, g, grid).reshape(2500, 5)
        logits = torch.randn(2500, 5, generator=g, device="cuda", dtype=torch.float64) * 3
        t = torch.randint(5, (2500,), generator=g, device="cuda")
        for p in (probs, logits):
            a.update(p, t)
            b.update(p, t)
        _states_equal(a, b)
        _close(a.compute(), b.compute())


def test_multiclass_binned_micro_functional(R, M):
    g = torch.Generator(device="cuda").manual_seed(21)
    grid = torch.linspace(0, 1, 21, device="cuda")
    fo, fr = M.functional.classification, R.functional.classification
    for p in (_near_grid(5 * 2500, g, grid).reshape(2500, 5),
              torch.randn(2500, 5, generator=g, device="cuda", dtype=torch.float64) * 3):
        t = torch.randint(5, (2500,), generator=g, device="cuda")
        for ignore in (None, 0):
            got = fo.multiclass_precision_recall_curve(p, t, 5, thresholds=21, average="micro", ignore_index=ignore)
            want = fr.multiclass_precision_recall_curve(p, t, 5, thresholds=21, average="micro", ignore_index=ignore)
            _close(got, want)
            _close(fo.multiclass_auroc(p, t, 5, thresholds=21, ignore_index=ignore),
                   fr.multiclass_auroc(p, t, 5, thresholds=21, ignore_index=ignore))


def test_binned_nan_is_refused(M):
    m = M.classification.BinaryPrecisionRecallCurve(thresholds=5).cuda()
    p = torch.tensor([0.2, float("nan"), 0.7], device="cuda", dtype=torch.float64)
    with pytest.raises(ValueError, match="NaN"):
        m.update(p, torch.tensor([0, 1, 1], device="cuda"))


def _ties_in_f32(n, gen):
    """float64 probabilities drawn from a few float32-indistinguishable clusters: base + k * 1e-13."""
    base = torch.rand(64, generator=gen, device="cuda", dtype=torch.float64) * 0.9 + 0.05
    k = torch.randint(4, (n,), generator=gen, device="cuda").double()
    return base[torch.randint(64, (n,), generator=gen, device="cuda")] + k * 1e-13


@pytest.mark.parametrize("ignore", [None, -1])
def test_binary_exact_curves(R, M, ignore):
    g = torch.Generator(device="cuda").manual_seed(17)
    for name in ("BinaryPrecisionRecallCurve", "BinaryROC", "BinaryAUROC", "BinaryAveragePrecision"):
        a, b = _pair(R, M, name, thresholds=None, ignore_index=ignore)
        for _ in range(3):
            p = _ties_in_f32(3000, g)
            t = torch.randint(2, (3000,), generator=g, device="cuda")
            if ignore is not None:
                t[::6] = ignore
            a.update(p, t)
            b.update(p, t)
        _states_equal(a, b)
        got, want = a.compute(), b.compute()
        _close(got, want)
        if isinstance(want, tuple):  # the thresholds are float64 values of the input, returned as they are
            assert got[2].dtype == want[2].dtype == torch.float64
            assert torch.equal(got[2], want[2])
            assert got[2].numel() > 64  # more distinct thresholds than float32 could tell apart


def test_binary_exact_curve_logits_and_nan(R, M):
    g = torch.Generator(device="cuda").manual_seed(18)
    a, b = _pair(R, M, "BinaryROC", thresholds=None)
    p = torch.randn(2000, generator=g, device="cuda", dtype=torch.float64) * 4
    p[::50] = float("nan")
    t = torch.randint(2, (2000,), generator=g, device="cuda")
    a.update(p, t)
    b.update(p, t)
    _states_equal(a, b)
    got, want = a.compute(), b.compute()
    # every NaN is a threshold of its own; in which order torch's sort leaves the NaNs among themselves is not defined
    # (its small-sort kernel is not stable), so the points AT the NaN thresholds are not compared, all others are
    assert torch.equal(got[2].isnan(), want[2].isnan()) and int(want[2].isnan().sum()) == 40
    rest = ~want[2].isnan()
    for x, y in zip(got, want):
        assert x.shape == y.shape
        torch.testing.assert_close(x[rest].double(), y[rest].double(), rtol=1e-6, atol=1e-7)


@pytest.mark.parametrize("average", ["macro", None])
def test_multilabel_and_multiclass_exact_scores(R, M, average):
    g = torch.Generator(device="cuda").manual_seed(19)
    for name in ("MultilabelAUROC", "MultilabelAveragePrecision"):
        a, b = _pair(R, M, name, 4, thresholds=None, average=average)
        for _ in range(2):
            p = _ties_in_f32(4 * 1500, g).reshape(1500, 4)
            t = torch.randint(2, (1500, 4), generator=g, device="cuda")
            a.update(p, t)
            b.update(p, t)
        _states_equal(a, b)
        _close(a.compute(), b.compute())
    for name in ("MulticlassAUROC", "MulticlassAveragePrecision", "MulticlassROC"):
        kw = {} if name == "MulticlassROC" else {"average": average}
        a, b = _pair(R, M, name, 5, thresholds=None, **kw)
        for scale in (1.0, 3.0):
            p = torch.randn(1500, 5, generator=g, device="cuda", dtype=torch.float64) * scale
            t = torch.randint(5, (1500,), generator=g, device="cuda")
            a.update(p, t)
            b.update(p, t)
        _states_equal(a, b)
        _close(a.compute(), b.compute())


def test_functional_entry_points(R, M):
    g = torch.Generator(device="cuda").manual_seed(20)
    p = torch.rand(4000, generator=g, device="cuda", dtype=torch.float64)
    t = torch.randint(2, (4000,), generator=g, device="cuda")
    fo, fr = M.functional.classification, R.functional.classification
    _close(fo.binary_auroc(p, t), fr.binary_auroc(p, t))
    _close(fo.binary_auroc(p, t, max_fpr=0.5), fr.binary_auroc(p, t, max_fpr=0.5))
    _close(fo.binary_average_precision(p, t, thresholds=11), fr.binary_average_precision(p, t, thresholds=11))
    _close(fo.binary_f1_score(p, t), fr.binary_f1_score(p, t))
    _close(fo.binary_matthews_corrcoef(p, t), fr.binary_matthews_corrcoef(p, t))
    pm = torch.randn(4000, 6, generator=g, device="cuda", dtype=torch.float64)
    tm = torch.randint(6, (4000,), generator=g, device="cuda")
    _close(fo.multiclass_accuracy(pm, tm, 6), fr.multiclass_accuracy(pm, tm, 6))
    _close(fo.multiclass_cohen_kappa(pm, tm, 6), fr.multiclass_cohen_kappa(pm, tm, 6))
    assert torch.equal(fo.multiclass_confusion_matrix(pm, tm, 6), fr.multiclass_confusion_matrix(pm, tm, 6))
