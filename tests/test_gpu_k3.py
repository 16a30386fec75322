"""GPU parity: K3 (exact curves) — conditional sigmoid, ordered ignore-filter, batched radix sort + scan —
against the oracle and the reference's golden vectors.  Counts (fps / tps) and thresholds must match bit for
bit; AUROC / AP within 1e-6 relative (float32 reductions on the device vs float64 in the oracle: 1e-5)."""
import ctypes

import numpy as np
import pytest
import torch
from conftest import golden_names, ignore_of, load_golden

import hotpath as O

pytestmark = pytest.mark.gpu


def _cuda(x, dtype=None):
    t = torch.from_numpy(np.ascontiguousarray(x))
    if dtype is not None:
        t = t.to(dtype)
    return t.cuda()


def cuda_sigmoid(x: np.ndarray) -> np.ndarray:
    return torch.from_numpy(np.ascontiguousarray(x, dtype=np.float32)).cuda().sigmoid().cpu().numpy()


def _check_curves(got, want):
    fps, tps, thr = got
    for i, (wf, wt, wh) in enumerate(want):
        assert np.array_equal(fps[i].cpu().numpy(), wf), f"fps of label {i}"
        assert np.array_equal(tps[i].cpu().numpy(), wt), f"tps of label {i}"
        assert np.array_equal(thr[i].float().cpu().numpy(), wh, equal_nan=True), f"thresholds of label {i}"


# ------------------------------------------------------------------------------------------ raw C ABI
def test_cabi_clf_curves_raw_pointers():
    from torchmetrics_forked_b200 import _native as nat

    lib = nat.lib()
    g = torch.Generator().manual_seed(5)
    n, L = 30_011, 3
    preds = (torch.rand(n, L, generator=g) * 64).round() / 64  # many ties
    target = torch.randint(2, (n, L), generator=g)
    dp, dt = preds.cuda(), target.cuda()
    nbytes = lib.tmb_clf_curve_scratch_bytes(n, L)
    scratch = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    fps, tps, thr = (torch.full((L, n), -7.0, device="cuda") for _ in range(3))
    counts = torch.zeros(L, dtype=torch.int64, device="cuda")
    err = torch.zeros(1, dtype=torch.int32, device="cuda")
    p = lambda t: ctypes.c_void_p(t.data_ptr())  # noqa: E731
    rc = lib.tmb_clf_curves(p(dp), 0, p(dt), 0, n, L, 0, 0, 1, p(scratch), nbytes, p(fps), p(tps), p(thr), p(counts),
                            p(err), ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0
    torch.cuda.synchronize()
    assert int(err.item()) == 0
    want = O.multilabel_clf_curves(preds.numpy(), target.numpy())
    for l in range(L):
        d = int(counts[l])
        assert d == len(want[l][0]) == 65
        assert np.array_equal(fps[l, :d].cpu().numpy(), want[l][0])
        assert np.array_equal(tps[l, :d].cpu().numpy(), want[l][1])
        assert np.array_equal(thr[l, :d].cpu().numpy(), want[l][2])
        assert float(fps[l, d]) == -7.0  # nothing is written past the curve
    assert lib.tmb_clf_curves(p(dp), 0, p(dt), 0, n, L, 0, 0, 1, p(scratch), 16, p(fps), p(tps), p(thr), p(counts),
                              p(err), None) == -5
    assert lib.tmb_clf_curves(p(dp), 0, p(dt), 0, 1 << 31, L, 0, 0, 1, p(scratch), nbytes, p(fps), p(tps), p(thr),
                              p(counts), p(err), None) == -6


# ------------------------------------------------------------------------------------------ golden vectors
@pytest.mark.parametrize("name", golden_names("clf_"))
def test_binary_exact_golden(name):
    from torchmetrics_forked_b200.classification import BinaryAUROC, BinaryAveragePrecision
    from torchmetrics_forked_b200.functional.classification.precision_recall_curve import _binary_clf_curve

    c = load_golden(name)
    preds, target = _cuda(c["preds"]), _cuda(c["target"])
    fps, tps, thr = _binary_clf_curve(preds, target)
    assert np.array_equal(fps.cpu().numpy(), c["fps"])
    assert np.array_equal(tps.cpu().numpy(), c["tps"])
    assert np.array_equal(thr.cpu().numpy(), c["thr"])
    if c["preds"].size > 1:
        for cls, key in ((BinaryAUROC, "auroc"), (BinaryAveragePrecision, "ap")):
            m = cls(thresholds=None).cuda()
            m.update(preds, target)
            np.testing.assert_allclose(m.compute().cpu().numpy(), c[key], rtol=1e-6, atol=1e-7)


@pytest.mark.parametrize("name", golden_names("mlexact_"))
def test_multilabel_exact_golden(name):
    from torchmetrics_forked_b200.classification import (
        MultilabelAUROC,
        MultilabelAveragePrecision,
        MultilabelPrecisionRecallCurve,
    )

    c = load_golden(name)
    L, ign = int(c["num_labels"]), ignore_of(c)
    preds, target = _cuda(c["preds"]), _cuda(c["target"])
    m = MultilabelPrecisionRecallCurve(L, thresholds=None, ignore_index=ign).cuda()
    m.update(preds, target)
    state = m.preds[0].cpu().numpy()
    want_state = O.exact_state(c["preds"], sigmoid_fn=cuda_sigmoid)
    assert np.array_equal(state, want_state)
    if "logits" not in name:
        assert np.array_equal(state, c["state_preds"])
    from torchmetrics_forked_b200 import kernels

    got = kernels.clf_curves(m.preds[0], m.target[0], ign, 1)
    _check_curves(got, O.multilabel_clf_curves(want_state, c["target"], ign))
    if "logits" not in name:
        for i in range(L):
            assert np.array_equal(got[0][i].cpu().numpy(), c[f"fps_{i}"])
            assert np.array_equal(got[1][i].cpu().numpy(), c[f"tps_{i}"])
            assert np.array_equal(got[2][i].cpu().numpy(), c[f"thr_{i}"])
    for avg in ("micro", "macro", "none"):
        for cls, key in ((MultilabelAUROC, "auroc"), (MultilabelAveragePrecision, "ap")):
            mm = cls(L, average=avg, thresholds=None, ignore_index=ign).cuda()
            mm.update(preds, target)
            np.testing.assert_allclose(mm.compute().cpu().numpy(), c[f"{key}_{avg}"], rtol=2e-6, atol=1e-6)


# ------------------------------------------------------------------------------------------ random cases vs the oracle
@pytest.mark.parametrize(
    ("n", "L", "levels"),
    [
        (1, 1, 0), (2, 1, 0), (33, 2, 4), (8191, 1, 0), (8192, 3, 16), (8193, 1, 1000),
        (100_003, 7, 0), (300_000, 2, 257), (1_000_001, 1, 0),
    ],
)  # fmt: skip
def test_clf_curves_random(n, L, levels):
    from torchmetrics_forked_b200 import kernels

    g = torch.Generator().manual_seed(n + L)
    preds = torch.rand(n, L, generator=g)
    if levels:
        preds = (preds * levels).round() / levels  # heavy ties
    target = torch.randint(2, (n, L), generator=g)
    got = kernels.clf_curves(preds.cuda(), target.cuda(), None, 1)
    _check_curves(got, O.multilabel_clf_curves(preds.numpy(), target.numpy()))


def test_clf_curves_special_values_and_ignore():
    from torchmetrics_forked_b200 import kernels

    g = torch.Generator().manual_seed(77)
    n, L = 50_000, 4
    preds = (torch.rand(n, L, generator=g) * 100).round() / 100
    preds[::97, 0] = float("nan")
    preds[::89, 1] = 0.0
    preds[5::89, 1] = -0.0
    preds[::83, 2] = 1.0
    preds[7::83, 2] = float("inf")
    preds[::101, 3] = 1e-42  # denormal: distinct from 0 only without flush-to-zero
    target = torch.randint(2, (n, L), generator=g)
    target[torch.rand(n, L, generator=g) < 0.3] = -1
    target[:, 3][target[:, 3] == -1] = 0
    target[: n // 2, 3] = -1  # a long ignored prefix
    # NaN (and inf) scores each form their own threshold, so the counts depend on the order argsort leaves
    # them in — which the reference does not define (unstable sort).  Give them one target value so that
    # every tie order yields the same curve.
    target[::97, 0] = 1
    target[7::83, 2] = 1
    got = kernels.clf_curves(preds.cuda(), target.cuda(), -1, 1)
    _check_curves(got, O.multilabel_clf_curves(preds.numpy(), target.numpy(), -1))
    # every entry of one label ignored -> empty curve
    target[:, 0] = -1
    got = kernels.clf_curves(preds.cuda(), target.cuda(), -1, 1)
    assert got[0][0].numel() == 0 and got[2][0].numel() == 0
    with pytest.raises(RuntimeError, match="non-negative"):
        kernels.clf_curves(-preds.abs().cuda() - 1, target.cuda(), None, 1)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_clf_curves_narrow_dtypes(dtype):
    from torchmetrics_forked_b200 import kernels

    g = torch.Generator().manual_seed(3)
    preds = torch.rand(20_000, 2, generator=g).to(dtype)
    target = torch.randint(2, (20_000, 2), generator=g)
    got = kernels.clf_curves(preds.cuda(), target.cuda(), None, 1)
    assert got[2][0].dtype == dtype
    _check_curves(got, O.multilabel_clf_curves(preds.float().numpy(), target.numpy()))


def test_sort_properties_at_scale():
    """Size-independent properties on a case too big for the Python oracle loop: thresholds strictly
    descending, counts monotone, totals add up, number of thresholds = number of distinct scores."""
    from torchmetrics_forked_b200 import kernels

    g = torch.Generator(device="cuda").manual_seed(9)
    n, L = 1 << 22, 3
    preds = torch.rand(n, L, generator=g, device="cuda")
    preds[:, 1] = (preds[:, 1] * 5000).round() / 5000
    target = torch.randint(2, (n, L), generator=g, device="cuda")
    fps, tps, thr = kernels.clf_curves(preds, target, None, 1)
    for l in range(L):
        assert bool((thr[l][1:] < thr[l][:-1]).all())
        assert bool((tps[l][1:] >= tps[l][:-1]).all()) and bool((fps[l][1:] >= fps[l][:-1]).all())
        assert float(tps[l][-1]) == float((target[:, l] == 1).sum())
        assert float(tps[l][-1] + fps[l][-1]) == n
        assert thr[l].numel() == torch.unique(preds[:, l]).numel()
        assert torch.equal(thr[l], torch.unique(preds[:, l]).flip(0))
    # AUROC of label 1 against a torch (not oracle) rank computation: U statistic with ties
    p, t = preds[:, 1].double(), target[:, 1]
    order = torch.argsort(p)
    uniq, inv, cnt = torch.unique(p, return_inverse=True, return_counts=True)
    start = torch.cumsum(cnt, 0) - cnt
    avg_rank = (start + (cnt + 1) / 2.0)[inv]
    n_pos = int((t == 1).sum())
    u = avg_rank[t == 1].sum() - n_pos * (n_pos + 1) / 2
    auc = float(u / (n_pos * (n - n_pos)))
    from torchmetrics_forked_b200.functional.classification.auroc import _auc_compute_without_check
    from torchmetrics_forked_b200.functional.classification.roc import _roc_from_curve

    fpr, tpr, _ = _roc_from_curve(fps[1], tps[1], thr[1])
    assert abs(float(_auc_compute_without_check(fpr.double(), tpr.double(), 1.0)) - auc) < 1e-9
    del order


# ------------------------------------------------------------------------------------------ update-time helpers
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16])
def test_sigmoid_normalize_matches_torch(dtype):
    from torchmetrics_forked_b200 import kernels

    g = torch.Generator().manual_seed(4)
    err = torch.zeros(1, dtype=torch.int32, device="cuda")
    logits = (torch.randn(1_000_003, generator=g) * 4).to(dtype).cuda()
    assert torch.equal(kernels.sigmoid_normalize(logits, None, None, err), logits.sigmoid())
    probs = torch.rand(1_000_003, generator=g).to(dtype).cuda()
    out = kernels.sigmoid_normalize(probs, None, None, err)
    assert torch.equal(out, probs) and out.data_ptr() != probs.data_ptr()
    # one NaN, or one value a hair above 1, flips the whole batch
    spiked = probs.clone()
    spiked[12345] = float("nan")
    out = kernels.sigmoid_normalize(spiked, None, None, err)
    assert torch.equal(out[:100], spiked[:100].sigmoid()) and bool(out[12345].isnan())
    spiked = probs.clone()
    spiked[-1] = 1.5
    assert torch.equal(kernels.sigmoid_normalize(spiked, None, None, err), spiked.sigmoid())
    # ... unless it sits at an ignored position and the caller asks for that exclusion (H2)
    target = torch.zeros(spiked.numel(), dtype=torch.long, device="cuda")
    target[-1] = -1
    assert torch.equal(kernels.sigmoid_normalize(spiked, target, -1, err), spiked)
    assert torch.equal(kernels.sigmoid_normalize(spiked, target, -2, err), spiked.sigmoid())
    assert kernels.sigmoid_normalize(probs[:0], None, None, err).numel() == 0


@pytest.mark.parametrize("tdtype", [torch.int64, torch.int32, torch.uint8])
@pytest.mark.parametrize("pdtype", [torch.float32, torch.bfloat16])
def test_compact_not_ignored(tdtype, pdtype):
    from torchmetrics_forked_b200 import kernels

    g = torch.Generator().manual_seed(6)
    for n in (1, 4095, 4096, 4097, 1_000_003):
        preds = torch.rand(n, generator=g).to(pdtype).cuda()
        target = torch.randint(3, (n,), generator=g).to(tdtype).cuda()
        p, t = kernels.compact_not_ignored(preds, target, 2)
        keep = target != 2
        assert torch.equal(p, preds[keep]) and torch.equal(t, target[keep])
    p, t = kernels.compact_not_ignored(preds, torch.full_like(target, 2), 2)
    assert p.numel() == 0 and t.numel() == 0


def test_binary_exact_with_ignore_index_and_logits():
    from torchmetrics_forked_b200.classification import BinaryAUROC, BinaryPrecisionRecallCurve

    g = torch.Generator().manual_seed(8)
    n = 40_000
    preds = torch.randn(n, generator=g)
    target = torch.randint(2, (n,), generator=g)
    target[torch.rand(n, generator=g) < 0.25] = -1
    m = BinaryPrecisionRecallCurve(thresholds=None, ignore_index=-1).cuda()
    m.update(preds[: n // 2].cuda(), target[: n // 2].cuda())
    m.update(preds[n // 2 :].cuda(), target[n // 2 :].cuda())
    keep = (target != -1).numpy()
    stored = np.concatenate([
        O.exact_state(preds[: n // 2].numpy()[keep[: n // 2]], sigmoid_fn=cuda_sigmoid),
        O.exact_state(preds[n // 2 :].numpy()[keep[n // 2 :]], sigmoid_fn=cuda_sigmoid),
    ])
    assert np.array_equal(torch.cat(m.preds).cpu().numpy(), stored)
    assert np.array_equal(torch.cat(m.target).cpu().numpy(), target.numpy()[keep])
    precision, recall, thr = m.compute()
    fps, tps, wthr = O.binary_clf_curve(stored, target.numpy()[keep])
    assert np.array_equal(thr.cpu().numpy(), wthr[::-1])
    np.testing.assert_allclose(precision.cpu().numpy()[:-1], (tps / (tps + fps))[::-1], rtol=1e-6)
    a = BinaryAUROC(thresholds=None, ignore_index=-1).cuda()
    a.update(preds.cuda(), target.cuda())
    whole = O.exact_state(preds.numpy()[keep], sigmoid_fn=cuda_sigmoid)
    f2, t2, _ = O.binary_clf_curve(whole, target.numpy()[keep])
    np.testing.assert_allclose(float(a.compute()), O.binary_auroc_exact(f2, t2), rtol=1e-5)


def test_multilabel_auroc_config4_scaled():
    """BASELINE config 4 shape (L = 100, fp32 logits, exact mode) at a size the oracle finishes quickly."""
    from torchmetrics_forked_b200.classification import MultilabelAUROC, MultilabelAveragePrecision

    g = torch.Generator().manual_seed(10)
    n, L = 6000, 100
    auroc = MultilabelAUROC(L, thresholds=None, average="none").cuda()
    ap = MultilabelAveragePrecision(L, thresholds=None, average="macro").cuda()
    P, T = [], []
    for _ in range(3):
        preds, target = torch.randn(n, L, generator=g), torch.randint(2, (n, L), generator=g)
        auroc.update(preds.cuda(), target.cuda())
        ap.update(preds.cuda(), target.cuda())
        P.append(O.exact_state(preds.numpy(), sigmoid_fn=cuda_sigmoid))
        T.append(target.numpy())
    curves = O.multilabel_clf_curves(np.concatenate(P), np.concatenate(T))
    np.testing.assert_allclose(auroc.compute().cpu().numpy(), [O.binary_auroc_exact(f, t) for f, t, _ in curves],
                               rtol=1e-5)
    np.testing.assert_allclose(float(ap.compute()),
                               np.mean([O.binary_average_precision_exact(f, t) for f, t, _ in curves]), rtol=1e-5)


# ------------------------------------------------------------------------------------------ chunked inputs, fused scores
def _ragged_chunks(n_rows, L, g, quant=None, ignore_frac=0.0, dtype=torch.float32, tdtype=torch.int64):
    """Per-update buffers of ragged sizes (one of them empty) plus their concatenation."""
    ps, ts = [], []
    for n in n_rows:
        p = torch.rand(n, L, generator=g)
        if quant:
            p = (p * quant).round() / quant
        t = torch.randint(2, (n, L), generator=g)
        if ignore_frac:
            t[torch.rand(n, L, generator=g) < ignore_frac] = -1
        ps.append(p.to(dtype).cuda())
        ts.append(t.to(tdtype).cuda())
    return ps, ts, torch.cat(ps), torch.cat(ts)


@pytest.mark.parametrize("L", [1, 2, 3, 31, 33, 100, 128, 130, 257])
def test_chunked_curves_equal_concatenated(L):
    """tmb_clf_curves_chunked reads the per-update buffers in place: same bits as the concatenated call, and as
    the oracle, for every label-tile shape of the pack kernel (L below / at / above 32 and 128)."""
    from torchmetrics_forked_b200 import kernels

    g = torch.Generator().manual_seed(100 + L)
    rows = [1, 127, 0, 128, 129, 300, 5] if L > 8 else [1, 4095, 0, 4097, 129, 9000, 5]
    ps, ts, P, T = _ragged_chunks(rows, L, g, quant=512, ignore_frac=0.1)
    got = kernels.clf_curves(ps, ts, -1, 1)
    one = kernels.clf_curves(P, T, -1, 1)
    for a, b in zip(got, one):
        for x, y in zip(a, b):
            assert torch.equal(x, y)
    _check_curves(got, O.multilabel_clf_curves(P.cpu().numpy(), T.cpu().numpy(), -1))


@pytest.mark.parametrize("pdtype,tdtype", [(torch.bfloat16, torch.uint8), (torch.float16, torch.int32)])
def test_chunked_curves_narrow_dtypes(pdtype, tdtype):
    from torchmetrics_forked_b200 import kernels

    g = torch.Generator().manual_seed(77)
    ps, ts, P, T = _ragged_chunks([700, 1, 2049], 5, g, dtype=pdtype, tdtype=tdtype)
    got = kernels.clf_curves(ps, ts, None, 1)
    _check_curves(got, O.multilabel_clf_curves(P.float().cpu().numpy(), T.cpu().numpy()))


def _emit_then_reduce(P, T, ignore_index):
    """The two-step path the fused kernel replaces: tmb_clf_curves + tmb_curve_reduce through the raw C ABI."""
    from torchmetrics_forked_b200 import _native as nat

    lib, p = nat.lib(), lambda t: ctypes.c_void_p(t.data_ptr())
    n, L = P.shape
    scratch = torch.empty(lib.tmb_clf_curve_scratch_bytes(n, L), dtype=torch.uint8, device="cuda")
    fps, tps, thr = (torch.empty(L, n, device="cuda") for _ in range(3))
    counts = torch.zeros(L, dtype=torch.int64, device="cuda")
    err = torch.zeros(1, dtype=torch.int32, device="cuda")
    has = int(ignore_index is not None)
    assert lib.tmb_clf_curves(p(P), nat.float_code(P), p(T), nat.label_code(T), n, L, has, ignore_index or 0, 1,
                              p(scratch), scratch.numel(), p(fps), p(tps), p(thr), p(counts), p(err), None) == 0
    out = torch.empty(L, 4, dtype=torch.float64, device="cuda")
    assert lib.tmb_curve_reduce(p(fps), p(tps), n, L, p(counts), p(out), None) == 0
    torch.cuda.synchronize()
    return out


@pytest.mark.parametrize("quant", [None, 4096, 3, 1])
def test_fused_scores_equal_emit_plus_reduce(quant):
    """k3_curve_scores (no curve in memory) against curve emit + reduce: same float32 curve points, float64 sums
    that differ only in summation order.  quant = 3 / 1 makes tie runs that span many 4096-key tiles, i.e. the
    look-back over tiles without a boundary; one label is ignored entirely (no curve points at all)."""
    from torchmetrics_forked_b200 import kernels

    g = torch.Generator().manual_seed(31)
    ps, ts, P, T = _ragged_chunks([9000, 0, 70_001, 4096], 4, g, quant=quant, ignore_frac=0.05)
    for t in ts:
        t[:, 2] = -1
    T = torch.cat(ts)
    fused = kernels.clf_curve_scores(ps, ts, -1, 1)
    want = _emit_then_reduce(P, T, -1)
    assert torch.equal(fused[:, 2:], want[:, 2:])
    torch.testing.assert_close(fused[:, :2], want[:, :2], rtol=1e-12, atol=0)
    assert fused[2].abs().sum() == 0
    # and against the oracle's float64 areas (the reference itself raises IndexError on a label without samples)
    keep = [0, 1, 3]
    curves = O.multilabel_clf_curves(P.cpu().numpy()[:, keep], T.cpu().numpy()[:, keep], -1)
    for l, (f, t, _) in zip(keep, curves):
        auroc = float(fused[l, 0] / (fused[l, 2] * fused[l, 3]))
        np.testing.assert_allclose(auroc, O.binary_auroc_exact(f, t), rtol=1e-6)
        np.testing.assert_allclose(float(fused[l, 1] / fused[l, 3]), O.binary_average_precision_exact(f, t), rtol=1e-6)


def test_fused_scores_single_tensor_and_empty():
    from torchmetrics_forked_b200 import kernels

    g = torch.Generator().manual_seed(32)
    _, _, P, T = _ragged_chunks([33_333], 7, g)
    torch.testing.assert_close(kernels.clf_curve_scores(P, T, None, 1), _emit_then_reduce(P, T, None), rtol=1e-12,
                               atol=0)
    empty = kernels.clf_curve_scores(P[:0], T[:0], None, 1)
    assert empty.shape == (7, 4) and float(empty.abs().sum()) == 0
    with pytest.raises(RuntimeError, match="negative|range|score"):
        bad = P.clone()
        bad[5, 3] = -0.5
        kernels.clf_curves(bad, T, None, 1)
    assert bool(kernels.clf_curve_scores(bad, T, None, 1).isnan().all())


@pytest.mark.parametrize("average", ["macro", "weighted", "micro", "none"])
def test_multilabel_metrics_over_many_updates(average):
    """Several updates -> compute() reads the per-update buffers through the chunk table; the state itself stays
    the list of buffers the reference keeps."""
    from torchmetrics_forked_b200.classification import MultilabelAUROC, MultilabelAveragePrecision, MultilabelROC
    from torchmetrics_forked_b200.functional.classification import multilabel_auroc, multilabel_average_precision

    g = torch.Generator().manual_seed(12)
    L = 6
    auroc = MultilabelAUROC(L, thresholds=None, average=average, ignore_index=-1).cuda()
    ap = MultilabelAveragePrecision(L, thresholds=None, average=average, ignore_index=-1).cuda()
    roc = MultilabelROC(L, thresholds=None, ignore_index=-1).cuda()
    ps, ts, P, T = _ragged_chunks([500, 4097, 33, 1200], L, g, quant=256, ignore_frac=0.1)
    for p, t in zip(ps, ts):
        auroc.update(p, t)
        ap.update(p, t)
        roc.update(p, t)
    assert len(auroc.preds) == 4 and all(torch.equal(a, b) for a, b in zip(auroc.preds, ps))
    torch.testing.assert_close(auroc.compute(), multilabel_auroc(P, T, L, average=average, ignore_index=-1),
                               rtol=1e-6, atol=0)
    torch.testing.assert_close(ap.compute(), multilabel_average_precision(P, T, L, average=average, ignore_index=-1),
                               rtol=1e-6, atol=0)
    fpr, tpr, thr = roc.compute()
    curves = O.multilabel_clf_curves(P.cpu().numpy(), T.cpu().numpy(), -1)
    for l in range(L):
        assert np.array_equal(thr[l].cpu().numpy()[1:], curves[l][2])
        np.testing.assert_allclose(tpr[l].cpu().numpy()[1:], curves[l][1] / curves[l][1][-1], rtol=1e-6)
    if average == "none":
        want = [O.binary_auroc_exact(f, t) for f, t, _ in curves]
        np.testing.assert_allclose(auroc.compute().cpu().numpy(), want, rtol=1e-6)


def test_binary_metrics_over_many_updates():
    from torchmetrics_forked_b200.classification import BinaryAUROC, BinaryAveragePrecision, BinaryROC

    g = torch.Generator().manual_seed(13)
    chunks = [(torch.rand(n, generator=g).cuda(), torch.randint(2, (n,), generator=g).cuda()) for n in (17, 5000, 1, 812)]
    P, T = torch.cat([p for p, _ in chunks]), torch.cat([t for _, t in chunks])
    au, ap, roc = BinaryAUROC(thresholds=None).cuda(), BinaryAveragePrecision(thresholds=None).cuda(), BinaryROC(thresholds=None).cuda()
    for p, t in chunks:
        au.update(p, t)
        ap.update(p, t)
        roc.update(p, t)
    f, t, h = O.binary_clf_curve(P.cpu().numpy(), T.cpu().numpy())
    np.testing.assert_allclose(float(au.compute()), O.binary_auroc_exact(f, t), rtol=1e-6)
    np.testing.assert_allclose(float(ap.compute()), O.binary_average_precision_exact(f, t), rtol=1e-6)
    assert np.array_equal(roc.compute()[2].cpu().numpy()[1:], h)


def test_fused_scores_float32_rounding_path_beyond_2p24():
    """From 2^24 rows on, float32(1 + idx) - float32(tps) rounds (as in the reference's float32 curves): the
    integer fast path is off and the fused kernel must still reproduce emit + reduce."""
    from torchmetrics_forked_b200 import kernels

    g = torch.Generator(device="cuda").manual_seed(33)
    n = (1 << 24) + 4099
    P = torch.rand(n, 1, generator=g, device="cuda")
    T = torch.randint(2, (n, 1), generator=g, device="cuda")
    fused = kernels.clf_curve_scores([P[:5_000_000], P[5_000_000:]], [T[:5_000_000], T[5_000_000:]], None, 1)
    want = _emit_then_reduce(P, T, None)
    assert torch.equal(fused[:, 2:], want[:, 2:])
    torch.testing.assert_close(fused[:, :2], want[:, :2], rtol=1e-12, atol=0)
    no_ap = kernels.clf_curve_scores(P, T, None, 1, want_ap=False)
    assert torch.equal(no_ap[:, 0], fused[:, 0]) and float(no_ap[0, 1]) == 0.0
    small = kernels.clf_curve_scores(P[:100_000], T[:100_000], None, 1, want_ap=False)
    torch.testing.assert_close(small[:, 0], _emit_then_reduce(P[:100_000].contiguous(), T[:100_000].contiguous(), None)[:, 0],
                               rtol=0, atol=0)


def test_compute_on_cpu_exact_mode_brings_the_buffers_back():
    """`compute_on_cpu=True` moves the list states to host memory after every update (reference metric.py:172-173);
    `compute()` of the exact-mode metrics must still work (the buffers return to the metric's device for the sort)
    and give the values of an ordinary metric."""
    from torchmetrics_forked_b200 import classification as M

    g = torch.Generator().manual_seed(91)
    pairs = [
        (M.MultilabelAUROC(4, thresholds=None, compute_on_cpu=True), M.MultilabelAUROC(4, thresholds=None), (300, 4)),
        (M.BinaryAveragePrecision(thresholds=None, compute_on_cpu=True), M.BinaryAveragePrecision(thresholds=None), (500,)),
        (M.BinaryROC(thresholds=None, compute_on_cpu=True), M.BinaryROC(thresholds=None), (400,)),
    ]
    for parked, plain, shape in pairs:
        parked, plain = parked.cuda(), plain.cuda()
        for _ in range(3):
            p, t = torch.randn(*shape, generator=g).cuda(), torch.randint(2, shape, generator=g).cuda()
            parked.update(p, t)
            plain.update(p, t)
        assert all(not c.is_cuda for c in parked.preds) and all(not c.is_cuda for c in parked.target)
        a, b = parked.compute(), plain.compute()
        if isinstance(a, tuple):
            assert all(torch.equal(x, y) for x, y in zip(a, b))
        else:
            assert torch.equal(a, b)
