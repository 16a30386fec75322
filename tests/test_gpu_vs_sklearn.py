"""Third-opinion oracle (SURVEY §8c tier 3): scikit-learn / scipy, the checkers the reference's own test-suite uses
(tests/unittests/classification/test_confusion_matrix.py:139-147, test_stat_scores.py:163-258, test_jaccard.py:132-147,
test_auroc.py:35-42 + 254-279, test_average_precision.py:44-75, test_precision_recall_curve.py:43-49).  Independent
of both the numpy oracle and the reference: a disagreement here would mean all three of our own checks share a
misreading of the semantics.  Tolerance 1e-6 absolute on derived floats (sklearn computes in float64, the metrics in
float32), exact on integer confusion matrices."""
import numpy as np
import pytest
import torch

sk = pytest.importorskip("sklearn.metrics")

pytestmark = pytest.mark.gpu

NUM_BATCHES, BATCH, C, L = 4, 512, 5, 4  # the reference's shapes scaled up a little (tests/unittests/conftest.py:28-33)


def _batches(make):
    g = torch.Generator().manual_seed(42)
    return [make(g) for _ in range(NUM_BATCHES)]


def _run(metric, batches):
    metric = metric.cuda()
    for p, t in batches:
        metric.update(p.cuda(), t.cuda())
    return metric.compute()


def _np(x):
    return x.detach().cpu().numpy()


def _softmax(x):
    e = np.exp(x - x.max(1, keepdims=True))
    return e / e.sum(1, keepdims=True)


@pytest.mark.parametrize("kind", ["logits", "probs", "labels"])
def test_multiclass_family_against_sklearn(kind):
    from torchmetrics_forked_b200 import classification as M

    def make(g):
        t = torch.randint(C, (BATCH,), generator=g)
        if kind == "labels":
            return torch.randint(C, (BATCH,), generator=g), t
        x = torch.randn(BATCH, C, generator=g)
        return (x.softmax(1) if kind == "probs" else x), t

    batches = _batches(make)
    yp = np.concatenate([_np(p) if kind == "labels" else _np(p).argmax(1) for p, _ in batches])
    yt = np.concatenate([_np(t) for _, t in batches])
    labels = list(range(C))
    assert np.array_equal(_np(_run(M.MulticlassConfusionMatrix(C), batches)), sk.confusion_matrix(yt, yp, labels=labels))
    for norm in ("true", "pred", "all"):
        got = _np(_run(M.MulticlassConfusionMatrix(C, normalize=norm), batches))
        np.testing.assert_allclose(got, sk.confusion_matrix(yt, yp, labels=labels, normalize=norm), atol=1e-6)
    np.testing.assert_allclose(_np(_run(M.MulticlassAccuracy(C, average="micro"), batches)), sk.accuracy_score(yt, yp), atol=1e-6)
    np.testing.assert_allclose(_np(_run(M.MulticlassAccuracy(C, average="macro"), batches)),
                               sk.balanced_accuracy_score(yt, yp), atol=1e-6)
    for avg in ("micro", "macro", "weighted", None):
        kw = dict(average=avg, labels=labels, zero_division=0)
        np.testing.assert_allclose(_np(_run(M.MulticlassF1Score(C, average=avg), batches)), sk.f1_score(yt, yp, **kw), atol=1e-6)
        np.testing.assert_allclose(_np(_run(M.MulticlassPrecision(C, average=avg), batches)), sk.precision_score(yt, yp, **kw), atol=1e-6)
        np.testing.assert_allclose(_np(_run(M.MulticlassRecall(C, average=avg), batches)), sk.recall_score(yt, yp, **kw), atol=1e-6)
        np.testing.assert_allclose(_np(_run(M.MulticlassJaccardIndex(C, average=avg), batches)),
                                   sk.jaccard_score(yt, yp, average=avg, labels=labels), atol=1e-6)
    np.testing.assert_allclose(_np(_run(M.MulticlassCohenKappa(C), batches)), sk.cohen_kappa_score(yt, yp), atol=1e-6)
    np.testing.assert_allclose(_np(_run(M.MulticlassMatthewsCorrCoef(C), batches)), sk.matthews_corrcoef(yt, yp), atol=1e-6)


@pytest.mark.parametrize("kind", ["logits", "probs"])
def test_binary_curves_and_areas_against_sklearn(kind):
    from torchmetrics_forked_b200 import classification as M

    def make(g):
        x = torch.randn(BATCH, generator=g) * 2
        x[::5] = (x[::5] * 2).round() / 2  # ties
        return (x.sigmoid() if kind == "probs" else x), torch.randint(2, (BATCH,), generator=g)

    batches = _batches(make)
    score = np.concatenate([_np(p.sigmoid() if kind == "logits" else p) for p, _ in batches]).astype(np.float64)
    yt = np.concatenate([_np(t) for _, t in batches])
    np.testing.assert_allclose(_np(_run(M.BinaryAUROC(thresholds=None), batches)), sk.roc_auc_score(yt, score), atol=1e-6)
    np.testing.assert_allclose(_np(_run(M.BinaryAveragePrecision(thresholds=None), batches)),
                               sk.average_precision_score(yt, score), atol=1e-6)
    fpr, tpr, thr = (_np(x) for x in _run(M.BinaryROC(thresholds=None), batches))
    sfpr, stpr, sthr = sk.roc_curve(yt, score, drop_intermediate=False)
    np.testing.assert_allclose(fpr, sfpr, atol=1e-6)
    np.testing.assert_allclose(tpr, stpr, atol=1e-6)
    np.testing.assert_allclose(thr[1:], sthr[1:], atol=1e-6)  # the first threshold is a convention (1.0 here, inf there)
    prec, rec, pthr = (_np(x) for x in _run(M.BinaryPrecisionRecallCurve(thresholds=None), batches))
    sprec, srec, spthr = sk.precision_recall_curve(yt, score)
    # sklearn >= 1.1 keeps the points below full recall; the reference keeps every distinct threshold: compare the
    # common (highest-threshold) part and the end point
    n = len(spthr)
    np.testing.assert_allclose(prec[-n - 1 :], sprec, atol=1e-6)
    np.testing.assert_allclose(rec[-n - 1 :], srec, atol=1e-6)
    np.testing.assert_allclose(pthr[-n:], spthr, atol=1e-6)
    # binned AUROC converges to the exact one (thresholds = every distinct score: the reference's own equivalence test,
    # tests/unittests/classification/test_auroc.py:129-138)
    distinct = torch.from_numpy(np.unique(score).astype(np.float32))
    binned = _run(M.BinaryAUROC(thresholds=distinct), [(torch.from_numpy(score.astype(np.float32)), torch.from_numpy(yt))])
    np.testing.assert_allclose(_np(binned), sk.roc_auc_score(yt, score), atol=1e-5)


@pytest.mark.parametrize("average", ["macro", "weighted", "none", "micro"])
def test_multilabel_and_multiclass_areas_against_sklearn(average):
    from torchmetrics_forked_b200 import classification as M

    ml = _batches(lambda g: (torch.randn(BATCH, L, generator=g), torch.randint(2, (BATCH, L), generator=g)))
    score = np.concatenate([_np(p.sigmoid()) for p, _ in ml]).astype(np.float64)
    yt = np.concatenate([_np(t) for _, t in ml])
    sk_avg = None if average == "none" else average
    np.testing.assert_allclose(_np(_run(M.MultilabelAUROC(L, average=average, thresholds=None), ml)),
                               sk.roc_auc_score(yt, score, average=sk_avg), atol=1e-6)
    np.testing.assert_allclose(_np(_run(M.MultilabelAveragePrecision(L, average=average, thresholds=None), ml)),
                               sk.average_precision_score(yt, score, average=sk_avg), atol=1e-6)
    if average == "micro":
        return
    mc = _batches(lambda g: (torch.randn(BATCH, C, generator=g), torch.randint(C, (BATCH,), generator=g)))
    probs = _softmax(np.concatenate([_np(p) for p, _ in mc]).astype(np.float64))
    ytc = np.concatenate([_np(t) for _, t in mc])
    got = _np(_run(M.MulticlassAUROC(C, average=average, thresholds=None), mc))
    if average == "none":
        want = [sk.roc_auc_score((ytc == c).astype(int), probs[:, c]) for c in range(C)]
    else:
        want = sk.roc_auc_score(ytc, probs, multi_class="ovr", average=average, labels=list(range(C)))
    np.testing.assert_allclose(got, want, atol=1e-6)


def test_binary_and_multilabel_stat_scores_against_sklearn():
    from torchmetrics_forked_b200 import classification as M

    b = _batches(lambda g: (torch.rand(BATCH, generator=g), torch.randint(2, (BATCH,), generator=g)))
    yp = np.concatenate([(_np(p) > 0.5).astype(int) for p, _ in b])
    yt = np.concatenate([_np(t) for _, t in b])
    assert np.array_equal(_np(_run(M.BinaryConfusionMatrix(), b)), sk.confusion_matrix(yt, yp, labels=[0, 1]))
    np.testing.assert_allclose(_np(_run(M.BinaryF1Score(), b)), sk.f1_score(yt, yp), atol=1e-6)
    np.testing.assert_allclose(_np(_run(M.BinaryAccuracy(), b)), sk.accuracy_score(yt, yp), atol=1e-6)
    np.testing.assert_allclose(_np(_run(M.BinaryJaccardIndex(), b)), sk.jaccard_score(yt, yp), atol=1e-6)
    ml = _batches(lambda g: (torch.randn(BATCH, L, generator=g), torch.randint(2, (BATCH, L), generator=g)))
    ypl = np.concatenate([(_np(p.sigmoid()) > 0.5).astype(int) for p, _ in ml])
    ytl = np.concatenate([_np(t) for _, t in ml])
    got = _np(_run(M.MultilabelConfusionMatrix(L), ml))
    assert np.array_equal(got, sk.multilabel_confusion_matrix(ytl, ypl))
    for avg in ("micro", "macro", "weighted", None):
        np.testing.assert_allclose(_np(_run(M.MultilabelF1Score(L, average=avg), ml)),
                                   sk.f1_score(ytl, ypl, average=avg, zero_division=0), atol=1e-6)
