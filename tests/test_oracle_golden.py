"""Pin the oracle (oracle/hotpath.py) against golden vectors produced by the unmodified reference
(oracle/gen_golden.py) and against the known-answer vectors hard-coded in the reference's own tests."""
import numpy as np
import pytest
from conftest import golden_names, ignore_of, load_golden

import hotpath as O


# ------------------------------------------------------------------------------------------ K1
@pytest.mark.parametrize("name", golden_names("mc_"))
def test_multiclass_confmat_and_stats(name):
    c = load_golden(name)
    C, ign = int(c["num_classes"]), ignore_of(c)
    assert np.array_equal(O.multiclass_confusion_matrix(c["preds"], c["target"], C, ign), c["confmat"])
    for avg in ("micro", "macro", "none"):
        tp, fp, tn, fn = O.multiclass_stat_scores(c["preds"], c["target"], C, avg, ign)
        if avg == "none":
            ref = c["stat_none"]
            assert np.array_equal(np.stack([tp, fp, tn, fn, tp + fn], -1), ref)
        np.testing.assert_allclose(O.accuracy_reduce(tp, fp, tn, fn, avg), c[f"acc_{avg}"], rtol=1e-6, atol=1e-7)
        np.testing.assert_allclose(O.fbeta_reduce(tp, fp, tn, fn, 1.0, avg), c[f"f1_{avg}"], rtol=1e-6, atol=1e-7)
    tp, fp, tn, fn = O.multiclass_stat_scores(c["preds"], c["target"], C, "micro", ign)
    assert np.array_equal(np.array([tp, fp, tn, fn, tp + fn]), c["stat_micro"])
    for avg in ("micro", "macro", "weighted", "none"):
        np.testing.assert_allclose(
            O.jaccard_reduce(c["confmat"], avg, ign), c[f"jaccard_{avg}"], rtol=1e-6, atol=1e-7
        )


# ------------------------------------------------------------------------------------------ K1b
@pytest.mark.parametrize("name", golden_names("bin_"))
def test_binary_counts(name):
    c = load_golden(name)
    ign, thr = ignore_of(c), float(c["threshold"])
    tp, fp, tn, fn = O.binary_stat_scores(c["preds"], c["target"], thr, ign)
    assert [tp, fp, tn, fn, tp + fn] == c["stat"].tolist()
    assert np.array_equal(O.binary_confusion_matrix(c["preds"], c["target"], thr, ign), c["confmat"])


@pytest.mark.parametrize("name", golden_names("ml_"))
def test_multilabel_counts(name):
    c = load_golden(name)
    L, ign, thr = int(c["num_labels"]), ignore_of(c), float(c["threshold"])
    tp, fp, tn, fn = O.multilabel_stat_scores(c["preds"], c["target"], L, thr, ign)
    assert np.array_equal(np.stack([tp, fp, tn, fn, tp + fn], -1), c["stat_none"])
    assert np.array_equal(np.stack([tp, fp, tn, fn, tp + fn], -1).sum(0), c["stat_micro"])
    assert np.array_equal(O.multilabel_confusion_matrix(c["preds"], c["target"], L, thr, ign), c["confmat"])


# ------------------------------------------------------------------------------------------ K2
@pytest.mark.parametrize("name", golden_names("bcurve_"))
def test_binary_binned_curve(name):
    c = load_golden(name)
    state = O.binary_curve_binned(c["preds"], c["target"], c["thresholds"], ignore_of(c))
    assert np.array_equal(state, c["state"])
    np.testing.assert_allclose(O.binary_auroc_binned(state), c["auroc"], rtol=1e-6, atol=1e-7)
    np.testing.assert_allclose(O.binary_average_precision_binned(state), c["ap"], rtol=1e-6, atol=1e-7)


@pytest.mark.parametrize("name", golden_names("mlcurve_"))
def test_multilabel_binned_curve(name):
    c = load_golden(name)
    L = int(c["num_labels"])
    state = O.multilabel_curve_binned(c["preds"], c["target"], L, c["thresholds"], ignore_of(c))
    assert np.array_equal(state, c["state"])
    per_label = [O.binary_auroc_binned(state[:, l]) for l in range(L)]
    np.testing.assert_allclose(per_label, c["auroc_none"], rtol=1e-6, atol=1e-7)
    np.testing.assert_allclose(np.mean(per_label), c["auroc_macro"], rtol=1e-6, atol=1e-7)
    np.testing.assert_allclose(O.binary_auroc_binned(state.sum(1)), c["auroc_micro"], rtol=1e-6, atol=1e-7)
    ap = [O.binary_average_precision_binned(state[:, l]) for l in range(L)]
    np.testing.assert_allclose(ap, c["ap_none"], rtol=1e-6, atol=1e-7)


# ------------------------------------------------------------------------------------------ K3
@pytest.mark.parametrize("name", golden_names("clf_"))
def test_binary_clf_curve(name):
    c = load_golden(name)
    fps, tps, thr = O.binary_clf_curve(c["preds"], c["target"])
    assert np.array_equal(fps, c["fps"]) and np.array_equal(tps, c["tps"])
    assert np.array_equal(thr, c["thr"])  # -0.0 == 0.0 compares equal, as torch.equal would
    if c["preds"].size > 1 and len(np.unique(c["target"])) == 2:
        np.testing.assert_allclose(O.binary_auroc_exact(fps, tps), c["auroc"], rtol=1e-6, atol=1e-7)
        np.testing.assert_allclose(O.binary_average_precision_exact(fps, tps), c["ap"], rtol=1e-6, atol=1e-7)


@pytest.mark.parametrize("name", golden_names("mlexact_"))
def test_multilabel_exact(name):
    c = load_golden(name)
    L, ign = int(c["num_labels"]), ignore_of(c)
    state = O.exact_state(c["preds"])
    assert np.array_equal(state, c["state_preds"])
    curves = O.multilabel_clf_curves(state, c["target"], ign)
    for i, (fps, tps, thr) in enumerate(curves):
        assert np.array_equal(fps, c[f"fps_{i}"]) and np.array_equal(tps, c[f"tps_{i}"])
        assert np.array_equal(thr, c[f"thr_{i}"])
    auroc = [O.binary_auroc_exact(f, t) for f, t, _ in curves]
    np.testing.assert_allclose(auroc, c["auroc_none"], rtol=1e-6, atol=1e-7)
    np.testing.assert_allclose(np.mean(auroc), c["auroc_macro"], rtol=1e-6, atol=1e-7)
    ap = [O.binary_average_precision_exact(f, t) for f, t, _ in curves]
    np.testing.assert_allclose(ap, c["ap_none"], rtol=1e-6, atol=1e-7)


# ------------------------------------------------------------------------------------------ known answers
# Hard-coded expectations copied as DATA from the reference's tests / doctests (file:line in each case).
def test_known_answer_confmat_doctest():
    # functional/classification/confusion_matrix.py:398-408
    preds = np.array([[0.16, 0.26, 0.58], [0.22, 0.61, 0.17], [0.71, 0.09, 0.20], [0.05, 0.82, 0.13]], np.float32)
    target = np.array([2, 1, 0, 0])
    assert O.multiclass_confusion_matrix(preds, target, 3).tolist() == [[1, 1, 0], [0, 1, 0], [0, 0, 1]]
    # :584-596 multilabel
    target = np.array([[0, 1, 0], [1, 0, 1]])
    preds = np.array([[0.11, 0.22, 0.84], [0.73, 0.33, 0.92]], np.float32)
    want = [[[1, 0], [0, 1]], [[1, 0], [1, 0]], [[0, 1], [0, 1]]]
    assert O.multilabel_confusion_matrix(preds, target, 3).tolist() == want
    # :213-219 binary
    assert O.binary_confusion_matrix(np.array([0.35, 0.85, 0.48, 0.01], np.float32), np.array([1, 1, 0, 0])).tolist() == [
        [2, 0], [1, 1]]


def test_known_answer_uint8_overflow():
    # tests/unittests/classification/test_confusion_matrix.py:238-247 and test_stat_scores.py:367-382:
    # uint8 inputs with 20 classes must not overflow in target * C + pred
    preds = np.arange(20, dtype=np.uint8)
    target = np.arange(20, dtype=np.uint8)
    assert np.array_equal(O.multiclass_confusion_matrix(preds, target, 20), np.eye(20, dtype=np.int64))
    tp, fp, tn, fn = O.multiclass_stat_scores(preds, target, 20, "micro")
    assert (tp, fp, tn, fn) == (20, 0, 380, 0)


def test_known_answer_stat_scores_doctest():
    # functional/classification/stat_scores.py:540-552 (multiclass doctest, float preds)
    target = np.array([2, 1, 0, 0])
    preds = np.array([[0.16, 0.26, 0.58], [0.22, 0.61, 0.17], [0.71, 0.09, 0.20], [0.05, 0.82, 0.13]], np.float32)
    tp, fp, tn, fn = O.multiclass_stat_scores(preds, target, 3, "micro")
    assert [tp, fp, tn, fn, tp + fn] == [3, 1, 7, 1, 4]
    tp, fp, tn, fn = O.multiclass_stat_scores(preds, target, 3, None)
    assert np.stack([tp, fp, tn, fn, tp + fn], -1).tolist() == [[1, 0, 2, 1, 2], [1, 1, 2, 0, 1], [1, 0, 3, 0, 1]]
    # :184-192 binary doctest
    assert O.binary_stat_scores(np.array([0.11, 0.22, 0.84, 0.73, 0.33, 0.92], np.float32),
                                np.array([0, 1, 0, 1, 0, 1])) == (2, 1, 2, 1)


def test_known_answer_jaccard_absent_class():
    # tests/unittests/classification/test_jaccard.py:347-367: a class absent from preds and target is
    # excluded from the macro average
    target = np.array([0, 1, 1, 0, 0])
    preds = np.array([0, 1, 1, 0, 1])
    confmat = O.multiclass_confusion_matrix(preds, target, 3)
    np.testing.assert_allclose(O.jaccard_reduce(confmat, "macro"), (2 / 3 + 2 / 3) / 2, rtol=1e-6)


def test_known_answer_binned_pr_curve_doctest():
    # functional/classification/precision_recall_curve.py:340-351
    preds = np.array([0, 0.5, 0.7, 0.8], np.float32)
    target = np.array([0, 1, 1, 0])
    state = O.binary_curve_binned(preds, target, np.linspace(0, 1, 5, dtype=np.float32))
    tps, fps, fns = state[:, 1, 1], state[:, 0, 1], state[:, 1, 0]
    precision = np.concatenate([O._safe_divide(tps, tps + fps), [1.0]])
    recall = np.concatenate([O._safe_divide(tps, tps + fns), [0.0]])
    np.testing.assert_allclose(precision, [0.5, 0.6667, 0.6667, 0.0, 0.0, 1.0], atol=1e-4)
    np.testing.assert_allclose(recall, [1.0, 1.0, 1.0, 0.0, 0.0, 0.0], atol=1e-4)


def test_known_answer_auroc_single_class_is_zero():
    # tests/unittests/classification/test_auroc.py:400-411: only one class present -> score 0
    fps, tps, _ = O.binary_clf_curve(np.random.RandomState(0).rand(20).astype(np.float32), np.zeros(20, np.int64))
    assert O.binary_auroc_exact(fps, tps) == 0.0


def test_argmax_rule_matches_torch():
    # SURVEY H7 probe vector
    import torch

    x = np.array([[1, 1, 0], [0, np.nan, 5], [np.nan, np.nan, 1], [-np.inf, -np.inf, -np.inf]], np.float32)
    assert O.argmax_first(x, 1).tolist() == [0, 1, 0, 0] == torch.from_numpy(x).argmax(1).tolist()


def test_nan_goes_below_every_threshold():
    # every `>=` against NaN is False
    preds = np.array([np.nan, 0.2, 0.9], np.float32)
    state = O.binary_curve_binned(np.clip(preds, 0, 1), np.array([1, 0, 1]), np.array([0.0, 0.5, 1.0], np.float32))
    # NaN makes the [0,1] test fail -> sigmoid; NaN stays NaN and lands in [target=1, pred=0] for every threshold
    assert state[:, 1, 0].tolist()[0] >= 1


def test_sync_semantics():
    a, b = np.arange(6).reshape(2, 3), np.ones((2, 3), np.int64)
    assert np.array_equal(O.sync_sum([a, b]), a + b)
    assert np.array_equal(O.sync_cat([a[:1], None, b]), np.concatenate([a[:1], b]))
    assert O.sync_cat([None, None]) is None
