"""SURVEY §8 row f1 on the CPU: the new reducers (pure torch expressions over the kernels' integer states) against
golden vectors produced by the unmodified reference (oracle/gen_golden_f1.py).  Tolerance 1e-6 relative; values
that are exact in the reference (0, 1, -1, 1e6 sentinels) must be exact here."""
import numpy as np
import pytest
import torch
from conftest import load_golden

from torchmetrics_forked_b200.functional.classification._fixed_point import (
    _precision_at_recall,
    _recall_at_precision,
    _specificity_at_sensitivity,
)
from torchmetrics_forked_b200.functional.classification._stat_family import (
    _hamming_distance_reduce,
    _precision_reduce,
    _recall_reduce,
    _specificity_reduce,
)
from torchmetrics_forked_b200.functional.classification.cohen_kappa import _cohen_kappa_reduce
from torchmetrics_forked_b200.functional.classification.matthews_corrcoef import _matthews_corrcoef_reduce

RTOL = 1e-6
REDUCERS = {"precision": _precision_reduce, "recall": _recall_reduce, "specificity": _specificity_reduce,
            "hamming": _hamming_distance_reduce}


def _close(got, want):
    np.testing.assert_allclose(np.asarray(got, dtype=np.float64), np.asarray(want, dtype=np.float64), rtol=RTOL,
                               atol=1e-7, equal_nan=True)


@pytest.mark.parametrize("name", sorted(REDUCERS))
def test_stat_score_reducers_match_reference(name):
    c = load_golden("f1_stat_reducers")
    for i in range(int(c["n_states"])):
        tp, fp, tn, fn = torch.from_numpy(c[f"state_{i}"])
        if tp.numel() == 1:
            _close(REDUCERS[name](tp, fp, tn, fn, average="binary").numpy(), c[f"{name}_{i}_binary"])
        for avg in ("micro", "macro", "weighted", "none"):
            for ml in (False, True):
                got = REDUCERS[name](tp, fp, tn, fn, average=avg, multilabel=ml)
                want = c[f"{name}_{i}_{avg}_{int(ml)}"]
                assert got.shape == want.shape, (name, i, avg, ml)
                _close(got.numpy(), want)


def test_confusion_matrix_reducers_match_reference():
    c = load_golden("f1_confmat_reducers")
    for i in range(int(c["n_mats"])):
        m = torch.from_numpy(c[f"mat_{i}"])
        got = _matthews_corrcoef_reduce(m)
        assert str(got.dtype) == str(c[f"mcc_dtype_{i}"]), i  # +-1 / 0 come back in the state's integer dtype
        _close(got.float().numpy(), c[f"mcc_{i}"])
        if m.ndim == 2:
            for w in ("none", "linear", "quadratic"):
                _close(_cohen_kappa_reduce(m, w).numpy(), c[f"kappa_{w}_{i}"])
            _close(_cohen_kappa_reduce(m, None).numpy(), c[f"kappa_none_{i}"])
    with pytest.raises(ValueError, match="weights"):
        _cohen_kappa_reduce(torch.eye(3), "cubic")


def test_operating_points_match_reference():
    c = load_golden("f1_fixed_points")
    for i in range(int(c["n_curves"])):
        p, r, t = (torch.from_numpy(c[f"{k}_{i}"]) for k in "prt")
        for j, lv in enumerate(c["levels"]):
            lv = float(lv)
            for key, fn, args in (("rap", _recall_at_precision, (p, r, t)), ("par", _precision_at_recall, (p, r, t)),
                                  ("sas", _specificity_at_sensitivity, (p[: len(t)], r[: len(t)], t))):
                value, threshold = fn(*args, lv)
                want = c[f"{key}_{i}_{j}"]
                assert float(value) == pytest.approx(want[0], rel=RTOL, abs=0), (key, i, j)
                assert float(threshold) == pytest.approx(want[1], rel=RTOL, abs=0), (key, i, j)


def test_task_wrappers_and_argument_errors():
    from torchmetrics_forked_b200 import classification as C
    from torchmetrics_forked_b200 import functional as F

    assert type(C.Precision(task="binary")).__name__ == "BinaryPrecision"
    assert type(C.Recall(task="multiclass", num_classes=3)).__name__ == "MulticlassRecall"
    assert type(C.Specificity(task="multilabel", num_labels=3)).__name__ == "MultilabelSpecificity"
    assert C.HammingDistance(task="binary").higher_is_better is False
    assert type(C.CohenKappa(task="multiclass", num_classes=3, weights="linear")).__name__ == "MulticlassCohenKappa"
    assert type(C.MatthewsCorrCoef(task="multilabel", num_labels=4)).__name__ == "MultilabelMatthewsCorrCoef"
    assert C.RecallAtFixedPrecision(task="binary", min_precision=0.5, thresholds=10).min_precision == 0.5
    assert C.PrecisionAtFixedRecall(task="multilabel", num_labels=3, min_recall=0.2).min_recall == 0.2
    assert C.SpecificityAtSensitivity(task="binary", min_sensitivity=0.7).min_sensitivity == 0.7
    with pytest.raises(ValueError, match="num_classes"):
        C.Precision(task="multiclass")
    with pytest.raises(ValueError, match="num_labels"):
        C.MatthewsCorrCoef(task="multilabel")
    with pytest.raises(ValueError):
        C.CohenKappa(task="multilabel", num_labels=3)
    with pytest.raises(ValueError, match="weight"):
        C.BinaryCohenKappa(weights="cubic")
    m = C.RecallAtFixedPrecision(task="multiclass", min_precision=0.5, num_classes=3, thresholds=4)
    assert type(m).__name__ == "MulticlassRecallAtFixedPrecision" and m.confmat.shape == (4, 3, 2, 2)
    assert F.specicity_at_sensitivity is F.specificity_at_sensitivity  # the reference's spelling is kept as an alias
    import inspect

    assert "min_recall" in inspect.signature(F.binary_precision_at_fixed_recall).parameters
    assert list(inspect.signature(F.multiclass_precision).parameters)[:5] == [
        "preds", "target", "num_classes", "average", "top_k"]
    # generated classes pickle like hand-written ones
    import pickle

    m = pickle.loads(pickle.dumps(C.MulticlassPrecision(4, average="weighted")))
    assert type(m) is C.MulticlassPrecision and m.average == "weighted"
