"""SURVEY §8 row f3 on the CPU: the oracle's top-k stat scores against the reference's golden vectors."""
import numpy as np
import pytest
from conftest import golden_names, ignore_of, load_golden

import hotpath as O


@pytest.mark.parametrize("name", golden_names("f3_topk_"))
def test_topk_stat_scores_oracle_matches_reference(name):
    c = load_golden(name)
    C, k, ign = int(c["num_classes"]), int(c["top_k"]), ignore_of(c)
    tp, fp, tn, fn = O.multiclass_stat_scores_topk(c["preds"], c["target"], C, k, ign)
    assert np.array_equal(np.stack([tp, fp, tn, fn, tp + fn], -1), c["stat_scores_none"])
    assert np.array_equal(np.stack([tp, fp, tn, fn, tp + fn], -1).sum(0), c["stat_scores_micro"])
    np.testing.assert_allclose(tp.sum() / (tp.sum() + fn.sum()), c["accuracy_micro"], rtol=1e-6)


@pytest.mark.parametrize("name", golden_names("f3_samplewise_"))
def test_samplewise_oracle_matches_reference(name):
    c = load_golden(name)
    ign = ignore_of(c)
    if "binary" in name:
        tp, fp, tn, fn = O.binary_stat_scores_samplewise(c["preds"], c["target"], 0.5, ign)
        assert np.array_equal(np.stack([tp, fp, tn, fn, tp + fn], 1), c["stat_scores"])
    elif "multilabel" in name:
        tp, fp, tn, fn = O.multilabel_stat_scores_samplewise(c["preds"], c["target"], int(c["num_labels"]), 0.5, ign)
        assert np.array_equal(np.stack([tp, fp, tn, fn, tp + fn], -1), c["stat_scores_none"])
    else:
        tp, fp, tn, fn = O.multiclass_stat_scores_samplewise(c["preds"], c["target"], int(c["num_classes"]), ign)
        assert np.array_equal(np.stack([tp, fp, tn, fn, tp + fn], -1), c["stat_scores_none"])
        assert np.array_equal(np.stack([tp, fp, tn, fn, tp + fn], -1).sum(1), c["stat_scores_micro"])


@pytest.mark.parametrize("name", golden_names("f3_swtopk_"))
def test_samplewise_topk_oracle_matches_reference(name):
    c = load_golden(name)
    C, k, ign = int(c["num_classes"]), int(c["top_k"]), ignore_of(c)
    tp, fp, tn, fn = O.multiclass_stat_scores_samplewise_topk(c["preds"], c["target"], C, k, ign)
    assert np.array_equal(np.stack([tp, fp, tn, fn, tp + fn], -1), c["stat_scores_none"])
    assert np.array_equal(np.stack([tp, fp, tn, fn, tp + fn], -1).sum(1), c["stat_scores_micro"])
