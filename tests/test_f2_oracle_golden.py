"""SURVEY §8 row f2 on the CPU: the oracle's multiclass-curve restatement against golden vectors of the reference
(oracle/gen_golden_f2.py).  The golden states were produced with torch's CPU softmax, which is the oracle's default
`softmax_fn`; integer states and exact curves must match bit for bit."""
import numpy as np
import pytest
from conftest import golden_names, ignore_of, load_golden

import hotpath as O


@pytest.mark.parametrize("name", golden_names("f2_mccurve_"))
def test_multiclass_curve_oracle_matches_reference(name):
    c = load_golden(name)
    C, ign = int(c["num_classes"]), ignore_of(c)
    assert np.array_equal(O.multiclass_curve_binned(c["preds"], c["target"], C, c["thresholds"], ign), c["state"])
    assert np.array_equal(
        O.multiclass_curve_binned(c["preds"], c["target"], C, c["thresholds"], ign, micro=True), c["state_micro"])
    probs, tgt = O.multiclass_rows(c["preds"], c["target"], C, ign)
    assert np.array_equal(probs, c["exact_preds"]) and np.array_equal(tgt, c["exact_target"])
    for k, (fps, tps, thr) in enumerate(O.multiclass_clf_curves(probs, tgt)):
        assert np.array_equal(fps, c[f"fps_{k}"]) and np.array_equal(tps, c[f"tps_{k}"])
        assert np.array_equal(thr, c[f"thr_{k}"])
    want_auroc = [O.binary_auroc_exact(c[f"fps_{k}"], c[f"tps_{k}"]) for k in range(C)]
    np.testing.assert_allclose(want_auroc, c["auroc_none_exact"], rtol=1e-5, atol=1e-7)
