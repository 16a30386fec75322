"""Pin both oracles against the LIVE reference on random inputs (build container only: /root/reference is not
on the GPU box, so this whole module skips there)."""
import numpy as np
import pytest
import torch

import hotpath as O
import torch_cpu as P
from ref_loader import load_reference, reference_available

pytestmark = pytest.mark.skipif(not reference_available(), reason="/root/reference is not mounted")


@pytest.fixture(scope="module")
def ref():
    load_reference()
    import torchmetrics.functional.classification as rf

    return rf


@pytest.mark.parametrize("seed", range(4))
def test_multiclass_random(ref, seed):
    g = torch.Generator().manual_seed(seed)
    C = [3, 17, 150, 1000][seed]
    shape = [(500, C), (64, C, 9), (10, C, 8, 8), (200, C)][seed]
    preds = torch.randn(*shape, generator=g)
    if seed == 3:
        preds = preds.bfloat16()
    target = torch.randint(C, (shape[0], *shape[2:]), generator=g)
    ign = [None, 2, None, 5][seed]
    want = ref.multiclass_confusion_matrix(preds, target, C, ignore_index=ign, validate_args=False).numpy()
    assert np.array_equal(O.multiclass_confusion_matrix(preds.float().numpy(), target.numpy(), C, ign), want)
    assert np.array_equal(P.multiclass_confusion_matrix_update(preds, target, C, ign).numpy(), want)
    for avg in ("micro", None):
        r = ref.multiclass_stat_scores(preds, target, C, average=avg, ignore_index=ign, validate_args=False).numpy()
        tp, fp, tn, fn = O.multiclass_stat_scores(preds.float().numpy(), target.numpy(), C, avg, ign)
        assert np.array_equal(np.stack([tp, fp, tn, fn, tp + fn], -1), r)
        tp, fp, tn, fn = P.multiclass_stat_scores_update(preds, target, C, avg, ign)
        assert np.array_equal(torch.stack([tp, fp, tn, fn, tp + fn], -1).numpy(), r)


@pytest.mark.parametrize("logits", [False, True])
def test_binary_curve_random(ref, logits):
    from torchmetrics.functional.classification.precision_recall_curve import (
        _binary_precision_recall_curve_format,
        _binary_precision_recall_curve_update,
    )

    g = torch.Generator().manual_seed(7)
    for n in (3000, 70_000):
        preds = torch.randn(n, generator=g) * 3 if logits else torch.rand(n, generator=g)
        target = torch.randint(2, (n,), generator=g)
        p, t, thr = _binary_precision_recall_curve_format(preds, target, 100, None)
        want = _binary_precision_recall_curve_update(p, t, thr).numpy()
        assert np.array_equal(O.binary_curve_binned(preds.numpy(), target.numpy(), thr.numpy()), want)
        assert np.array_equal(P.binary_curve_update(preds, target, thr).numpy(), want)


def test_clf_curve_random(ref):
    from torchmetrics.functional.classification.precision_recall_curve import _binary_clf_curve

    g = torch.Generator().manual_seed(9)
    preds = (torch.rand(20_000, generator=g) * 50).round() / 50
    target = torch.randint(2, (20_000,), generator=g)
    want = _binary_clf_curve(preds, target)
    got = O.binary_clf_curve(preds.numpy(), target.numpy())
    port = P.binary_clf_curve(preds, target)
    for w, a, b in zip(want, got, port):
        assert np.array_equal(w.numpy(), a) and torch.equal(w, b)
    auroc = ref.binary_auroc(preds, target, validate_args=False)
    np.testing.assert_allclose(O.binary_auroc_exact(got[0], got[1]), auroc.numpy(), rtol=1e-6)


def test_multilabel_exact_port(ref):
    g = torch.Generator().manual_seed(10)
    preds = torch.randn(2000, 6, generator=g)
    target = torch.randint(2, (2000, 6), generator=g)
    want = ref.multilabel_auroc(preds, target, 6, validate_args=False)
    p, t = P.multilabel_exact_update(preds, target, 6)
    np.testing.assert_allclose(P.multilabel_auroc_exact_compute(p, t, 6).numpy(), want.numpy(), rtol=1e-6)
