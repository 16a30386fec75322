"""GPU parity for SURVEY §8 row f3: stat scores with `top_k > 1` (top-k selection kernel) against the reference's
golden vectors, the oracle, and — for ties and NaNs — torch.topk on the device."""
import numpy as np
import pytest
import torch
from conftest import golden_names, ignore_of, load_golden

import hotpath as O

pytestmark = pytest.mark.gpu
AVERAGES = ("micro", "macro", "weighted", "none")


@pytest.mark.parametrize("name", golden_names("f3_topk_"))
def test_topk_golden(name):
    from torchmetrics_forked_b200 import classification as C
    from torchmetrics_forked_b200 import functional as F

    c = load_golden(name)
    nC, k, ign = int(c["num_classes"]), int(c["top_k"]), ignore_of(c)
    preds, target = torch.from_numpy(c["preds"]).cuda(), torch.from_numpy(c["target"]).cuda()
    for avg in AVERAGES:
        got = F.multiclass_stat_scores(preds, target, nC, average=avg, top_k=k, ignore_index=ign)
        want = c[f"stat_scores_{avg}"]
        if avg in ("micro", "none"):
            assert np.array_equal(got.cpu().numpy(), want), (name, avg)
        else:
            np.testing.assert_allclose(got.cpu().numpy(), want, rtol=1e-6)
        for fn in ("accuracy", "precision", "recall", "f1_score", "specificity", "hamming_distance"):
            got = getattr(F, f"multiclass_{fn}")(preds, target, nC, average=avg, top_k=k, ignore_index=ign)
            np.testing.assert_allclose(got.cpu().numpy(), c[f"{fn}_{avg}"], rtol=1e-6, atol=1e-7, err_msg=f"{fn} {avg}")
    # module API, two updates; the micro state stays per class for top_k > 1 (reference :318-320)
    m = C.MulticlassAccuracy(nC, top_k=k, average="micro", ignore_index=ign).cuda()
    s = C.StatScores(task="multiclass", num_classes=nC, top_k=k, average="none", ignore_index=ign).cuda()
    h = preds.shape[0] // 3
    for sl in (slice(0, h), slice(h, None)):
        m.update(preds[sl], target[sl])
        s.update(preds[sl], target[sl])
    assert m.tp.shape == (nC,)
    np.testing.assert_allclose(float(m.compute()), c["accuracy_micro"], rtol=1e-6)
    assert np.array_equal(s.compute().cpu().numpy(), c["stat_scores_none"])


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16])
@pytest.mark.parametrize("C,k", [(2, 2), (5, 2), (8, 3), (33, 7), (100, 5), (264, 9), (520, 5), (1000, 5), (1024, 20)])
def test_topk_ties_and_special_values_follow_torch_topk(C, k, dtype):
    """Heavily tied scores (16-bit dtypes, quantised values), NaN and +-inf entries: the selected sets must be the
    ones torch.topk picks on this device, and the oracle's rule (value desc, index asc, NaN first) must agree."""
    from torchmetrics_forked_b200 import functional as F

    g = torch.Generator().manual_seed(C * 31 + k)
    n = 3001 if C <= 100 else 257
    preds = ((torch.randn(n, C, generator=g) * 2).round() / 2).to(dtype)  # few distinct values: ties everywhere
    preds[0, 0] = float("nan")
    preds[1, C - 1] = float("inf")
    preds[2, :] = float("-inf")
    preds[3, :] = 0.0
    if C > 1:
        preds[3, 1] = -0.0
    preds[4, C // 2] = -float("nan")  # sign bit set
    preds[5, :] = float("inf")
    preds[5, 0] = float("nan")
    target = torch.randint(C, (n,), generator=g)
    target[::11] = -1
    dp, dt = preds.cuda(), target.cuda()
    got = F.multiclass_stat_scores(dp, dt, C, average="none", top_k=k, ignore_index=-1).cpu().numpy()
    # torch.topk on the device -> k-hot -> the reference's one-hot arithmetic
    keep = dt != -1
    member = torch.zeros(n, C, dtype=torch.bool, device="cuda").scatter_(1, dp.topk(k, dim=1).indices, True)[keep]
    hot = torch.nn.functional.one_hot(dt[keep], C).bool()
    want = torch.stack([(hot & member).sum(0), (~hot & member).sum(0), (~hot & ~member).sum(0), (hot & ~member).sum(0),
                        hot.sum(0)], -1).cpu().numpy()
    assert np.array_equal(got, want), (C, k, dtype)
    tp, fp, tn, fn = O.multiclass_stat_scores_topk(preds.float().numpy(), target.numpy(), C, k, -1)
    assert np.array_equal(got, np.stack([tp, fp, tn, fn, tp + fn], -1))


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("C,k", [(136, 3), (264, 5), (520, 9), (1000, 5), (1000, 1), (1024, 20), (1000, 32), (1000, 40)])
def test_topk_natural_scores_follow_torch_topk(C, k, dtype):
    """Unquantised scores: the rows where the k largest sit in k lanes, the rows where two of them share a lane, and
    (16-bit dtypes) the occasional tie across the k-th place all take the lane-maxima route of the top-k kernels; k = 32
    is its limit, k = 40 the full ranking.  A few rows are made adversarial: the k largest in consecutive classes (one
    lane), a plateau of equal scores around the k-th place."""
    from torchmetrics_forked_b200 import functional as F

    g = torch.Generator().manual_seed(C * 7 + k)
    n = 4099
    preds = torch.randn(n, C, generator=g).to(dtype)
    preds[0, 8:8 + min(k + 2, C - 8)] = 9.0 + torch.arange(min(k + 2, C - 8)).to(dtype) * 0.25  # one lane holds them all
    preds[1, ::3] = 7.0  # a plateau far wider than k
    preds[2, 5:5 + k + 1] = 6.5  # a plateau of k + 1: one of them stays out
    preds[3, C - k:] = 8.0  # exactly k equal winners at the end of the row
    target = torch.randint(C, (n,), generator=g)
    dp, dt = preds.cuda(), target.cuda()
    got = F.multiclass_stat_scores(dp, dt, C, average="none", top_k=k).cpu().numpy()
    member = torch.zeros(n, C, dtype=torch.bool, device="cuda").scatter_(1, dp.topk(k, dim=1).indices, True)
    hot = torch.nn.functional.one_hot(dt, C).bool()
    want = torch.stack([(hot & member).sum(0), (~hot & member).sum(0), (~hot & ~member).sum(0), (hot & ~member).sum(0),
                        hot.sum(0)], -1).cpu().numpy()
    assert np.array_equal(got, want), (C, k, dtype)


def test_topk_argument_handling():
    from torchmetrics_forked_b200 import classification as C
    from torchmetrics_forked_b200 import functional as F

    p, t = torch.randn(10, 4).cuda(), torch.randint(4, (10,)).cuda()
    with pytest.raises(ValueError, match="top_k"):
        F.multiclass_accuracy(p, t, 4, top_k=5)
    with pytest.raises(ValueError, match="top_k"):
        F.multiclass_accuracy(t, t, 4, top_k=2)  # label-valued preds cannot be ranked
    C.MulticlassStatScores(4, multidim_average="samplewise", top_k=2)  # the combination is built (round 2)
    bad = t.clone()
    bad[0] = 9
    with pytest.raises(RuntimeError, match="target"):
        F.multiclass_accuracy(p, bad, 4, top_k=2)


@pytest.mark.parametrize("name", golden_names("f3_swtopk_"))
def test_samplewise_topk_golden(name):
    """`top_k > 1` together with `multidim_average="samplewise"` (round 2): per-sample counts bit for bit against the
    reference's golden vectors and the oracle, derived scores within 1e-6, module API over two updates."""
    from torchmetrics_forked_b200 import classification as C
    from torchmetrics_forked_b200 import functional as F

    c = load_golden(name)
    nC, k, ign = int(c["num_classes"]), int(c["top_k"]), ignore_of(c)
    preds, target = torch.from_numpy(c["preds"]).cuda(), torch.from_numpy(c["target"]).cuda()
    kw = dict(top_k=k, ignore_index=ign, multidim_average="samplewise")
    tp, fp, tn, fn = O.multiclass_stat_scores_samplewise_topk(c["preds"], c["target"], nC, k, ign)
    want = np.stack([tp, fp, tn, fn, tp + fn], -1)
    assert np.array_equal(want, c["stat_scores_none"])
    assert np.array_equal(F.multiclass_stat_scores(preds, target, nC, average="none", **kw).cpu().numpy(), want)
    for dt in (torch.bfloat16, torch.float16):  # narrow dtypes: compare with the oracle on the rounded scores
        q = preds.to(dt)
        o = O.multiclass_stat_scores_samplewise_topk(q.float().cpu().numpy(), c["target"], nC, k, ign)
        got = F.multiclass_stat_scores(q, target, nC, average="none", **kw).cpu().numpy()
        # ties at the k-th place (possible after rounding) go to the lower class in both
        assert np.array_equal(got, np.stack([o[0], o[1], o[2], o[3], o[0] + o[3]], -1)), dt
    for avg in AVERAGES:
        for fn_ in ("stat_scores", "accuracy", "precision", "recall", "f1_score", "specificity", "hamming_distance"):
            got = getattr(F, f"multiclass_{fn_}")(preds, target, nC, average=avg, **kw)
            np.testing.assert_allclose(got.cpu().numpy(), c[f"{fn_}_{avg}"], rtol=1e-6, atol=1e-7, err_msg=f"{fn_} {avg}")
    m = C.MulticlassStatScores(nC, average="none", **kw).cuda()
    h = preds.shape[0] // 2
    m.update(preds[:h], target[:h])
    m.update(preds[h:], target[h:])
    assert isinstance(m.tp, list) and len(m.tp) == 2
    assert np.array_equal(m.compute().cpu().numpy(), want)


def test_topk_wide_rows_follow_torch_topk():
    """C > 1024 (the row no longer fits the register kernels): ties, NaN and infinities as torch.topk on the device,
    every dtype; the shared-memory histogram (C = 1500) and the global-atomic one (C = 20000) both."""
    from torchmetrics_forked_b200 import functional as F

    for C_, k, n in ((1500, 5, 129), (4099, 3, 33), (20_000, 2, 17)):
        g = torch.Generator().manual_seed(C_)
        for dtype in (torch.float32, torch.bfloat16, torch.float16):
            preds = ((torch.randn(n, C_, generator=g) * 2).round() / 2).to(dtype)
            preds[0, 0] = float("nan")
            preds[1, C_ - 1] = float("inf")
            preds[2, :] = float("-inf")
            target = torch.randint(C_, (n,), generator=g)
            got = F.multiclass_stat_scores(preds.cuda(), target.cuda(), C_, average="none", top_k=k).cpu().numpy()
            tp, fp, tn, fn = O.multiclass_stat_scores_topk(preds.float().numpy(), target.numpy(), C_, k)
            assert np.array_equal(got, np.stack([tp, fp, tn, fn, tp + fn], -1)), (C_, dtype)
            idx = preds.cuda().topk(k, dim=1).indices.cpu()
            hits = (idx == target[:, None]).any(1).sum()
            assert int(got[:, 0].sum()) == int(hits)


# ------------------------------------------------------------------------------------------ samplewise
def cuda_sigmoid(x: np.ndarray) -> np.ndarray:
    return torch.from_numpy(np.ascontiguousarray(x, dtype=np.float32)).cuda().sigmoid().cpu().numpy()


@pytest.mark.parametrize("name", golden_names("f3_samplewise_"))
def test_samplewise_golden(name):
    """`multidim_average="samplewise"`: per-sample counts bit for bit (logits: against the oracle with the device
    sigmoid, H5), derived scores within 1e-6, functional and module API (list states over two updates)."""
    from torchmetrics_forked_b200 import classification as C
    from torchmetrics_forked_b200 import functional as F

    c = load_golden(name)
    ign = ignore_of(c)
    task = name.split("_")[2]
    preds, target = torch.from_numpy(c["preds"]).cuda(), torch.from_numpy(c["target"]).cuda()
    logits = "logits" in name and task != "multiclass"  # argmax needs no sigmoid: multiclass is exact on any device
    kw = {"ignore_index": ign, "multidim_average": "samplewise"}
    if task == "binary":
        tp, fp, tn, fn = O.binary_stat_scores_samplewise(c["preds"], c["target"], 0.5, ign, cuda_sigmoid)
        want = np.stack([tp, fp, tn, fn, tp + fn], 1)
        got = F.binary_stat_scores(preds, target, **kw)
        assert np.array_equal(got.cpu().numpy(), want)
        if not logits:
            assert np.array_equal(want, c["stat_scores"])
            for fn_ in ("accuracy", "precision", "recall", "f1_score", "specificity", "hamming_distance"):
                np.testing.assert_allclose(getattr(F, f"binary_{fn_}")(preds, target, **kw).cpu().numpy(), c[fn_],
                                           rtol=1e-6, atol=1e-7)
        m = C.BinaryStatScores(**kw).cuda()
        args = {}
    else:
        n_key = "num_classes" if task == "multiclass" else "num_labels"
        n_cls = int(c[n_key])
        if task == "multiclass":
            tp, fp, tn, fn = O.multiclass_stat_scores_samplewise(c["preds"], c["target"], n_cls, ign)
        else:
            tp, fp, tn, fn = O.multilabel_stat_scores_samplewise(c["preds"], c["target"], n_cls, 0.5, ign, cuda_sigmoid)
        want = np.stack([tp, fp, tn, fn, tp + fn], -1)
        got = getattr(F, f"{task}_stat_scores")(preds, target, n_cls, average="none", **kw)
        assert np.array_equal(got.cpu().numpy(), want)
        if not logits:
            assert np.array_equal(want, c["stat_scores_none"])
            for avg in AVERAGES:
                for fn_ in ("stat_scores", "accuracy", "precision", "recall", "f1_score", "specificity", "hamming_distance"):
                    got = getattr(F, f"{task}_{fn_}")(preds, target, n_cls, average=avg, **kw)
                    np.testing.assert_allclose(got.cpu().numpy(), c[f"{fn_}_{avg}"], rtol=1e-6, atol=1e-7,
                                               err_msg=f"{fn_} {avg}")
        m = getattr(C, f"{task.capitalize()}StatScores")(n_cls, average="none", **kw).cuda()
    h = preds.shape[0] // 2
    m.update(preds[:h], target[:h])
    m.update(preds[h:], target[h:])
    assert isinstance(m.tp, list) and len(m.tp) == 2
    if not logits:  # two updates take two sigmoid decisions; with probabilities / argmax they agree with one
        assert np.array_equal(m.compute().cpu().numpy(), want)


def test_samplewise_many_samples_go_through_in_slices():
    """More samples than one 256 MB scratch holds ([3C] int64 per sample): the wrapper slices the batch."""
    from torchmetrics_forked_b200 import functional as F

    g = torch.Generator().manual_seed(3)
    C, n = 1000, 12_000  # 11184 samples per slice at C = 1000
    preds, target = torch.randint(C, (n, 6), generator=g), torch.randint(C, (n, 6), generator=g)
    got = F.multiclass_stat_scores(preds.cuda(), target.cuda(), C, average="micro", multidim_average="samplewise")
    tp = (preds == target).sum(1)
    want = torch.stack([tp, 6 - tp, 6 * C - 12 + tp, 6 - tp, torch.full_like(tp, 6)], 1)
    assert torch.equal(got.cpu(), want)
