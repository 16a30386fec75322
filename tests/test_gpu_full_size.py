"""GPU parity at BASELINE.json's full sizes, through size-independent properties and independent torch computations
on the device (the Python oracle would take minutes at these sizes):

* config 2  MulticlassConfusionMatrix(1000), bf16 [2^22, 1000]: row sums = bincount(target), column sums =
            bincount(torch.argmax), diagonal = matches, total = N
* config 3  BinaryPrecisionRecallCurve(thresholds=200), 2^28 fp32 logits: every threshold's 2x2 matrix sums to N,
            counts are monotone in the threshold, and each cell equals a torch.bucketize histogram of torch.sigmoid
* config 4  MultilabelAUROC(100, thresholds=None), 2^24 x 100 split over 8 updates: AUROC of sampled labels equals the
            rank statistic (Mann-Whitney U with ties) computed by torch.sort in float64
* config 5  MulticlassJaccardIndex(150), [16, 150, 512, 512]: confusion matrix equals torch.bincount of
            (target * C + argmax)
"""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _need(gib: float):
    free, _ = torch.cuda.mem_get_info()
    if free < gib * (1 << 30):
        pytest.skip(f"needs {gib} GiB of free device memory")


def test_config2_full_size_confusion_matrix():
    from torchmetrics_forked_b200.classification import MulticlassConfusionMatrix

    _need(24)
    g = torch.Generator(device="cuda").manual_seed(2)
    n, C = 1 << 22, 1000
    preds = torch.empty(n, C, dtype=torch.bfloat16, device="cuda")
    for s in range(0, n, 1 << 19):
        preds[s : s + (1 << 19)] = torch.randn(1 << 19, C, generator=g, device="cuda").bfloat16()
    target = torch.randint(C, (n,), generator=g, device="cuda")
    m = MulticlassConfusionMatrix(C, validate_args=False).cuda()
    m.update(preds, target)
    cm = m.confmat
    am = torch.cat([preds[s : s + (1 << 19)].argmax(1) for s in range(0, n, 1 << 19)])
    assert int(cm.sum()) == n
    assert torch.equal(cm.sum(1), torch.bincount(target, minlength=C))
    assert torch.equal(cm.sum(0), torch.bincount(am, minlength=C))
    assert torch.equal(cm, torch.bincount(target * C + am, minlength=C * C).reshape(C, C))


def test_config3_full_size_binned_curve():
    from torchmetrics_forked_b200.classification import BinaryPrecisionRecallCurve

    _need(16)
    g = torch.Generator(device="cuda").manual_seed(3)
    n, T = 1 << 28, 200
    logits = torch.randn(n, generator=g, device="cuda") * 3
    target = torch.randint(2, (n,), generator=g, device="cuda")
    m = BinaryPrecisionRecallCurve(thresholds=T, validate_args=False).cuda()
    m.update(logits, target)
    cm = m.confmat  # [T, 2, 2], cell [t, target, pred >= thr_t]
    assert torch.equal(cm.sum((1, 2)), torch.full((T,), n, device="cuda"))
    assert bool((cm[1:, :, 1] <= cm[:-1, :, 1]).all())  # fewer predicted positives as the threshold grows
    thr = torch.linspace(0, 1, T).cuda()
    want = torch.zeros(T, 2, 2, dtype=torch.long, device="cuda")
    for s in range(0, n, 1 << 26):  # torch.sigmoid on the device is what the reference would run (H5)
        q = logits[s : s + (1 << 26)].sigmoid()
        k = torch.bucketize(q, thr, right=True)  # number of thresholds <= q
        for cls in (0, 1):
            hist = torch.bincount(k[target[s : s + (1 << 26)] == cls], minlength=T + 1)
            ge = hist.flip(0).cumsum(0).flip(0)[1:]  # samples with q >= thr_t
            want[:, cls, 1] += ge
            want[:, cls, 0] += hist.sum() - ge
    assert torch.equal(cm, want)


def test_config4_full_size_exact_auroc():
    from torchmetrics_forked_b200.classification import MultilabelAUROC

    _need(60)
    g = torch.Generator(device="cuda").manual_seed(4)
    n, L, upd = 1 << 21, 100, 8
    m = MultilabelAUROC(L, thresholds=None, average="none", validate_args=False).cuda()
    keep = {}
    for _ in range(upd):
        p = torch.randn(n, L, generator=g, device="cuda")
        p[:, 7] = (p[:, 7] * 8).round() / 8  # one heavily tied label
        t = torch.randint(2, (n, L), generator=g, device="cuda")
        m.update(p, t)
        for l in (0, 7, 99):
            keep.setdefault(l, []).append((p[:, l].sigmoid(), t[:, l].clone()))
    got = m.compute()
    for l, parts in keep.items():
        s = torch.cat([a for a, _ in parts]).double()
        y = torch.cat([b for _, b in parts])
        uniq, inv, cnt = torch.unique(s, return_inverse=True, return_counts=True)
        rank = (torch.cumsum(cnt, 0) - cnt + (cnt + 1) / 2.0)[inv]  # average ranks, ties shared
        n_pos = int((y == 1).sum())
        u = rank[y == 1].sum() - n_pos * (n_pos + 1) / 2
        auc = float(u / (n_pos * (s.numel() - n_pos)))
        assert abs(float(got[l]) - auc) < 1e-6, (l, float(got[l]), auc)


def test_config5_full_size_jaccard_state():
    from torchmetrics_forked_b200.classification import MulticlassJaccardIndex

    _need(8)
    g = torch.Generator(device="cuda").manual_seed(5)
    C = 150
    for dtype in (torch.float32, torch.bfloat16):
        preds = torch.randn(16, C, 512, 512, generator=g, device="cuda").to(dtype)
        target = torch.randint(C, (16, 512, 512), generator=g, device="cuda")
        target[torch.rand(16, 512, 512, generator=g, device="cuda") < 0.05] = 255
        m = MulticlassJaccardIndex(C, ignore_index=255, validate_args=False).cuda()
        m.update(preds, target)
        am, keep = preds.argmax(1), target != 255
        want = torch.bincount(target[keep] * C + am[keep], minlength=C * C).reshape(C, C)
        assert torch.equal(m.confmat, want), dtype
