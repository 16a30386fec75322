"""Differential fuzzing against the UNMODIFIED reference on the same GPU.

Seeded random configurations of the metric classes on the path (task, arguments, shapes, dtypes, `ignore_index`,
`thresholds`, averaging, `top_k`, `multidim_average`, extra dimensions, kind of scores), two updates each, through this
package and through `oracle/_ref/reference.zip` on CUDA: integer states `torch.equal`, stored scores equal, `compute()`
within 1e-5 relative / 1e-6 absolute (float32 reductions in another order; the exact-mode AUROC / AP reductions run in
float64 here, DESIGN.md deviation 4).  A configuration the reference itself rejects must be rejected here too.

The generator avoids what DESIGN.md §2 lists as deliberate deviations or reference defects: NaN scores (order of NaN
thresholds), `MulticlassPrecisionRecallCurve(average="micro")` as a class (the reference cannot hold that state),
labels outside the class range (validation differs by design), and a `multidim_average="samplewise"` multilabel batch
of one sample (the reference's `.squeeze()` drops the sample dimension there: deviation 12, found by this test).
"""
import os
import random

import pytest
import torch

import ref_loader

pytestmark = pytest.mark.gpu

if not ref_loader.staged_available() and not ref_loader.reference_available():
    pytest.skip("oracle/_ref is not staged (run __graft_entry__.build() where /root/reference exists)",
                allow_module_level=True)

SCALE = int(os.environ.get("TMB_FUZZ_SCALE", "1"))  # more seeds for a longer hunt: TMB_FUZZ_SCALE=5 pytest ...
N_CASES = 600 * SCALE
OUTCOME = {"compared": 0, "rejected_by_both": 0, "skipped": 0}
STAT_FAMILY = ["StatScores", "Accuracy", "Precision", "Recall", "F1Score", "Specificity", "HammingDistance",
               "JaccardIndex", "ConfusionMatrix", "CohenKappa", "MatthewsCorrCoef"]
CURVE_FAMILY = ["PrecisionRecallCurve", "ROC", "AUROC", "AveragePrecision"]
NO_AVERAGE = {"ConfusionMatrix", "CohenKappa", "MatthewsCorrCoef", "PrecisionRecallCurve", "ROC"}
NO_SAMPLEWISE = {"JaccardIndex", "ConfusionMatrix", "CohenKappa", "MatthewsCorrCoef"}


@pytest.fixture(scope="module")
def R():
    return ref_loader.load_staged_reference()


@pytest.fixture(scope="module")
def M():
    import torchmetrics_forked_b200 as ours
    import torchmetrics_forked_b200.classification  # noqa: F401

    return ours


def _draw(seed):
    """One configuration: (class name, constructor kwargs, batch maker)."""
    rnd = random.Random(seed)
    task = rnd.choice(["Binary", "Multiclass", "Multilabel"])
    curve = rnd.random() < 0.4
    base = rnd.choice(CURVE_FAMILY if curve else STAT_FAMILY)
    name = task + base
    kw = {}
    dtype = rnd.choice([torch.float32, torch.float32, torch.bfloat16, torch.float16, torch.float64])
    n = rnd.choice([0, 1, 7, 33, 257, 1500])
    target_dtype = rnd.choice([torch.int64, torch.int64, torch.int32, torch.int16, torch.int8, torch.uint8, torch.float32,
                               torch.bool])
    strided = rnd.random() < 0.25  # scores and targets as non-contiguous views
    extra = rnd.choice([(), (), (3,), (2, 5)])
    C = rnd.choice([2, 3, 5, 17, 40, 130, 300]) if task == "Multiclass" else rnd.choice([1, 2, 5, 11, 70])
    if task == "Multilabel" and C == 1 and not curve and (base == "StatScores" or seed >= 50_000):
        C = 2  # `.squeeze()` on the counts of a single label (deviation 12); the functional face returns them as they are
    kind = rnd.choice(["probs", "logits", "labels"]) if not curve else rnd.choice(["probs", "logits"])
    ignore = rnd.choice([None, None, -1, 0, 1])
    if task == "Multiclass":
        kw["num_classes"] = C
        if ignore is not None and ignore >= C:
            ignore = None
    elif task == "Multilabel":
        kw["num_labels"] = C
    kw["ignore_index"] = ignore
    if curve:
        kw["thresholds"] = rnd.choice([None, None, 5, 64, 200, [0.0, 0.3, 0.31, 0.9, 1.0], [0.5]])
        if base in ("AUROC", "AveragePrecision") and task != "Binary":
            kw["average"] = rnd.choice(["macro", "weighted", None] + (["micro"] if task == "Multilabel" else []))
        if base == "AUROC" and task == "Binary" and rnd.random() < 0.3:
            kw["max_fpr"] = rnd.choice([0.2, 0.5, 1.0])
        n = max(n, 33)  # a curve of one sample has nothing to say
    else:
        if task != "Multiclass":
            kw["threshold"] = rnd.choice([0.5, 0.5, 0.3, 0.71])
        if base not in NO_AVERAGE and task != "Binary":
            kw["average"] = rnd.choice(["micro", "macro", "weighted", "none"])
        if base not in NO_SAMPLEWISE and extra and rnd.random() < 0.5:
            kw["multidim_average"] = "samplewise"
            if task == "Multilabel":
                n = max(n, 7)  # a samplewise multilabel batch of ONE sample: DESIGN.md deviation 12 (reference defect)
        if task == "Multiclass" and base not in NO_SAMPLEWISE and kind != "labels" and rnd.random() < 0.3:
            kw["top_k"] = rnd.randint(2, min(C, 6)) if C > 2 else 1
        if base == "CohenKappa":
            kw["weights"] = rnd.choice([None, "linear", "quadratic"])
        if base in ("ConfusionMatrix",):
            kw["normalize"] = rnd.choice([None, "true", "pred", "all"])
        if base == "JaccardIndex" and task != "Binary":
            kw["average"] = rnd.choice(["micro", "macro", "weighted", "none"])

    def batch(gen):
        if task == "Binary":
            pshape, tshape = (n, *extra), (n, *extra)
        elif task == "Multilabel":
            pshape, tshape = (n, C, *extra), (n, C, *extra)
        else:
            pshape, tshape = ((n, *extra) if kind == "labels" else (n, C, *extra)), (n, *extra)
        hi = C if task == "Multiclass" else 2
        target = torch.randint(hi, tshape, generator=gen, device="cuda")
        if ignore is not None:
            target = torch.where(torch.rand(tshape, generator=gen, device="cuda") < 0.15, torch.full_like(target, ignore),
                                 target)
        if kind == "labels":
            preds = torch.randint(hi, pshape, generator=gen, device="cuda")
        else:
            x = torch.randn(pshape, generator=gen, device="cuda") * 2
            if kind == "probs":
                x = x.softmax(1) if task == "Multiclass" else x.sigmoid()
            preds = x.to(dtype)
            if rnd.random() < 0.3:  # ties: quantised scores
                preds = (preds.float() * 8).round().div(8).clamp(0, 1).to(dtype) if kind == "probs" else (
                    (preds.float() * 2).round() / 2).to(dtype)
        # the reference takes targets of any dtype that holds the values (its formats end in `.long()` / `== 1`)
        lo_v, hi_v = min(int(target.min()) if target.numel() else 0, 0), max(hi - 1, ignore if ignore is not None else 0)
        # (bool targets only without `ignore_index`: the reference marks ignored positions with -1, which a bool tensor
        # stores as True — deviation 13)
        fits = {torch.int8: hi_v < 128, torch.uint8: lo_v >= 0 and hi_v < 256,
                torch.bool: lo_v >= 0 and hi_v < 2 and ignore is None, torch.int16: hi_v < 32768}.get(target_dtype, True)
        if fits:
            target = target.to(target_dtype)
        if strided and preds.numel():
            wide_p = torch.zeros((preds.shape[0], 2, *preds.shape[1:]), dtype=preds.dtype, device="cuda")
            wide_t = torch.zeros((target.shape[0], 2, *target.shape[1:]), dtype=target.dtype, device="cuda")
            wide_p[:, 1], wide_t[:, 1] = preds, target
            preds, target = wide_p[:, 1], wide_t[:, 1]
            assert not preds.is_contiguous() or preds.shape[0] <= 1
        return preds, target

    return name, kw, batch


def _states(metric):
    out = {}
    for k in metric._defaults:
        v = getattr(metric, k)
        if isinstance(v, list):
            v = torch.cat([c.reshape(-1) for c in v]) if v else torch.zeros(0)
        out[k] = v
    return out


def _close(a, b, path="value"):
    if isinstance(b, (tuple, list)):
        assert isinstance(a, (tuple, list)) and len(a) == len(b), path
        for i, (x, y) in enumerate(zip(a, b)):
            _close(x, y, f"{path}[{i}]")
        return
    assert a.shape == b.shape, (path, a.shape, b.shape)
    torch.testing.assert_close(a.double(), b.double(), rtol=1e-5, atol=1e-6, equal_nan=True, msg=lambda m: f"{path}: {m}")


@pytest.mark.parametrize("seed", range(N_CASES))
def test_random_configuration(R, M, seed):
    name, kw, batch = _draw(seed)
    try:
        theirs = getattr(R.classification, name)(**kw).cuda()
    except Exception as ref_err:  # noqa: BLE001 - the reference rejects the configuration: so must we
        with pytest.raises(type(ref_err)):
            getattr(M.classification, name)(**kw)
        OUTCOME["rejected_by_both"] += 1
        return
    ours = getattr(M.classification, name)(**kw).cuda()
    gen = torch.Generator(device="cuda").manual_seed(seed)
    for _ in range(2):
        preds, target = batch(gen)
        try:
            theirs.update(preds, target)
        except Exception as ref_err:  # noqa: BLE001
            # torch ops of the reference that do not take the dtype (`one_hot` wants int64, the -1 marker does not fit
            # an unsigned target, half scores in `bucketize` ...) or an empty batch (`reshape(0, -1)`): where the
            # reference has no answer there is nothing to compare; this package may well have one
            OUTCOME["skipped"] += 1
            pytest.skip(f"the reference's update raises here: {type(ref_err).__name__}: {str(ref_err)[:80]}")
        ours.update(preds, target)
    mine, ref = _states(ours), _states(theirs)
    for k, want in ref.items():
        got = mine[k]
        assert got.shape == want.shape, (name, kw, k, got.shape, want.shape)
        assert torch.equal(got.to(want.dtype), want), (name, kw, k)
    try:
        want = theirs.compute()
    except Exception as ref_err:  # noqa: BLE001 - e.g. a label without samples: IndexError in the reference
        OUTCOME["skipped"] += 1
        pytest.skip(f"the reference's compute() raises here: {type(ref_err).__name__}: {str(ref_err)[:80]}")
    _close(ours.compute(), want, f"{name}({kw})")
    OUTCOME["compared"] += 1


def test_zz_most_configurations_were_compared():
    """The generator must not degenerate into configurations both sides reject."""
    total = sum(OUTCOME.values())
    if total < N_CASES // 2:
        pytest.skip("run together with test_random_configuration")
    print("fuzz outcome:", OUTCOME)
    assert OUTCOME["compared"] >= 0.7 * total, OUTCOME


FIXED_POINT = [("PrecisionAtFixedRecall", "min_recall"), ("RecallAtFixedPrecision", "min_precision"),
               ("SpecificityAtSensitivity", "min_sensitivity")]


@pytest.mark.parametrize("seed", range(120 * SCALE))
def test_random_fixed_point_metric(R, M, seed):
    """The *AtFixed* family (row f1): exact and binned curves behind a constrained arg-max."""
    rnd = random.Random(10_000 + seed)
    task = rnd.choice(["Binary", "Multiclass", "Multilabel"])
    base, arg = rnd.choice(FIXED_POINT)
    C = rnd.choice([3, 5, 12]) if task == "Multiclass" else rnd.choice([2, 4, 9])
    kw = {arg: rnd.choice([0.1, 0.5, 0.75, 0.9]), "thresholds": rnd.choice([None, None, 7, 100, [0.1, 0.5, 0.8]]),
          "ignore_index": rnd.choice([None, None, -1])}
    if task == "Multiclass":
        kw["num_classes"] = C
    elif task == "Multilabel":
        kw["num_labels"] = C
    ours, theirs = getattr(M.classification, task + base)(**kw).cuda(), getattr(R.classification, task + base)(**kw).cuda()
    gen = torch.Generator(device="cuda").manual_seed(seed)
    n = rnd.choice([40, 300, 2000])
    dtype = rnd.choice([torch.float32, torch.float32, torch.bfloat16, torch.float64])
    for _ in range(2):
        if task == "Binary":
            p, t = torch.rand(n, generator=gen, device="cuda"), torch.randint(2, (n,), generator=gen, device="cuda")
        elif task == "Multilabel":
            p = torch.rand(n, C, generator=gen, device="cuda")
            t = torch.randint(2, (n, C), generator=gen, device="cuda")
        else:
            p = torch.randn(n, C, generator=gen, device="cuda").softmax(1)
            t = torch.randint(C, (n,), generator=gen, device="cuda")
        if rnd.random() < 0.5:
            p = (p * 16).round() / 16  # ties
        if kw["ignore_index"] is not None:
            t = torch.where(torch.rand(t.shape, generator=gen, device="cuda") < 0.1, torch.full_like(t, -1), t)
        p = p.to(dtype)
        try:
            theirs.update(p, t)
        except Exception as ref_err:  # noqa: BLE001
            pytest.skip(f"the reference's update raises here: {str(ref_err)[:80]}")
        ours.update(p, t)
    try:
        want = theirs.compute()
    except Exception as ref_err:  # noqa: BLE001
        pytest.skip(f"the reference's compute() raises here: {type(ref_err).__name__}: {str(ref_err)[:80]}")
    _close(ours.compute(), want, f"{task + base}({kw})")


@pytest.mark.parametrize("seed", range(80 * SCALE))
def test_random_collection_forward_update_reset(R, M, seed):
    """MetricCollection with compute groups (rows b, f4): forward() batch values, accumulation across forward and
    update, reset, forward again — the sequence in which the in-place forward and the shared group states interact."""
    rnd = random.Random(20_000 + seed)
    task = rnd.choice(["Binary", "Multiclass", "Multilabel"])
    C = rnd.choice([3, 10, 50]) if task == "Multiclass" else rnd.choice([2, 6])
    common = {"ignore_index": rnd.choice([None, None, 0])}
    if task == "Multiclass":
        common["num_classes"] = C
    elif task == "Multilabel":
        common["num_labels"] = C
    names = rnd.sample(["Accuracy", "Precision", "Recall", "F1Score", "Specificity", "ConfusionMatrix", "JaccardIndex",
                        "StatScores", "MatthewsCorrCoef"], k=rnd.randint(2, 5))
    groups = rnd.choice([True, False])

    def build(pkg):
        members = {}
        for nm in names:
            kw = dict(common)
            if task != "Binary" and nm in ("Accuracy", "Precision", "Recall", "F1Score", "Specificity", "StatScores"):
                kw["average"] = rnd_avg[nm]
            members[nm] = getattr(pkg.classification, task + nm)(**kw)
        return pkg.MetricCollection(members, compute_groups=groups).cuda()

    rnd_avg = {nm: rnd.choice(["micro", "macro", "none"]) for nm in names}
    ours, theirs = build(M), build(R)
    gen = torch.Generator(device="cuda").manual_seed(seed)
    n = rnd.choice([16, 200])

    def batch():
        if task == "Binary":
            return torch.rand(n, generator=gen, device="cuda"), torch.randint(2, (n,), generator=gen, device="cuda")
        if task == "Multilabel":
            return (torch.rand(n, C, generator=gen, device="cuda"),
                    torch.randint(2, (n, C), generator=gen, device="cuda"))
        return torch.randn(n, C, generator=gen, device="cuda"), torch.randint(C, (n,), generator=gen, device="cuda")

    def same(a, b, where):
        assert a.keys() == b.keys(), where
        for k in b:
            _close(a[k], b[k], f"{where}:{k}")

    steps = [rnd.choice(["forward", "update"]) for _ in range(4)] + ["compute", "reset", "forward", "update", "compute"]
    for i, step in enumerate(steps):
        if step in ("forward", "update"):
            p, t = batch()
            got, want = getattr(ours, step)(p, t), getattr(theirs, step)(p, t)
            if step == "forward":
                same(got, want, f"step {i} forward")
        elif step == "compute":
            same(ours.compute(), theirs.compute(), f"step {i} compute")
        else:
            ours.reset()
            theirs.reset()


_SNAKE = {"StatScores": "stat_scores", "Accuracy": "accuracy", "Precision": "precision", "Recall": "recall",
          "F1Score": "f1_score", "Specificity": "specificity", "HammingDistance": "hamming_distance",
          "JaccardIndex": "jaccard_index", "ConfusionMatrix": "confusion_matrix", "CohenKappa": "cohen_kappa",
          "MatthewsCorrCoef": "matthews_corrcoef", "PrecisionRecallCurve": "precision_recall_curve", "ROC": "roc",
          "AUROC": "auroc", "AveragePrecision": "average_precision"}


@pytest.mark.parametrize("seed", range(300 * SCALE))
def test_random_functional_call(R, M, seed):
    """The functional face (`torchmetrics.functional.classification.*`) and the task wrappers (`Accuracy(task=...)`):
    the same generator, one batch, `validate_args` on or off."""
    import torchmetrics_forked_b200.functional.classification  # noqa: F401

    name, kw, batch = _draw(50_000 + seed)
    task = next(t for t in ("Binary", "Multiclass", "Multilabel") if name.startswith(t))
    base = name[len(task):]
    fn = f"{task.lower()}_{_SNAKE[base]}"
    if not hasattr(R.functional.classification, fn):
        assert not hasattr(M.functional.classification, fn)
        return
    kw = dict(kw, validate_args=bool(seed % 2))
    preds, target = batch(torch.Generator(device="cuda").manual_seed(seed))
    try:
        want = getattr(R.functional.classification, fn)(preds, target, **kw)
    except Exception as ref_err:  # noqa: BLE001 - see test_random_configuration
        pytest.skip(f"the reference raises here: {type(ref_err).__name__}: {str(ref_err)[:80]}")
    _close(getattr(M.functional.classification, fn)(preds, target, **kw), want, f"{fn}({kw})")
    if seed % 3 == 0 and hasattr(R, base):  # the task wrapper class builds the same metric
        wkw = {k: v for k, v in kw.items() if k != "validate_args"}
        try:
            theirs = getattr(R, base)(task=task.lower(), **wkw).cuda()
        except Exception as ref_err:  # noqa: BLE001 - e.g. a wrapper that insists on num_classes for the binary task
            with pytest.raises(type(ref_err)):
                getattr(M, base)(task=task.lower(), **wkw)
            return
        ours = getattr(M, base)(task=task.lower(), **wkw).cuda()
        assert type(ours).__name__ == type(theirs).__name__ == name
        try:
            want = theirs(preds, target)
        except Exception as ref_err:  # noqa: BLE001 - `validate_args=True` here: e.g. float-typed targets of a curve
            with pytest.raises(type(ref_err)):
                ours(preds, target)
            return
        _close(ours(preds, target), want, f"{base}(task={task.lower()!r}, {wkw})")


def _misaligned(x, off):
    """The same values in storage that starts `off` elements past a 256-byte boundary (kernel dispatch looks at the
    alignment of the pointer and of the rows)."""
    if off == 0:
        return x
    flat = torch.empty(x.numel() + off, dtype=x.dtype, device=x.device)
    flat[off:] = x.reshape(-1)
    return flat[off:].view(x.shape)


@pytest.mark.parametrize("seed", range(160 * SCALE))
def test_random_shapes_through_the_kernel_dispatch(R, M, seed):
    """Mid-size inputs (1e5 - 4e6 scores) of random width, dtype, pointer alignment and target dtype: the shapes at
    which the launchers switch between their kernels (thread per row, lanes per row, warp per row, pipelined rows,
    strided planes, top-k by lane maxima, binary lanes, multilabel tiles, fused softmax rows ...)."""
    rnd = random.Random(70_000 + seed)
    kind = rnd.choice(["confmat", "stats", "topk", "strided", "binary_curve", "multilabel_curve", "multilabel_stats",
                       "multiclass_curve", "exact_auroc"])
    dtype = rnd.choice([torch.float32, torch.bfloat16, torch.float16])
    tdt = rnd.choice([torch.int64, torch.int64, torch.int32, torch.uint8])
    off = rnd.choice([0, 0, 1, 3, 4, 8])
    ignore = rnd.choice([None, None, 0, -1])
    gen = torch.Generator(device="cuda").manual_seed(seed)
    total = rnd.choice([100_000, 700_000, 4_000_000])

    def labels(shape, hi):
        t = torch.randint(hi, shape, generator=gen, device="cuda")
        if ignore is not None:
            t = torch.where(torch.rand(shape, generator=gen, device="cuda") < 0.1, torch.full_like(t, ignore), t)
        ok = tdt != torch.uint8 or (hi <= 256 and ignore != -1)
        return t.to(tdt) if ok else t

    if kind in ("confmat", "stats", "topk", "multiclass_curve", "exact_auroc"):
        C = rnd.choice([2, 3, 7, 8, 10, 16, 31, 32, 33, 64, 100, 128, 136, 200, 256, 264, 500, 512, 1000, 1024, 1100])
        if kind in ("multiclass_curve", "exact_auroc"):
            C = min(C, rnd.choice([10, 100, 1100]))
        if kind == "exact_auroc":
            total = min(total, 700_000)
            C = min(C, 100)
        n = max(total // C, 3)
        p = _misaligned((torch.randn(n, C, generator=gen, device="cuda") * 3).to(dtype), off)
        if rnd.random() < 0.3:
            p = _misaligned(((p.float() * 2).round() / 2).to(dtype), off)  # ties
        t = labels((n,), C)
        ign = ignore if ignore is None or ignore < C else None
        if kind == "confmat":
            name, kw = "MulticlassConfusionMatrix", {"num_classes": C, "ignore_index": ign}
        elif kind == "stats":
            name, kw = "MulticlassStatScores", {"num_classes": C, "ignore_index": ign, "average": rnd.choice(["micro", "none"])}
        elif kind == "topk":
            name, kw = "MulticlassStatScores", {"num_classes": C, "ignore_index": ign, "average": "none",
                                                "top_k": rnd.randint(2, min(C, 7)) if C > 2 else 1}
        elif kind == "multiclass_curve":
            name, kw = "MulticlassPrecisionRecallCurve", {"num_classes": C, "ignore_index": ign,
                                                          "thresholds": rnd.choice([1, 7, 200, 300])}
        else:
            name, kw = "MulticlassAUROC", {"num_classes": C, "ignore_index": ign, "thresholds": None, "average": "macro"}
    elif kind == "strided":
        C, hw = rnd.choice([2, 5, 21, 150]), rnd.choice([(7, 9), (64, 64), (31, 128)])
        n = max(total // (C * hw[0] * hw[1]), 1)
        p = _misaligned(torch.randn(n, C, *hw, generator=gen, device="cuda").to(dtype), off)
        t = labels((n, *hw), C)
        name, kw = "MulticlassJaccardIndex", {"num_classes": C, "ignore_index": ignore if ignore in (None, 0) else None}
    elif kind == "binary_curve":
        p = _misaligned((torch.randn(total, generator=gen, device="cuda") * 3).to(dtype), off)
        if rnd.random() < 0.5:
            p = _misaligned(p.float().sigmoid().to(dtype), off)
        t = labels((total,), 2)
        name, kw = "BinaryPrecisionRecallCurve", {"ignore_index": ignore, "thresholds": rnd.choice([1, 2, 50, 200, 257, 1000])}
    else:
        L = rnd.choice([1, 2, 3, 4, 5, 16, 100, 300])
        n = max(total // L, 3)
        p = _misaligned((torch.randn(n, L, generator=gen, device="cuda") * 3).to(dtype), off)
        t = labels((n, L), 2)
        if kind == "multilabel_curve":
            name, kw = "MultilabelPrecisionRecallCurve", {"num_labels": L, "ignore_index": ignore,
                                                          "thresholds": rnd.choice([1, 50, 200, 600])}
        else:
            name, kw = "MultilabelStatScores", {"num_labels": max(L, 2), "ignore_index": ignore, "average": "none"}
            if L == 1:
                p, t = p.expand(n, 2).contiguous(), t.expand(n, 2).contiguous()
    theirs = getattr(R.classification, name)(**kw, validate_args=False).cuda()
    ours = getattr(M.classification, name)(**kw, validate_args=False).cuda()
    try:
        theirs.update(p, t)
    except Exception as ref_err:  # noqa: BLE001
        pytest.skip(f"the reference's update raises here: {type(ref_err).__name__}: {str(ref_err)[:80]}")
    ours.update(p, t)
    mine, ref = _states(ours), _states(theirs)
    for k, want in ref.items():
        assert mine[k].shape == want.shape and torch.equal(mine[k].to(want.dtype), want), (name, kw, p.shape, p.dtype, k)
    if kind == "exact_auroc":
        _close(ours.compute(), theirs.compute(), f"{name}({kw})")


@pytest.mark.parametrize("seed", range(150 * SCALE))
def test_random_lifecycle(R, M, seed):
    """One metric through a random life: update / forward / compute / reset mixed with clone, pickle, deepcopy,
    `state_dict` round trips (with `persistent(True)`), `.half()` / `.float()` / `.double()` and a move to the CPU and
    back — the same steps on the reference's object, every returned value compared."""
    import copy
    import pickle

    name, kw, batch = _draw(30_000 + seed)
    rnd = random.Random(31_000 + seed)
    try:
        theirs = getattr(R.classification, name)(**kw).cuda()
    except Exception:  # noqa: BLE001
        return
    ours = getattr(M.classification, name)(**kw).cuda()
    gen = torch.Generator(device="cuda").manual_seed(seed)
    steps = [rnd.choice(["update", "update", "forward", "compute", "reset", "clone", "pickle", "deepcopy", "state_dict",
                         "float", "cpu_and_back"]) for _ in range(8)] + ["update", "compute"]
    for i, step in enumerate(steps):
        where = f"{name}({kw}) step {i} {step} of {steps}"
        if step in ("update", "forward"):
            preds, target = batch(gen)
            try:
                want = getattr(theirs, step)(preds, target)
            except Exception as ref_err:  # noqa: BLE001
                pytest.skip(f"the reference raises here: {type(ref_err).__name__}: {str(ref_err)[:80]}")
            got = getattr(ours, step)(preds, target)
            if step == "forward":
                _close(got, want, where)
        elif step == "compute":
            try:
                want = theirs.compute()
            except Exception as ref_err:  # noqa: BLE001
                pytest.skip(f"the reference's compute() raises here: {type(ref_err).__name__}: {str(ref_err)[:80]}")
            _close(ours.compute(), want, where)
        elif step == "reset":
            ours.reset()
            theirs.reset()
        elif step == "clone":
            ours, theirs = ours.clone(), theirs.clone()
        elif step == "pickle":
            ours, theirs = pickle.loads(pickle.dumps(ours)), pickle.loads(pickle.dumps(theirs))
        elif step == "deepcopy":
            ours, theirs = copy.deepcopy(ours), copy.deepcopy(theirs)
        elif step == "state_dict":
            ours.persistent(True)
            theirs.persistent(True)
            a, b = ours.state_dict(), theirs.state_dict()
            assert a.keys() == b.keys(), where
            fresh_o, fresh_t = getattr(M.classification, name)(**kw).cuda(), getattr(R.classification, name)(**kw).cuda()
            fresh_o.persistent(True)
            fresh_t.persistent(True)
            try:
                fresh_t.load_state_dict(b)
            except Exception as ref_err:  # noqa: BLE001 - list states of the reference do not always load back
                with pytest.raises(type(ref_err)):
                    fresh_o.load_state_dict(a)
                continue
            fresh_o.load_state_dict(a)
            fresh_o._update_count, fresh_t._update_count = ours._update_count, theirs._update_count
            ours, theirs = fresh_o, fresh_t
        elif step == "float":
            how = rnd.choice(["half", "float", "double"])
            ours, theirs = getattr(ours, how)(), getattr(theirs, how)()
        else:
            ours, theirs = ours.cpu().cuda(), theirs.cpu().cuda()


_BAD_ARGS = [("average", "mean"), ("average", 3), ("threshold", 1.5), ("threshold", "high"), ("thresholds", -4),
             ("thresholds", [0.2, 1.3]), ("thresholds", 0.5), ("num_classes", 1), ("num_classes", 2.5), ("num_labels", 0),
             ("top_k", 0), ("top_k", 10_000), ("ignore_index", "last"), ("ignore_index", 1.5), ("multidim_average", "rows"),
             ("normalize", "both"), ("weights", "cubic"), ("max_fpr", 1.5), ("max_fpr", 0)]


@pytest.mark.parametrize("seed", range(200 * SCALE))
def test_random_invalid_use_is_rejected_like_the_reference(R, M, seed):
    """Arguments and inputs the reference rejects (`validate_args=True`): the same exception type here, from the
    constructor, `update` or `compute` — bad argument values, shapes that do not match, integer scores for a curve,
    float targets, labels outside the task's range.  (Where the reference lets a bad input pass, nothing is asked.)"""
    name, kw, batch = _draw(40_000 + seed)
    rnd = random.Random(41_000 + seed)
    mode = rnd.choice(["argument", "argument", "shape", "rank", "dtype", "label"])
    kw = dict(kw)
    if mode == "argument":
        key, value = rnd.choice(_BAD_ARGS)
        kw[key] = value

    def attempt(pkg, device="cuda"):
        metric = getattr(pkg.classification, name)(**kw).to(device)
        preds, target = (x.to(device) for x in batch(torch.Generator(device="cuda").manual_seed(seed)))
        if mode == "shape" and preds.shape[0] > 1:
            preds = preds[:-1]
        elif mode == "rank":
            preds = preds.unsqueeze(-1).expand(*preds.shape, 2)
        elif mode == "dtype":
            if rnd_mode == 0:
                preds = (preds * 3).long() if preds.is_floating_point() else preds.double() / 3 + 0.1
            else:
                target = target.float() + 0.5
        elif mode == "label" and target.numel():
            target = target.long().clone()
            target.view(-1)[0] = 7_000 if rnd_mode == 0 else -9
        metric.update(preds, target)
        return metric.compute()

    rnd_mode = rnd.randint(0, 1)
    try:
        # a bad label (or an integer "score") that gets past the reference's checks ends in an index assert of a torch
        # kernel; on the GPU that is a device-side assert, which poisons the CUDA context of the whole process: the
        # reference meets these inputs on the CPU
        attempt(R, "cpu")
    except Exception as ref_err:  # noqa: BLE001
        kind = type(ref_err)
        diagnosis = kind is ValueError or (kind is RuntimeError and any(
            phrase in str(ref_err) for phrase in ("Detected", "expected to have the same shape", "Expected", "must match")))
        if not diagnosis:  # an IndexError, a torch kernel's complaint about its operands, ...: the reference crashed
            pytest.skip(f"the reference fails without a diagnosis here: {kind.__name__}: {str(ref_err)[:80]}")
        with pytest.raises(kind):
            attempt(M)
        return
    pytest.skip("the reference accepts this")
