"""GPU parity: K1 (argmax + histogram) and K2 / K1b (threshold binning) against the oracle.

Every test calls the CUDA path — either the raw C ABI through ctypes or the public functional / Metric API,
which are thin wrappers over the same entry points — and compares with `oracle/hotpath.py` on the same
inputs.  Integer states must match bit for bit; derived floats within 1e-6 relative.
"""
import ctypes

import numpy as np
import pytest
import torch
from conftest import golden_names, ignore_of, load_golden

import hotpath as O

pytestmark = pytest.mark.gpu

RTOL, ATOL = 1e-6, 1e-7  # tolerance for derived float metrics (north_star: 1e-6 relative)


def _cuda(x, dtype=None):
    t = torch.from_numpy(np.ascontiguousarray(x))
    if dtype is not None:
        t = t.to(dtype)
    return t.cuda()


def cuda_sigmoid(x: np.ndarray) -> np.ndarray:
    """What the reference would execute on this device: torch's CUDA sigmoid kernel."""
    return torch.from_numpy(np.ascontiguousarray(x, dtype=np.float32)).cuda().sigmoid().cpu().numpy()


@pytest.fixture(scope="module")
def F():
    import torchmetrics_forked_b200.functional as fn

    return fn


# ------------------------------------------------------------------------------------------ raw C ABI
def test_cabi_mc_argmax_hist_raw_pointers():
    from torchmetrics_forked_b200 import _native as nat

    lib = nat.lib()
    g = torch.Generator().manual_seed(1)
    preds = torch.randn(1000, 12, generator=g)
    target = torch.randint(12, (1000,), generator=g)
    dp, dt = preds.cuda(), target.cuda()
    out = torch.zeros(12, 12, dtype=torch.int64, device="cuda")
    err = torch.zeros(1, dtype=torch.int32, device="cuda")
    rc = lib.tmb_mc_argmax_hist(
        ctypes.c_void_p(dp.data_ptr()), 0, ctypes.c_void_p(dt.data_ptr()), 0, 1000, 12, 1, 0, 0, 0,
        ctypes.c_void_p(out.data_ptr()), ctypes.c_void_p(err.data_ptr()),
        ctypes.c_void_p(torch.cuda.current_stream().cuda_stream),
    )
    assert rc == 0
    torch.cuda.synchronize()
    assert int(err.item()) == 0
    assert np.array_equal(out.cpu().numpy(), O.multiclass_confusion_matrix(preds.numpy(), target.numpy(), 12))
    # accumulation is in place: a second call doubles the state
    lib.tmb_mc_argmax_hist(
        ctypes.c_void_p(dp.data_ptr()), 0, ctypes.c_void_p(dt.data_ptr()), 0, 1000, 12, 1, 0, 0, 0,
        ctypes.c_void_p(out.data_ptr()), ctypes.c_void_p(err.data_ptr()),
        ctypes.c_void_p(torch.cuda.current_stream().cuda_stream),
    )
    assert np.array_equal(out.cpu().numpy(), 2 * O.multiclass_confusion_matrix(preds.numpy(), target.numpy(), 12))
    # argument errors come back as negative codes, nothing is launched
    assert lib.tmb_mc_argmax_hist(None, 7, None, 0, 10, 12, 1, 0, 0, 0, None, None, None) == -1
    assert lib.tmb_mc_argmax_hist(None, 0, None, 0, 10, 12, 1, 0, 0, 9, ctypes.c_void_p(out.data_ptr()),
                                  ctypes.c_void_p(err.data_ptr()), None) == -4


def test_cabi_thresh_bin_raw_pointers():
    from torchmetrics_forked_b200 import _native as nat

    lib = nat.lib()
    g = torch.Generator().manual_seed(2)
    n, T = 100_003, 200
    preds = torch.rand(n, generator=g)
    target = torch.randint(2, (n,), generator=g)
    thr = torch.linspace(0, 1, T)
    dp, dt, dthr = preds.cuda(), target.cuda(), thr.cuda()
    elems = lib.tmb_thresh_bin_scratch_elems(1, T)
    scratch = torch.zeros(elems, dtype=torch.int64, device="cuda")
    state = torch.zeros(T, 2, 2, dtype=torch.int64, device="cuda")
    err = torch.zeros(1, dtype=torch.int32, device="cuda")
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = lib.tmb_thresh_bin_update(
        ctypes.c_void_p(dp.data_ptr()), 0, 0, ctypes.c_void_p(dt.data_ptr()), 0, n, 1, 1,
        ctypes.c_void_p(dthr.data_ptr()), T, 0, 0, 0, 0, ctypes.c_void_p(scratch.data_ptr()), elems,
        ctypes.c_void_p(err.data_ptr()), st,
    )
    assert rc == 0
    rc = lib.tmb_thresh_bin_finalize(
        ctypes.c_void_p(scratch.data_ptr()), elems, 1, T, None, 0, ctypes.c_void_p(state.data_ptr()), None, None,
        None, st,
    )
    assert rc == 0
    torch.cuda.synchronize()
    assert np.array_equal(state.cpu().numpy(), O.binary_curve_binned(preds.numpy(), target.numpy(), thr.numpy()))
    assert int(scratch.abs().sum().item()) == 0, "finalize must leave the scratch zeroed"
    assert lib.tmb_thresh_bin_update(None, 0, 0, None, 0, 5, 1, 1, ctypes.c_void_p(dthr.data_ptr()), T, 0, 0, 0, 0,
                                     ctypes.c_void_p(scratch.data_ptr()), 3, ctypes.c_void_p(err.data_ptr()),
                                     st) == -5


# ------------------------------------------------------------------------------------------ K1 on the golden fixtures
@pytest.mark.parametrize("name", golden_names("mc_"))
def test_multiclass_golden(name, F):
    c = load_golden(name)
    C, ign = int(c["num_classes"]), ignore_of(c)
    is_bf16 = "bf16" in name
    preds = _cuda(c["preds"], torch.bfloat16 if is_bf16 else None)
    target = _cuda(c["target"])
    got = F.multiclass_confusion_matrix(preds, target, C, ignore_index=ign)
    assert np.array_equal(got.cpu().numpy(), c["confmat"])
    for avg in ("micro", "macro", "none"):
        assert np.array_equal(
            F.multiclass_stat_scores(preds, target, C, average=avg, ignore_index=ign).cpu().numpy(), c[f"stat_{avg}"]
        ) or avg == "macro"
        np.testing.assert_allclose(
            F.multiclass_stat_scores(preds, target, C, average=avg, ignore_index=ign).cpu().numpy(),
            c[f"stat_{avg}"], rtol=RTOL, atol=ATOL,
        )
        np.testing.assert_allclose(
            F.multiclass_accuracy(preds, target, C, average=avg, ignore_index=ign).cpu().numpy(), c[f"acc_{avg}"],
            rtol=RTOL, atol=ATOL,
        )
        np.testing.assert_allclose(
            F.multiclass_f1_score(preds, target, C, average=avg, ignore_index=ign).cpu().numpy(), c[f"f1_{avg}"],
            rtol=RTOL, atol=ATOL,
        )
    for avg in ("micro", "macro", "weighted", "none"):
        np.testing.assert_allclose(
            F.multiclass_jaccard_index(preds, target, C, average=avg, ignore_index=ign).cpu().numpy(),
            c[f"jaccard_{avg}"], rtol=RTOL, atol=ATOL,
        )


# ------------------------------------------------------------------------------------------ K1 layouts, dtypes, alignment
@pytest.mark.parametrize(
    ("n", "C", "dtype"),
    [
        (100_000, 5, torch.float32),  # config 1 shape: thread-per-row kernel
        (100_000, 10, torch.float32),
        (20_000, 63, torch.float16),
        (8192, 1000, torch.bfloat16),  # config 2 row shape: warp-per-row, 16-byte vectors
        (4096, 1000, torch.float32),
        (3000, 150, torch.float32),  # 600-byte rows: 8-byte vectors
        (3000, 37, torch.bfloat16),  # 74-byte rows: 2-byte granularity
        (1500, 333, torch.float16),
        (257, 4099, torch.bfloat16),  # rows longer than one register tile
        # 17 .. 64 vectors per row: several rows per warp (k1_rows_sub), 8 or 16 lanes per row, ragged row counts
        (3001, 136, torch.bfloat16),
        (3001, 100, torch.float32),
        (2999, 256, torch.bfloat16),
        (1777, 200, torch.float32),
        (4001, 72, torch.float32),
        (2003, 132, torch.float16),
        (999, 512, torch.bfloat16),
        # rows of a multiple of four words in the thread-per-row kernel: padded shared-memory tile
        (5003, 64, torch.float32),  # 64 x 64 confusion matrix = the largest privatised histogram, next to the tile
        (5003, 128, torch.bfloat16),
        (7001, 32, torch.float32),
        (7001, 8, torch.bfloat16),
        (7001, 4, torch.float32),
        (4099, 16, torch.float16),
    ],
)
@pytest.mark.parametrize("tdtype", [torch.int64, torch.uint8, torch.int32])
def test_multiclass_rows_random(n, C, dtype, tdtype, F):
    if tdtype == torch.uint8 and C > 255:
        pytest.skip("labels do not fit uint8")
    g = torch.Generator().manual_seed(n + C)
    preds = torch.randn(n, C, generator=g).to(dtype)
    target = torch.randint(C, (n,), generator=g)
    # diagonal-heavy like real classifiers, and a few ties / NaNs
    hot = torch.rand(n, generator=g) < 0.7
    preds[hot, target[hot]] += 4
    preds[5, :] = 1.0
    preds[6, C // 2] = float("nan")
    want = O.multiclass_confusion_matrix(preds.float().numpy(), target.numpy(), C)
    got = F.multiclass_confusion_matrix(preds.cuda(), target.to(tdtype).cuda(), C, validate_args=False)
    assert np.array_equal(got.cpu().numpy(), want)
    tp, fp, tn, fn = O.multiclass_stat_scores(preds.float().numpy(), target.numpy(), C, None)
    got = F.multiclass_stat_scores(preds.cuda(), target.to(tdtype).cuda(), C, average=None, validate_args=False)
    assert np.array_equal(got.cpu().numpy(), np.stack([tp, fp, tn, fn, tp + fn], -1))
    tp, fp, tn, fn = O.multiclass_stat_scores(preds.float().numpy(), target.numpy(), C, "micro")
    got = F.multiclass_stat_scores(preds.cuda(), target.to(tdtype).cuda(), C, average="micro", validate_args=False)
    assert got.cpu().tolist() == [tp, fp, tn, fn, tp + fn]


@pytest.mark.parametrize("offset", [1, 2, 3])
def test_multiclass_rows_misaligned_base(offset, F):
    """A sliced view shifts the base pointer off 16 bytes: the launcher must pick a narrower vector."""
    g = torch.Generator().manual_seed(offset)
    C = 100
    big = torch.randn(2000 * C + 8, generator=g).cuda()
    preds = big[offset : offset + 2000 * C].view(2000, C)
    assert preds.data_ptr() % 16 != 0 and preds.is_contiguous()
    target = torch.randint(C, (2000,), generator=g)
    want = O.multiclass_confusion_matrix(preds.cpu().numpy(), target.numpy(), C)
    assert np.array_equal(F.multiclass_confusion_matrix(preds, target.cuda(), C).cpu().numpy(), want)


@pytest.mark.parametrize(
    ("shape", "dtype"),
    [
        ((4, 150, 64, 64), torch.float32),  # config 5 layout, scaled
        ((3, 150, 33, 31), torch.float32),  # odd plane size -> 4-byte vectors
        ((2, 19, 128, 128), torch.bfloat16),
        ((8, 5, 3), torch.float32),  # the reference tests' EXTRA_DIM shape
        ((2, 21, 7, 5, 3), torch.float16),
    ],
)
def test_multiclass_strided_random(shape, dtype, F):
    g = torch.Generator().manual_seed(sum(shape))
    C = shape[1]
    preds = torch.randn(*shape, generator=g).to(dtype)
    target = torch.randint(C, (shape[0], *shape[2:]), generator=g)
    target.view(-1)[::17] = C - 1
    want = O.multiclass_confusion_matrix(preds.float().numpy(), target.numpy(), C, ignore_index=C - 1)
    got = F.multiclass_confusion_matrix(preds.cuda(), target.cuda(), C, ignore_index=C - 1)
    assert np.array_equal(got.cpu().numpy(), want)
    # non-contiguous input (channels-last style) is accepted and gives the same answer
    perm = preds.cuda().movedim(1, -1).contiguous().movedim(-1, 1)
    assert not perm.is_contiguous()
    assert np.array_equal(F.multiclass_confusion_matrix(perm, target.cuda(), C, ignore_index=C - 1).cpu().numpy(), want)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
def test_multiclass_strided_ties_nan_inf(dtype, F):
    """torch.argmax's rule on the class-strided layout: first maximum wins, a NaN beats everything (first NaN
    wins), all -inf gives class 0 — including the packed two-pixels-per-register path of the 16-bit dtypes."""
    g = torch.Generator().manual_seed(17)
    C = 40
    preds = torch.randn(3, C, 32, 64, generator=g).to(dtype)  # 16-bit rounding alone already makes ties
    preds[0, :, 0, :8] = 0.5  # every class equal -> 0
    preds[0, 7, 1, :16] = float("nan")  # single NaN -> 7
    preds[0, 3, 2, :16] = float("nan")
    preds[0, 30, 2, :16] = float("nan")  # two NaNs -> the first, 3
    preds[1, :, 3, :] = float("-inf")  # all -inf -> 0
    preds[1, 5, 4, :] = float("inf")
    preds[1, 9, 4, :] = float("inf")  # two +inf -> 5
    preds[2, 11, 5, :] = float("inf")
    preds[2, 20, 5, :] = float("nan")  # NaN beats +inf -> 20
    preds[2, 0, 6, :] = -0.0
    preds[2, 1:, 6, :] = -1.0
    preds[2, 13, 6, :] = 0.0  # -0.0 == +0.0 -> the first, 0
    target = torch.randint(C, (3, 32, 64), generator=g)
    want = O.multiclass_confusion_matrix(preds.float().numpy(), target.numpy(), C)
    assert np.array_equal(F.multiclass_confusion_matrix(preds.cuda(), target.cuda(), C).cpu().numpy(), want)
    assert torch.equal(preds.cuda().argmax(1).cpu(), torch.from_numpy(O.argmax_first(preds.float().numpy(), 1)))


def test_multiclass_empty_and_errors(F):
    empty = F.multiclass_confusion_matrix(torch.zeros(0, 4).cuda(), torch.zeros(0, dtype=torch.long).cuda(), 4)
    assert empty.shape == (4, 4) and int(empty.sum()) == 0
    preds = torch.randn(10, 4).cuda()
    bad = torch.tensor([0, 1, 2, 3, 4, 0, 1, 2, 3, 7]).cuda()
    with pytest.raises(RuntimeError):
        F.multiclass_confusion_matrix(preds, bad, 4)
    with pytest.raises(RuntimeError):
        F.multiclass_confusion_matrix(torch.tensor([0, 9]).cuda(), torch.tensor([0, 1]).cuda(), 4)
    with pytest.raises(RuntimeError, match="no CPU"):
        F.multiclass_confusion_matrix(torch.randn(10, 4), torch.zeros(10, dtype=torch.long), 4)
    with pytest.raises(ValueError):
        F.multiclass_confusion_matrix(preds, torch.zeros(10, 2, dtype=torch.long).cuda(), 4)


# ------------------------------------------------------------------------------------------ K1b
@pytest.mark.parametrize("name", golden_names("bin_"))
def test_binary_counts_golden(name, F):
    c = load_golden(name)
    ign, thr = ignore_of(c), float(c["threshold"])
    preds, target = _cuda(c["preds"]), _cuda(c["target"])
    want_stat = O.binary_stat_scores(c["preds"], c["target"], thr, ign, sigmoid_fn=cuda_sigmoid)
    got = F.binary_stat_scores(preds, target, thr, ignore_index=ign).cpu().tolist()
    assert got[:4] == list(want_stat)
    want_cm = O.binary_confusion_matrix(c["preds"], c["target"], thr, ign, sigmoid_fn=cuda_sigmoid)
    assert np.array_equal(F.binary_confusion_matrix(preds, target, thr, ignore_index=ign).cpu().numpy(), want_cm)
    if "logits" not in name and "multidim" not in name:
        # no sigmoid involved: the CPU reference's golden output is bit exact too
        assert got == c["stat"].tolist()
        assert np.array_equal(want_cm, c["confmat"])


@pytest.mark.parametrize("name", golden_names("ml_"))
def test_multilabel_counts_golden(name, F):
    c = load_golden(name)
    L, ign, thr = int(c["num_labels"]), ignore_of(c), float(c["threshold"])
    preds, target = _cuda(c["preds"]), _cuda(c["target"])
    tp, fp, tn, fn = O.multilabel_stat_scores(c["preds"], c["target"], L, thr, ign, sigmoid_fn=cuda_sigmoid)
    got = F.multilabel_stat_scores(preds, target, L, thr, average=None, ignore_index=ign).cpu().numpy()
    assert np.array_equal(got, np.stack([tp, fp, tn, fn, tp + fn], -1))
    want_cm = O.multilabel_confusion_matrix(c["preds"], c["target"], L, thr, ign, sigmoid_fn=cuda_sigmoid)
    assert np.array_equal(F.multilabel_confusion_matrix(preds, target, L, thr, ignore_index=ign).cpu().numpy(), want_cm)
    if "logits" not in name:
        assert np.array_equal(got, c["stat_none"]) and np.array_equal(want_cm, c["confmat"])


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16])
@pytest.mark.parametrize("logits", [False, True])
def test_binary_counts_random(dtype, logits, F):
    g = torch.Generator().manual_seed(11)
    n = 300_007
    preds = (torch.randn(n, generator=g) * 3 if logits else torch.rand(n, generator=g)).to(dtype)
    target = torch.randint(2, (n,), generator=g)
    sig = lambda x: torch.from_numpy(x).to(dtype).cuda().sigmoid().float().cpu().numpy()  # noqa: E731
    thr = 0.3
    thr_eff = float(torch.tensor(thr, dtype=dtype))  # the reference compares in the dtype of preds
    want = O.binary_stat_scores(preds.float().numpy(), target.numpy(), thr_eff, None, sigmoid_fn=sig)
    got = F.binary_stat_scores(preds.cuda(), target.cuda(), thr).cpu().tolist()
    assert got[:4] == list(want)


def test_multilabel_counts_random_layouts(F):
    g = torch.Generator().manual_seed(12)
    for shape in [(5000, 100), (64, 7, 33), (12, 3, 16, 16)]:
        L = shape[1]
        preds = torch.randn(*shape, generator=g)
        target = torch.randint(2, shape, generator=g)
        target.view(-1)[::13] = -1
        tp, fp, tn, fn = O.multilabel_stat_scores(preds.numpy(), target.numpy(), L, 0.5, -1, sigmoid_fn=cuda_sigmoid)
        got = F.multilabel_stat_scores(preds.cuda(), target.cuda(), L, average=None, ignore_index=-1).cpu().numpy()
        assert np.array_equal(got, np.stack([tp, fp, tn, fn, tp + fn], -1)), shape


def test_binary_target_validation(F):
    with pytest.raises(RuntimeError):
        F.binary_stat_scores(torch.rand(100).cuda(), torch.full((100,), 3).cuda())


# ------------------------------------------------------------------------------------------ K2 curves
@pytest.mark.parametrize("name", golden_names("bcurve_"))
def test_binary_curve_golden(name, F):
    from torchmetrics_forked_b200.classification import BinaryAUROC, BinaryAveragePrecision, BinaryPrecisionRecallCurve

    c = load_golden(name)
    ign = ignore_of(c)
    thresholds = torch.from_numpy(c["thresholds"])
    want = O.binary_curve_binned(c["preds"], c["target"], c["thresholds"], ign, sigmoid_fn=cuda_sigmoid)
    m = BinaryPrecisionRecallCurve(thresholds=thresholds, ignore_index=ign).cuda()
    m.update(_cuda(c["preds"]), _cuda(c["target"]))
    assert np.array_equal(m.confmat.cpu().numpy(), want)
    if "logits" not in name and "nan" not in name:
        assert np.array_equal(want, c["state"])  # no sigmoid: identical to the CPU reference run
    else:
        assert np.abs(want - c["state"]).sum() <= 8  # CPU-vs-CUDA sigmoid may move a boundary sample (H5)
    for cls, key in ((BinaryAUROC, "auroc"), (BinaryAveragePrecision, "ap")):
        mm = cls(thresholds=thresholds, ignore_index=ign).cuda()
        mm.update(_cuda(c["preds"]), _cuda(c["target"]))
        np.testing.assert_allclose(mm.compute().cpu().numpy(), c[key], rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("name", golden_names("switch16_"))
def test_16bit_scores_follow_the_reference_loop_or_vectorised_compare(name):
    """bf16 / fp16 scores near the thresholds: the reference's per-threshold loop compares against thresholds rounded
    to the score dtype, its vectorised form against float32 ones, and it picks by the number of samples LEFT after the
    `ignore_index` filter (precision_recall_curve.py:203, 473).  Goldens from the unmodified reference on both sides
    of the switch, with and without ignored samples (oracle/gen_golden_r2.py)."""
    from torchmetrics_forked_b200.classification import BinaryPrecisionRecallCurve, MulticlassPrecisionRecallCurve

    c = load_golden(name)
    dt = getattr(torch, str(c["dtype"]))
    ign = ignore_of(c)
    preds = torch.from_numpy(c["preds"]).to(dt).cuda()  # exact: the stored float32 values are dtype values
    target = torch.from_numpy(c["target"]).cuda()
    if "binary" in name:
        m = BinaryPrecisionRecallCurve(thresholds=int(c["thresholds"]), ignore_index=ign).cuda()
    else:
        m = MulticlassPrecisionRecallCurve(int(c["num_classes"]), thresholds=int(c["thresholds"]), ignore_index=ign).cuda()
    m.update(preds, target)
    assert np.array_equal(m.confmat.cpu().numpy(), c["state"]), (name, int(c["kept"]))


@pytest.mark.parametrize("name", golden_names("mlcurve_"))
def test_multilabel_curve_golden(name, F):
    from torchmetrics_forked_b200.classification import MultilabelAUROC, MultilabelPrecisionRecallCurve

    c = load_golden(name)
    L, ign = int(c["num_labels"]), ignore_of(c)
    thresholds = torch.from_numpy(c["thresholds"])
    want = O.multilabel_curve_binned(c["preds"], c["target"], L, c["thresholds"], ign, sigmoid_fn=cuda_sigmoid)
    m = MultilabelPrecisionRecallCurve(L, thresholds=thresholds, ignore_index=ign).cuda()
    m.update(_cuda(c["preds"]), _cuda(c["target"]))
    assert np.array_equal(m.confmat.cpu().numpy(), want)
    if "logits" not in name:
        assert np.array_equal(want, c["state"])
    for avg in ("micro", "macro", "none"):
        mm = MultilabelAUROC(L, average=avg, thresholds=thresholds, ignore_index=ign).cuda()
        mm.update(_cuda(c["preds"]), _cuda(c["target"]))
        np.testing.assert_allclose(mm.compute().cpu().numpy(), c[f"auroc_{avg}"], rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("tdtype", [torch.int64, torch.uint8, torch.int32, torch.int8, torch.bool, torch.int16])
@pytest.mark.parametrize("pdtype", [torch.float32, torch.bfloat16])
def test_binary_curve_random_dtypes(tdtype, pdtype, F):
    from torchmetrics_forked_b200.classification import BinaryPrecisionRecallCurve

    g = torch.Generator().manual_seed(21)
    n, T = 1_000_003, 200
    preds = (torch.randn(n, generator=g) * 3).to(pdtype)
    target = torch.randint(2, (n,), generator=g)
    m = BinaryPrecisionRecallCurve(thresholds=T, validate_args=False).cuda()
    m.update(preds.cuda(), target.to(tdtype).cuda())
    sig = lambda x: torch.from_numpy(x).to(pdtype).cuda().sigmoid().float().cpu().numpy()  # noqa: E731
    thr = torch.linspace(0, 1, T)
    if pdtype != torch.float32:
        thr = thr.to(pdtype).float()  # above 50 000 samples the reference compares in the dtype of preds
    want = O.binary_curve_binned(preds.float().numpy(), target.numpy(), thr.numpy(), None, sigmoid_fn=sig)
    assert np.array_equal(m.confmat.cpu().numpy(), want)


def test_binary_curve_probs_and_accumulation(F):
    """Inputs already in [0,1] take the raw pass only; two updates accumulate; forward() returns batch values."""
    from torchmetrics_forked_b200.classification import BinaryPrecisionRecallCurve

    g = torch.Generator().manual_seed(22)
    m = BinaryPrecisionRecallCurve(thresholds=64).cuda()
    total = np.zeros((64, 2, 2), np.int64)
    thr = torch.linspace(0, 1, 64).numpy()  # NOT np.linspace: the two differ in the last bit
    for i in range(3):
        preds = torch.rand(70_001 + i, generator=g)
        target = torch.randint(2, (70_001 + i,), generator=g)
        m.update(preds.cuda(), target.cuda())
        total += O.binary_curve_binned(preds.numpy(), target.numpy(), thr)
    assert np.array_equal(m.confmat.cpu().numpy(), total)


def test_binary_curve_h1_single_outlier(F):
    """One logit hidden in a batch of probabilities flips the whole batch to sigmoid (H1)."""
    from torchmetrics_forked_b200.classification import BinaryPrecisionRecallCurve

    g = torch.Generator().manual_seed(23)
    n = 400_000
    preds = torch.rand(n, generator=g)
    preds[n - 7] = 1.5
    target = torch.randint(2, (n,), generator=g)
    m = BinaryPrecisionRecallCurve(thresholds=50).cuda()
    m.update(preds.cuda(), target.cuda())
    want = O.binary_curve_binned(preds.numpy(), target.numpy(), torch.linspace(0, 1, 50).numpy(), sigmoid_fn=cuda_sigmoid)
    assert np.array_equal(m.confmat.cpu().numpy(), want)
    # ... unless that logit sits at an ignored position (H2, binary curves filter first)
    target[n - 7] = -1
    m = BinaryPrecisionRecallCurve(thresholds=50, ignore_index=-1).cuda()
    m.update(preds.cuda(), target.cuda())
    want = O.binary_curve_binned(preds.numpy(), target.numpy(), torch.linspace(0, 1, 50).numpy(), -1, sigmoid_fn=cuda_sigmoid)
    assert np.array_equal(m.confmat.cpu().numpy(), want)


@pytest.mark.parametrize("T", [2, 3, 1000, 5000])
def test_binary_curve_threshold_counts(T, F):
    """Few thresholds, and more thresholds than the lane-private histograms can hold (generic path)."""
    from torchmetrics_forked_b200.classification import BinaryPrecisionRecallCurve

    g = torch.Generator().manual_seed(T)
    preds = torch.rand(200_000, generator=g)
    target = torch.randint(2, (200_000,), generator=g)
    thr = torch.sort(torch.rand(T, generator=g)).values
    m = BinaryPrecisionRecallCurve(thresholds=thr).cuda()
    m.update(preds.cuda(), target.cuda())
    assert np.array_equal(m.confmat.cpu().numpy(), O.binary_curve_binned(preds.numpy(), target.numpy(), thr.numpy()))


def test_multilabel_curve_random(F):
    from torchmetrics_forked_b200.classification import MultilabelPrecisionRecallCurve

    g = torch.Generator().manual_seed(31)
    for shape, T in [((20_000, 100), 200), ((300, 5, 40), 33)]:
        L = shape[1]
        preds = torch.randn(*shape, generator=g)
        target = torch.randint(2, shape, generator=g)
        m = MultilabelPrecisionRecallCurve(L, thresholds=T).cuda()
        m.update(preds.cuda(), target.cuda())
        want = O.multilabel_curve_binned(preds.numpy(), target.numpy(), L, torch.linspace(0, 1, T).numpy(),
                                         sigmoid_fn=cuda_sigmoid)
        assert np.array_equal(m.confmat.cpu().numpy(), want), shape


def test_sigmoid_bits_match_torch_cuda():
    """SURVEY Appendix B probe 1: the kernels' sigmoid must be torch's CUDA sigmoid bit for bit — checked
    through the binned state with one threshold per distinct sigmoid value of a dense logit sweep."""
    from torchmetrics_forked_b200.classification import BinaryPrecisionRecallCurve

    x = torch.linspace(-12, 12, 2_000_001)
    s = x.cuda().sigmoid()
    thr = torch.unique(s)[::997].cpu()  # thresholds that ARE sigmoid outputs: any 1-ulp drift moves a count
    target = torch.ones_like(x, dtype=torch.long)
    m = BinaryPrecisionRecallCurve(thresholds=thr).cuda()
    m.update(x.cuda(), target.cuda())
    want = (s.unsqueeze(1) >= thr.cuda().unsqueeze(0)).sum(0).cpu().numpy()
    assert np.array_equal(m.confmat[:, 1, 1].cpu().numpy(), want)


@pytest.mark.parametrize("T", [200, 2, 1000])
def test_binary_curve_saturated_and_special_logits(T):
    """The filtered sigmoid must stay exact where the approximation is weakest: saturated logits (q within
    1e-9 of the first / last threshold), huge magnitudes, infinities, NaN, denormal-producing inputs."""
    from torchmetrics_forked_b200.classification import BinaryPrecisionRecallCurve

    g = torch.Generator().manual_seed(99 + T)
    n = 2_000_003
    x = torch.randn(n, generator=g) * 15
    x[::1001] = float("inf")
    x[7::1003] = float("-inf")
    x[11::1009] = float("nan")
    x[13::1013] = 88.5
    x[17::1019] = -88.5
    x[19::1021] = -103.9
    x[23::1031] = 0.0
    target = torch.randint(2, (n,), generator=g)
    m = BinaryPrecisionRecallCurve(thresholds=T, validate_args=False).cuda()
    m.update(x.cuda(), target.cuda())
    want = O.binary_curve_binned(x.numpy(), target.numpy(), torch.linspace(0, 1, T).numpy(), sigmoid_fn=cuda_sigmoid)
    assert np.array_equal(m.confmat.cpu().numpy(), want)


def test_binary_curve_uneven_thresholds_with_logits():
    """Thresholds off the even grid take the compare-against-the-buffer path of the filtered sigmoid."""
    from torchmetrics_forked_b200.classification import BinaryPrecisionRecallCurve

    g = torch.Generator().manual_seed(123)
    n = 1_000_001
    x = torch.randn(n, generator=g) * 4
    target = torch.randint(2, (n,), generator=g)
    thr = torch.sort(torch.rand(150, generator=g) ** 3).values
    m = BinaryPrecisionRecallCurve(thresholds=thr, validate_args=False).cuda()
    m.update(x.cuda(), target.cuda())
    want = O.binary_curve_binned(x.numpy(), target.numpy(), thr.numpy(), sigmoid_fn=cuda_sigmoid)
    assert np.array_equal(m.confmat.cpu().numpy(), want)


# ------------------------------------------------------------------------------------------ Metric classes end to end
def test_collection_config1_matches_oracle():
    """BASELINE config 1: MetricCollection([MulticlassAccuracy, MulticlassConfusionMatrix]), 10 updates."""
    import torchmetrics_forked_b200 as tm
    from torchmetrics_forked_b200.classification import MulticlassAccuracy, MulticlassConfusionMatrix

    for C in (5, 10):
        g = torch.Generator().manual_seed(C)
        coll = tm.MetricCollection([MulticlassAccuracy(C, average="micro"), MulticlassConfusionMatrix(C)]).cuda()
        allp, allt = [], []
        for _ in range(10):
            preds = torch.randn(10_000, C, generator=g).softmax(-1)
            target = torch.randint(C, (10_000,), generator=g)
            coll.update(preds.cuda(), target.cuda())
            allp.append(preds)
            allt.append(target)
        out = coll.compute()
        P, T = torch.cat(allp).numpy(), torch.cat(allt).numpy()
        assert np.array_equal(out["MulticlassConfusionMatrix"].cpu().numpy(), O.multiclass_confusion_matrix(P, T, C))
        tp, fp, tn, fn = O.multiclass_stat_scores(P, T, C, "micro")
        np.testing.assert_allclose(out["MulticlassAccuracy"].cpu().numpy(), O.accuracy_reduce(tp, fp, tn, fn, "micro"),
                                   rtol=RTOL)
        # forward returns the batch value and keeps accumulating
        coll.reset()
        batch = coll(allp[0].cuda(), allt[0].cuda())
        assert np.array_equal(batch["MulticlassConfusionMatrix"].cpu().numpy(),
                              O.multiclass_confusion_matrix(allp[0].numpy(), allt[0].numpy(), C))
        coll(allp[1].cuda(), allt[1].cuda())
        both = O.multiclass_confusion_matrix(torch.cat(allp[:2]).numpy(), torch.cat(allt[:2]).numpy(), C)
        assert np.array_equal(coll.compute()["MulticlassConfusionMatrix"].cpu().numpy(), both)


def test_compute_groups_share_state_in_place():
    """H11: members of a compute group alias one state tensor, so the kernels must add in place."""
    import torchmetrics_forked_b200 as tm
    from torchmetrics_forked_b200.classification import MulticlassAccuracy, MulticlassF1Score

    g = torch.Generator().manual_seed(3)
    coll = tm.MetricCollection([MulticlassAccuracy(7), MulticlassF1Score(7)]).cuda()
    P, T = [], []
    for _ in range(3):
        preds, target = torch.randn(5000, 7, generator=g), torch.randint(7, (5000,), generator=g)
        coll.update(preds.cuda(), target.cuda())
        P.append(preds)
        T.append(target)
    assert coll.compute_groups == {0: ["MulticlassAccuracy", "MulticlassF1Score"]}
    out = coll.compute()
    tp, fp, tn, fn = O.multiclass_stat_scores(torch.cat(P).numpy(), torch.cat(T).numpy(), 7, "macro")
    np.testing.assert_allclose(out["MulticlassAccuracy"].cpu().numpy(), O.accuracy_reduce(tp, fp, tn, fn, "macro"), rtol=RTOL)
    np.testing.assert_allclose(out["MulticlassF1Score"].cpu().numpy(), O.fbeta_reduce(tp, fp, tn, fn, 1.0, "macro"), rtol=RTOL)


def test_deferred_data_error_surfaces_at_compute():
    from torchmetrics_forked_b200.classification import MulticlassConfusionMatrix

    m = MulticlassConfusionMatrix(4, validate_args=False).cuda()
    m.update(torch.randn(10, 4).cuda(), torch.tensor([0, 1, 2, 3, 9, 0, 1, 2, 3, 0]).cuda())
    with pytest.raises(RuntimeError, match="outside"):
        m.compute()


def test_jaccard_config5_scaled():
    from torchmetrics_forked_b200.classification import MulticlassJaccardIndex

    g = torch.Generator().manual_seed(5)
    preds = torch.randn(2, 150, 96, 96, generator=g)
    target = torch.randint(150, (2, 96, 96), generator=g)
    m = MulticlassJaccardIndex(150).cuda()
    m.update(preds.cuda(), target.cuda())
    confmat = O.multiclass_confusion_matrix(preds.numpy(), target.numpy(), 150)
    assert np.array_equal(m.confmat.cpu().numpy(), confmat)
    np.testing.assert_allclose(m.compute().cpu().numpy(), O.jaccard_reduce(confmat, "macro"), rtol=RTOL)


@pytest.mark.parametrize("L", [4, 5, 7, 100, 251])
@pytest.mark.parametrize("kind", ["logits", "probs", "outlier"])
def test_multilabel_curve_shared_histogram_kernel(L, kind, F):
    """k2_multi ([N, L] row-major, CTA-wide shared histogram): label wrap-around inside a vector of 4 (L not a
    multiple of 4), ragged element counts, ignore_index, uneven thresholds, both passes of the sigmoid protocol
    and accumulation over two updates, against the oracle bit for bit."""
    from torchmetrics_forked_b200.classification import MultilabelPrecisionRecallCurve

    g = torch.Generator().manual_seed(1000 + L)
    thr_even = torch.linspace(0, 1, 50)
    thr_uneven = torch.sort(torch.rand(37, generator=g)).values
    for thr, tdtype, pdtype in [(thr_even, torch.int64, torch.float32), (thr_uneven, torch.uint8, torch.float32),
                                (thr_even, torch.int32, torch.bfloat16)]:
        m = MultilabelPrecisionRecallCurve(L, thresholds=thr, ignore_index=None if tdtype == torch.uint8 else -1).cuda()
        ign = m.ignore_index
        want = None
        for n in (3001, 77):  # n * L is odd for odd L: exercises the ragged batch and the n % 4 tail
            x = torch.randn(n, L, generator=g) * 2 if kind != "probs" else torch.rand(n, L, generator=g)
            if kind == "outlier":
                x = torch.rand(n, L, generator=g)
                x[n // 2, L // 2] = 1.25  # one value outside [0,1] flips the whole batch to sigmoid (H1)
            x = x.to(pdtype)
            t = torch.randint(2, (n, L), generator=g)
            if ign is not None:
                t[torch.rand(n, L, generator=g) < 0.2] = ign
            m.update(x.cuda(), t.to(tdtype).cuda())
            part = O.multilabel_curve_binned(x.float().numpy(), t.numpy(), L, thr.numpy(), ign,
                                             sigmoid_fn=lambda a, d=pdtype: torch.from_numpy(a).to(d).cuda().sigmoid().float().cpu().numpy())
            want = part if want is None else want + part
        assert np.array_equal(m.confmat.cpu().numpy(), want), (L, kind, tdtype, pdtype)


@pytest.mark.parametrize("case", ["logits", "labels", "strided_bf16", "ignore", "wide"])
def test_collection_shares_one_argmax(case):
    """a confusion-matrix family head and the stat-score family heads of a MetricCollection are
    fed by ONE pass over the logits (mode TMB_MC_BOTH + the shared finalize kernels).  Every state must equal the
    state of the same metric updated on its own, bit for bit; data errors still surface; other metrics of the
    collection are untouched."""
    from torchmetrics_forked_b200 import MetricCollection
    from torchmetrics_forked_b200 import classification as M

    g = torch.Generator().manual_seed(77)
    C = 300 if case == "wide" else 7
    ign = 2 if case == "ignore" else None

    def make():
        return [M.MulticlassAccuracy(C, average="micro", ignore_index=ign), M.MulticlassF1Score(C, average="macro", ignore_index=ign),
                M.MulticlassRecall(C, average="none", ignore_index=ign), M.MulticlassConfusionMatrix(C, ignore_index=ign),
                M.MulticlassJaccardIndex(C, ignore_index=ign), M.MulticlassAccuracy(C, top_k=2 if case != "labels" else 1, ignore_index=ign)]

    names = ["acc", "f1", "rec", "cm", "iou", "top2"]
    coll = MetricCollection(dict(zip(names, make()))).cuda()
    alone = {k: m.cuda() for k, m in zip(names, make())}
    for step in range(4):
        if case == "labels":
            preds, target = torch.randint(C, (501,), generator=g), torch.randint(C, (501,), generator=g)
        elif case == "strided_bf16":
            preds, target = torch.randn(6, C, 5, 9, generator=g).bfloat16(), torch.randint(C, (6, 5, 9), generator=g)
        else:
            preds, target = torch.randn(1003, C, generator=g), torch.randint(C, (1003,), generator=g)
        coll.update(preds.cuda(), target.cuda())
        for m in alone.values():
            m.update(preds.cuda(), target.cuda())
        if step == 0:
            continue
        plan = coll._shared_plan_cache
        assert plan is not None and plan[0] == "cm" and sorted(plan[1]) == ["acc", "f1"]
    # "rec" (average none) shares f1's [C] states, "iou" shares cm's, "top2" is its own group (top_k = 2)
    for k in names:
        a, b = coll[k], alone[k]
        for state in a._defaults:
            assert torch.equal(getattr(a, state), getattr(b, state)), (case, k, state)
        assert a._update_count == b._update_count == 4
    out = coll.compute()
    for k in names:
        assert torch.equal(out[k], alone[k].compute()), (case, k)
    # a bad target is still reported, by whichever participant validates
    bad_p, bad_t = (torch.randint(C, (10,)), torch.full((10,), C)) if case == "labels" else (torch.randn(10, C), torch.full((10,), C))
    if case != "strided_bf16":
        with pytest.raises(RuntimeError, match="target"):
            coll.update(bad_p.cuda(), bad_t.cuda())


def test_update_path_is_cuda_graph_capturable():
    """The update path enqueues kernels on the current stream and nothing else (no host read-back, no allocation after
    the first call): a MetricCollection update — including the shared-argmax path and a binned curve — can be captured
    once and replayed on new data written into the static input buffers."""
    from torchmetrics_forked_b200 import MetricCollection
    from torchmetrics_forked_b200 import classification as M

    g = torch.Generator().manual_seed(5)
    C, n = 10, 4096

    def make():
        return MetricCollection([M.MulticlassAccuracy(C, average="micro", validate_args=False),
                                 M.MulticlassConfusionMatrix(C, validate_args=False),
                                 M.MulticlassAUROC(C, thresholds=20, validate_args=False)]).cuda()

    batches = [(torch.randn(n, C, generator=g).cuda(), torch.randint(C, (n,), generator=g).cuda()) for _ in range(5)]
    eager, graphed = make(), make()
    for p, t in batches:
        eager.update(p, t)
    sp, st = batches[0][0].clone(), batches[0][1].clone()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):  # warm-up outside capture: group discovery, scratch / error-word allocation
        graphed.update(sp, st)
        graphed.update(sp, st)
    torch.cuda.current_stream().wait_stream(side)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        graphed.update(sp, st)

    def zero_states_in_place():  # reset() would hand the metrics NEW tensors; the graph holds the addresses of these
        for m in graphed.values(copy_state=False):
            for state in m._defaults:
                getattr(m, state).zero_()

    zero_states_in_place()
    for p, t in batches:
        sp.copy_(p)
        st.copy_(t)
        graph.replay()
    torch.cuda.synchronize()
    for name in ("MulticlassAccuracy", "MulticlassConfusionMatrix", "MulticlassAUROC"):
        for state in eager[name]._defaults:
            assert torch.equal(getattr(eager[name], state), getattr(graphed[name], state)), (name, state)


def test_forward_accumulates_in_place_and_returns_batch_values():
    """SURVEY §8 row f4, `forward()`: batch states are views of one zeroed buffer, the kernels fill them, one fused add
    folds them into the global tensors IN PLACE.  Batch values equal the functional results, the accumulated state
    equals update-only accumulation, also through a MetricCollection's forward."""
    from torchmetrics_forked_b200 import MetricCollection
    from torchmetrics_forked_b200 import classification as M
    from torchmetrics_forked_b200 import functional as F

    g = torch.Generator().manual_seed(21)
    C = 12
    acc, ref = M.MulticlassAccuracy(C, average="macro").cuda(), M.MulticlassAccuracy(C, average="macro").cuda()
    curve, curve_ref = M.BinaryAUROC(thresholds=50).cuda(), M.BinaryAUROC(thresholds=50).cuda()
    coll = MetricCollection([M.MulticlassPrecision(C), M.MulticlassRecall(C), M.MulticlassConfusionMatrix(C)]).cuda()
    tp_tensor = acc.tp
    for _ in range(3):
        p, t = torch.randn(700, C, generator=g).cuda(), torch.randint(C, (700,), generator=g).cuda()
        assert torch.equal(acc(p, t), F.multiclass_accuracy(p, t, C, average="macro"))
        ref.update(p, t)
        bp, bt = torch.randn(900, generator=g).cuda(), torch.randint(2, (900,), generator=g).cuda()
        assert torch.equal(curve(bp, bt), F.binary_auroc(bp, bt, thresholds=50))
        curve_ref.update(bp, bt)
        out = coll(p, t)
        assert torch.equal(out["MulticlassRecall"], F.multiclass_recall(p, t, C))
    assert acc.tp is tp_tensor  # same tensor object, accumulated in place
    for s in ("tp", "fp", "tn", "fn"):
        assert torch.equal(getattr(acc, s), getattr(ref, s))
    assert torch.equal(curve.confmat, curve_ref.confmat) and acc._update_count == 3
    assert torch.equal(acc.compute(), ref.compute())
    assert torch.equal(coll["MulticlassPrecision"].tp, ref.tp) and torch.equal(coll["MulticlassRecall"].fn, ref.fn)


def test_forward_on_grouped_collection_matches_standalone_metrics():
    """update -> forward and reset -> forward on a collection whose compute groups are formed (Precision + Recall share
    tp/fp/tn/fn, ConfusionMatrix + Jaccard share confmat): every state takes each batch exactly once."""
    from torchmetrics_forked_b200 import MetricCollection
    from torchmetrics_forked_b200 import classification as M

    g = torch.Generator().manual_seed(33)
    C = 7

    def make():
        return [M.MulticlassPrecision(C), M.MulticlassRecall(C), M.MulticlassConfusionMatrix(C),
                M.MulticlassJaccardIndex(C)]

    coll = MetricCollection(make()).cuda()
    alone = [m.cuda() for m in make()]
    batches = [(torch.randn(500, C, generator=g).cuda(), torch.randint(C, (500,), generator=g).cuda()) for _ in range(5)]

    def check():
        out = coll.compute()
        for m in alone:
            assert torch.equal(out[type(m).__name__], m.compute()), type(m).__name__
        assert torch.equal(coll["MulticlassRecall"].tp, alone[1].tp)
        assert torch.equal(coll["MulticlassJaccardIndex"].confmat, alone[3].confmat)

    coll.update(*batches[0])
    [m.update(*batches[0]) for m in alone]
    assert len(coll.compute_groups) == 2
    batch_vals = coll(*batches[1])          # update -> forward
    for m in alone:
        assert torch.equal(batch_vals[type(m).__name__], m(*batches[1]))
    check()
    coll.update(*batches[2])                # forward -> update links the members again
    [m.update(*batches[2]) for m in alone]
    check()
    coll.reset()
    [m.reset() for m in alone]
    for b in batches[3:]:                   # reset -> forward, forward
        coll(*b)
        [m(*b) for m in alone]
    check()


def test_collection_shares_one_argmax_without_confusion_matrix():
    """Stat-score heads that only differ in their averaging (micro keeps [1] states, macro [C]) share one pass too."""
    from torchmetrics_forked_b200 import MetricCollection
    from torchmetrics_forked_b200 import classification as M

    g = torch.Generator().manual_seed(78)
    C = 9

    def make():
        return {"acc": M.MulticlassAccuracy(C, average="micro"), "f1": M.MulticlassF1Score(C, average="macro"),
                "prec": M.MulticlassPrecision(C, average="weighted"), "spec": M.MulticlassSpecificity(C, average="micro")}

    coll, alone = MetricCollection(make()).cuda(), {k: m.cuda() for k, m in make().items()}
    for _ in range(3):
        p, t = torch.randn(777, C, generator=g).cuda(), torch.randint(C, (777,), generator=g).cuda()
        coll.update(p, t)
        for m in alone.values():
            m.update(p, t)
    plan = coll._shared_plan_cache
    assert plan is not None and plan[0] is None and sorted(plan[1]) == ["acc", "f1"]  # heads of the two groups
    out = coll.compute()
    for k, m in alone.items():
        for state in m._defaults:
            assert torch.equal(getattr(coll[k], state), getattr(m, state)), (k, state)
        assert torch.equal(out[k], m.compute()) and coll[k]._update_count == 3
