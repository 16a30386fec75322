"""ORACLE — CPU restatement (numpy) of the reference's classification-update hot path.

TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s CPU-baseline
legs may import this module, and only as the checker: the product path (`torchmetrics_forked_b200`) never
touches it and has no CPU fallback.

What it restates: torchmetrics 1.3.0dev (Viola-Siemens/torchmetrics-forked), paths relative to
/root/reference/src/torchmetrics/.  Every function names the reference lines it follows.  The reference is
pure Python on torch ATen, so "restating" means re-deriving each stage with integer / byte arithmetic in
numpy rather than calling the same ATen op.

Parity status: PINNED.  `tests/test_oracle_golden.py` checks this file against (a) golden vectors produced by
running the unmodified reference in the build container (`oracle/gen_golden.py` -> `tests/golden/*.npz`),
(b) the hard-coded known-answer vectors of the reference's own tests and doctests, and, when
/root/reference is mounted, (c) the live reference on random inputs (`tests/test_oracle_vs_reference.py`).

Floating point note: sigmoid bits differ between implementations.  Functions that may apply
sigmoid take `sigmoid_fn`; the default is torch's CPU kernel (what the reference executes on CPU), GPU
parity tests pass torch's CUDA kernel (what the reference executes on the GPU).
"""
from typing import Callable, List, Optional, Sequence, Tuple

import numpy as np

Array = np.ndarray
SigmoidFn = Callable[[Array], Array]


# ------------------------------------------------------------------------------------------ helpers
def torch_cpu_sigmoid(x: Array) -> Array:
    """`Tensor.sigmoid()` on CPU for a contiguous float32 array — the op the reference calls."""
    import torch

    return torch.from_numpy(np.ascontiguousarray(x, dtype=np.float32)).sigmoid().numpy()


def numpy_sigmoid(x: Array) -> Array:
    """Plain float32 `1 / (1 + exp(-x))`; differs from torch by an ulp on a fraction of inputs (H5)."""
    x = x.astype(np.float32)
    one = np.float32(1.0)
    with np.errstate(over="ignore"):
        return (one / (one + np.exp(-x))).astype(np.float32)


def argmax_first(x: Array, axis: int) -> Array:
    """`torch.argmax(dim)`: lowest index of the maximum, a NaN counts as the maximum (first NaN wins).

    Restated without relying on np.argmax's NaN behaviour: scan the axis keeping (best value, best index).
    """
    x = np.moveaxis(np.asarray(x), axis, 0)
    best_v = x[0].copy()
    best_i = np.zeros(x.shape[1:], dtype=np.int64)
    best_nan = np.isnan(best_v)
    for c in range(1, x.shape[0]):
        v = x[c]
        v_nan = np.isnan(v)
        with np.errstate(invalid="ignore"):
            take = (~best_nan) & (v_nan | (v > best_v))
        best_v = np.where(take, v, best_v)
        best_i = np.where(take, c, best_i)
        best_nan = best_nan | v_nan
    return best_i


def bincount(x: Array, minlength: int) -> Array:
    """`_bincount` (utilities/data.py:169-199): occurrences of each value, at least `minlength` entries."""
    x = np.asarray(x, dtype=np.int64).reshape(-1)
    if x.size and x.min() < 0:
        raise RuntimeError("bincount only supports 1-d non-negative integral inputs.")
    return np.bincount(x, minlength=minlength).astype(np.int64)


def _needs_sigmoid(x: Array) -> bool:
    """`not torch.all((preds >= 0) * (preds <= 1))` — true for any NaN as well."""
    with np.errstate(invalid="ignore"):
        return not bool(np.all((x >= 0) & (x <= 1)))


# ------------------------------------------------------------------------------------------ K1: multiclass
def multiclass_labels(preds: Array, target: Array, ignore_index: Optional[int]) -> Tuple[Array, Array]:
    """`_multiclass_confusion_matrix_format` (functional/classification/confusion_matrix.py:306-330):
    argmax over dim 1 when preds has one more dim than target, flatten, drop ignored targets."""
    preds = np.asarray(preds)
    target = np.asarray(target)
    if preds.ndim == target.ndim + 1:
        preds = argmax_first(preds, 1)
    preds = preds.reshape(-1).astype(np.int64)
    target = target.reshape(-1).astype(np.int64)
    if ignore_index is not None:
        keep = target != ignore_index
        preds, target = preds[keep], target[keep]
    return preds, target


def multiclass_confusion_matrix(preds: Array, target: Array, num_classes: int, ignore_index: Optional[int] = None) -> Array:
    """`_multiclass_confusion_matrix_update` (:333-337): bincount(target * C + pred).reshape(C, C)."""
    p, t = multiclass_labels(preds, target, ignore_index)
    bins = bincount(t * num_classes + p, num_classes**2)
    if bins.size != num_classes**2:
        raise RuntimeError(f"shape '[{num_classes}, {num_classes}]' is invalid for input of size {bins.size}")
    return bins.reshape(num_classes, num_classes)


def multiclass_stat_scores(
    preds: Array, target: Array, num_classes: int, average: Optional[str] = "macro", ignore_index: Optional[int] = None
) -> Tuple[Array, Array, Array, Array]:
    """`_multiclass_stat_scores_format/_update`, global, top_k=1 (functional/classification/stat_scores.py
    :325-341, :394-418) -> tp, fp, tn, fn (shape [C], or 0-d for average="micro")."""
    p, t = multiclass_labels(preds, target, ignore_index)
    if average == "micro":
        tp = np.int64((p == t).sum())
        fp = np.int64((p != t).sum())
        fn = np.int64((p != t).sum())
        tn = np.int64(num_classes * p.size - (fp + fn + tp))
        return tp, fp, tn, fn
    confmat = bincount(t * num_classes + p, num_classes**2)
    if confmat.size != num_classes**2:
        raise RuntimeError("label out of range")
    confmat = confmat.reshape(num_classes, num_classes)
    tp = np.diag(confmat).copy()
    fp = confmat.sum(0) - tp
    fn = confmat.sum(1) - tp
    tn = confmat.sum() - (fp + fn + tp)
    return tp, fp, tn, fn


# ------------------------------------------------------------------------------------------ K1b: binary / multilabel counts
def _threshold_labels(preds: Array, threshold: float, flag_over: Array, sigmoid_fn: SigmoidFn) -> Array:
    """Float preds -> 0/1 via conditional sigmoid and strict `> threshold`; integer preds pass through.

    `flag_over` selects the elements the [0,1] test looks at (H2).  The comparison is done in float32 with
    the threshold rounded to float32 (the reference compares a float32 tensor with a Python float)."""
    if not np.issubdtype(preds.dtype, np.floating):
        return preds.astype(np.int64)
    q = preds.astype(np.float32)
    if _needs_sigmoid(q[flag_over]):
        q = sigmoid_fn(q)
    with np.errstate(invalid="ignore"):
        return (q > np.float32(threshold)).astype(np.int64)


def binary_stat_scores(
    preds: Array, target: Array, threshold: float = 0.5, ignore_index: Optional[int] = None,
    sigmoid_fn: SigmoidFn = torch_cpu_sigmoid,
) -> Tuple[int, int, int, int]:  # fmt: skip
    """`_binary_stat_scores_format/_update` (stat_scores.py:90-131): sigmoid test over ALL elements,
    ignored targets become -1 and fall out of all four counts."""
    preds = np.asarray(preds).reshape(-1)
    target = np.asarray(target).reshape(-1).astype(np.int64)
    p = _threshold_labels(preds, threshold, np.ones(preds.shape, bool), sigmoid_fn)
    if ignore_index is not None:
        target = np.where(target == ignore_index, -1, target)
    tp = int(((target == p) & (target == 1)).sum())
    fn = int(((target != p) & (target == 1)).sum())
    fp = int(((target != p) & (target == 0)).sum())
    tn = int(((target == p) & (target == 0)).sum())
    return tp, fp, tn, fn


def multilabel_stat_scores(
    preds: Array, target: Array, num_labels: int, threshold: float = 0.5, ignore_index: Optional[int] = None,
    sigmoid_fn: SigmoidFn = torch_cpu_sigmoid,
) -> Tuple[Array, Array, Array, Array]:  # fmt: skip
    """`_multilabel_stat_scores_format/_update` (stat_scores.py:647-681), global: label axis is dim 1."""
    preds = np.asarray(preds)
    target = np.asarray(target).astype(np.int64)
    p = _threshold_labels(preds, threshold, np.ones(preds.shape, bool), sigmoid_fn)
    p = p.reshape(p.shape[0], num_labels, -1)
    t = target.reshape(target.shape[0], num_labels, -1)
    if ignore_index is not None:
        t = np.where(t == ignore_index, -1, t)
    tp = ((t == p) & (t == 1)).sum((0, 2))
    fn = ((t != p) & (t == 1)).sum((0, 2))
    fp = ((t != p) & (t == 0)).sum((0, 2))
    tn = ((t == p) & (t == 0)).sum((0, 2))
    return tp.astype(np.int64), fp.astype(np.int64), tn.astype(np.int64), fn.astype(np.int64)


def binary_confusion_matrix(
    preds: Array, target: Array, threshold: float = 0.5, ignore_index: Optional[int] = None,
    sigmoid_fn: SigmoidFn = torch_cpu_sigmoid,
) -> Array:  # fmt: skip
    """`_binary_confusion_matrix_format/_update` (confusion_matrix.py:118-153): ignored samples are removed
    BEFORE the sigmoid test; bincount(2 * target + pred).reshape(2, 2)."""
    preds = np.asarray(preds).reshape(-1)
    target = np.asarray(target).reshape(-1).astype(np.int64)
    if ignore_index is not None:
        keep = target != ignore_index
        preds, target = preds[keep], target[keep]
    p = _threshold_labels(preds, threshold, np.ones(preds.shape, bool), sigmoid_fn)
    return bincount(target * 2 + p, 4).reshape(2, 2)


def multilabel_confusion_matrix(
    preds: Array, target: Array, num_labels: int, threshold: float = 0.5, ignore_index: Optional[int] = None,
    sigmoid_fn: SigmoidFn = torch_cpu_sigmoid,
) -> Array:  # fmt: skip
    """`_multilabel_confusion_matrix_format/_update` (confusion_matrix.py:486-526): sigmoid test over all
    elements, ignored positions dropped, bincount(2t + p + 4*label).reshape(L, 2, 2)."""
    preds = np.asarray(preds)
    target = np.asarray(target).astype(np.int64)
    p = _threshold_labels(preds, threshold, np.ones(preds.shape, bool), sigmoid_fn)
    p = np.moveaxis(p, 1, -1).reshape(-1, num_labels)
    t = np.moveaxis(target, 1, -1).reshape(-1, num_labels)
    mapping = 2 * t + p + 4 * np.arange(num_labels)
    if ignore_index is not None:
        mapping = mapping[t != ignore_index]
    return bincount(mapping.reshape(-1), 4 * num_labels).reshape(num_labels, 2, 2)


# ------------------------------------------------------------------------------------------ K2: binned curves
def _bin_counts(q: Array, t: Array, thresholds: Array) -> Array:
    """[T, 2, 2] with cell [j, target, q >= thr_j]; restated through the bin index k = #{j: thr_j <= q}
    (what `preds >= thresholds[i]` sums to, precision_recall_curve.py:244-249) and suffix sums, so that it is
    an independent derivation from the per-threshold loop of the reference."""
    thr = np.asarray(thresholds, dtype=np.float32)
    order = np.argsort(thr, kind="stable")
    thr_sorted = thr[order]
    T = thr.size
    with np.errstate(invalid="ignore"):
        k = np.searchsorted(thr_sorted, q, side="right")  # NaN sorts last in numpy -> fix below
    k = np.where(np.isnan(q), 0, k)
    out = np.zeros((T, 2, 2), dtype=np.int64)
    for cls in (0, 1):
        hist = np.bincount(k[t == cls], minlength=T + 1).astype(np.int64)
        above = hist[::-1].cumsum()[::-1]  # above[k] = #{samples with bin >= k}
        ge = above[1:]  # threshold j is passed by the samples with bin >= j + 1
        out[order, cls, 1] = ge
        out[order, cls, 0] = hist.sum() - ge
    return out


def binary_curve_binned(
    preds: Array, target: Array, thresholds: Array, ignore_index: Optional[int] = None,
    sigmoid_fn: SigmoidFn = torch_cpu_sigmoid,
) -> Array:  # fmt: skip
    """`_binary_precision_recall_curve_format/_update` (precision_recall_curve.py:162-250) -> [T, 2, 2].
    Ignored samples are removed before the sigmoid test (:176-184).  Any target != 1 counts as negative in
    the loop variant (:241); validated inputs only hold 0/1."""
    q = np.asarray(preds).reshape(-1).astype(np.float32)
    t = np.asarray(target).reshape(-1).astype(np.int64)
    if ignore_index is not None:
        keep = t != ignore_index
        q, t = q[keep], t[keep]
    if _needs_sigmoid(q):
        q = sigmoid_fn(q)
    return _bin_counts(q, (t == 1).astype(np.int64), thresholds)


def multilabel_curve_binned(
    preds: Array, target: Array, num_labels: int, thresholds: Array, ignore_index: Optional[int] = None,
    sigmoid_fn: SigmoidFn = torch_cpu_sigmoid,
) -> Array:  # fmt: skip
    """`_multilabel_precision_recall_curve_format/_update` (precision_recall_curve.py:739-793)
    -> [T, L, 2, 2]; sigmoid test over all elements, ignored positions dropped from every cell."""
    preds = np.asarray(preds).astype(np.float32)
    target = np.asarray(target).astype(np.int64)
    q = np.moveaxis(preds, 1, -1).reshape(-1, num_labels)
    t = np.moveaxis(target, 1, -1).reshape(-1, num_labels)
    if _needs_sigmoid(q):
        q = sigmoid_fn(np.ascontiguousarray(q))
    T = np.asarray(thresholds).size
    out = np.zeros((T, num_labels, 2, 2), dtype=np.int64)
    for l in range(num_labels):
        ql, tl = q[:, l], t[:, l]
        if ignore_index is not None:
            keep = tl != ignore_index
            ql, tl = ql[keep], tl[keep]
        out[:, l] = _bin_counts(ql, tl, thresholds)
    return out


# ------------------------------------------------------------------------------------------ samplewise stat scores (row f3)
def _counts(t: Array, p: Array, axes) -> Tuple[Array, Array, Array, Array]:
    tp = ((t == p) & (t == 1)).sum(axes)
    fn = ((t != p) & (t == 1)).sum(axes)
    fp = ((t != p) & (t == 0)).sum(axes)
    tn = ((t == p) & (t == 0)).sum(axes)
    return tp.astype(np.int64), fp.astype(np.int64), tn.astype(np.int64), fn.astype(np.int64)


def binary_stat_scores_samplewise(preds: Array, target: Array, threshold: float = 0.5,
                                  ignore_index: Optional[int] = None,
                                  sigmoid_fn: SigmoidFn = torch_cpu_sigmoid) -> Tuple[Array, Array, Array, Array]:
    """`_binary_stat_scores_update` with `multidim_average="samplewise"` (stat_scores.py:116-129, sum_dim = 1): `[N]`
    counts; the sigmoid decision still looks at the whole batch."""
    preds = np.asarray(preds)
    p = _threshold_labels(preds, threshold, np.ones(preds.shape, bool), sigmoid_fn).reshape(preds.shape[0], -1)
    t = np.asarray(target).astype(np.int64).reshape(preds.shape[0], -1)
    if ignore_index is not None:
        t = np.where(t == ignore_index, -1, t)
    return _counts(t, p, 1)


def multilabel_stat_scores_samplewise(preds: Array, target: Array, num_labels: int, threshold: float = 0.5,
                                      ignore_index: Optional[int] = None,
                                      sigmoid_fn: SigmoidFn = torch_cpu_sigmoid) -> Tuple[Array, Array, Array, Array]:
    """`_multilabel_stat_scores_update`, samplewise (stat_scores.py:672-679, sum_dim = 2): `[N, L]` counts."""
    preds = np.asarray(preds)
    p = _threshold_labels(preds, threshold, np.ones(preds.shape, bool), sigmoid_fn)
    p = p.reshape(p.shape[0], num_labels, -1)
    t = np.asarray(target).astype(np.int64).reshape(p.shape[0], num_labels, -1)
    if ignore_index is not None:
        t = np.where(t == ignore_index, -1, t)
    return _counts(t, p, 2)


def multiclass_stat_scores_samplewise(preds: Array, target: Array, num_classes: int,
                                      ignore_index: Optional[int] = None) -> Tuple[Array, Array, Array, Array]:
    """`_multiclass_stat_scores_update`, samplewise, `top_k = 1` (stat_scores.py:363-393 with sum_dim = [1]): `[N, C]`
    counts from the one-hot prediction / target; ignored positions leave all four."""
    preds = np.asarray(preds)
    target = np.asarray(target).astype(np.int64)
    labels = argmax_first(preds.astype(np.float32), 1) if preds.ndim == target.ndim + 1 else preds.astype(np.int64)
    n = target.shape[0]
    labels, t = labels.reshape(n, -1), target.reshape(n, -1)
    classes = np.arange(num_classes)[None, None, :]
    keep = np.ones(t.shape, bool) if ignore_index is None else t != ignore_index
    hot_t = (t[..., None] == classes) & keep[..., None]
    hot_p = (labels[..., None] == classes) & keep[..., None]
    tp = (hot_t & hot_p).sum(1)
    fn = (hot_t & ~hot_p).sum(1)
    fp = (~hot_t & hot_p).sum(1)
    tn = (keep[..., None] & ~hot_t & ~hot_p).sum(1)
    return tp.astype(np.int64), fp.astype(np.int64), tn.astype(np.int64), fn.astype(np.int64)


# ------------------------------------------------------------------------------------------ top-k stat scores (row f3)
def topk_sets(rows: Array, k: int) -> Array:
    """`select_topk` (utilities/data.py:115-139) as a boolean [M, C] membership matrix: the k largest scores of a row,
    NaN ranking highest, -0.0 below +0.0, ties broken towards the lower class index (what torch.topk's CUDA
    radix select + gather picks; the CPU order only differs when there IS a tie at the k-th place)."""
    rows = np.ascontiguousarray(rows, dtype=np.float32)
    bits = rows.view(np.uint32).astype(np.int64)
    # torch's radix-select key (TopKTypeConfig<float>::convert): flip all bits of negatives, the sign bit of the
    # rest, NaN -> all ones; so -0.0 ranks below +0.0
    key = np.where(bits & 0x80000000, bits ^ 0xFFFFFFFF, bits | 0x80000000)
    key = np.where(np.isnan(rows), 0xFFFFFFFF, key)
    order = np.argsort(-key, axis=1, kind="stable")  # stable: equal keys keep ascending class order
    member = np.zeros(rows.shape, dtype=bool)
    np.put_along_axis(member, order[:, :k], True, axis=1)
    return member


def multiclass_stat_scores_topk(preds: Array, target: Array, num_classes: int, top_k: int,
                                ignore_index: Optional[int] = None) -> Tuple[Array, Array, Array, Array]:
    """`_multiclass_stat_scores_update` with `top_k > 1`, global (stat_scores.py:363-393): k-hot prediction against
    the one-hot target, per class; rows whose target is ignored leave all four counts."""
    q = np.moveaxis(np.asarray(preds).astype(np.float32), 1, -1).reshape(-1, num_classes)
    t = np.asarray(target).astype(np.int64).reshape(-1)
    if ignore_index is not None:
        keep = t != ignore_index
        q, t = q[keep], t[keep]
    member = topk_sets(q, top_k)
    hot = t[:, None] == np.arange(num_classes)[None, :]
    tp = (hot & member).sum(0).astype(np.int64)
    fn = (hot & ~member).sum(0).astype(np.int64)
    fp = (~hot & member).sum(0).astype(np.int64)
    tn = (~hot & ~member).sum(0).astype(np.int64)
    return tp, fp, tn, fn


def multiclass_stat_scores_samplewise_topk(preds: Array, target: Array, num_classes: int, top_k: int,
                                           ignore_index: Optional[int] = None) -> Tuple[Array, Array, Array, Array]:
    """`_multiclass_stat_scores_update` with `top_k > 1` AND `multidim_average="samplewise"` (stat_scores.py:363-393,
    `sum_dim = [1]`): per sample `[N, C]` counts of the k-hot prediction against the one-hot target over the sample's
    positions; positions whose target is ignored leave all four counts."""
    q = np.asarray(preds).astype(np.float32)
    n = q.shape[0]
    q = np.moveaxis(q.reshape(n, num_classes, -1), 1, -1)  # [N, S, C]
    t = np.asarray(target).astype(np.int64).reshape(n, -1)
    out = [np.zeros((n, num_classes), np.int64) for _ in range(4)]
    for i in range(n):
        keep = np.ones(t.shape[1], bool) if ignore_index is None else t[i] != ignore_index
        member = topk_sets(q[i][keep], top_k)
        hot = t[i][keep][:, None] == np.arange(num_classes)[None, :]
        out[0][i] = (hot & member).sum(0)
        out[3][i] = (hot & ~member).sum(0)
        out[1][i] = (~hot & member).sum(0)
        out[2][i] = (~hot & ~member).sum(0)
    return out[0], out[1], out[2], out[3]


# ------------------------------------------------------------------------------------------ K5: multiclass curves (row f2)
def torch_cpu_softmax(x: Array) -> Array:
    """Row softmax with the bits of torch's CPU kernel (what the golden vectors of the CPU reference contain)."""
    import torch

    return torch.from_numpy(np.ascontiguousarray(x, dtype=np.float32)).softmax(1).numpy()


def multiclass_rows(preds: Array, target: Array, num_classes: int, ignore_index: Optional[int] = None,
                    softmax_fn: SigmoidFn = torch_cpu_softmax) -> Tuple[Array, Array]:
    """`_multiclass_precision_recall_curve_format` (precision_recall_curve.py:423-458) up to the softmax:
    `[M', C]` float32 probabilities and `[M']` targets — extra dimensions folded into the rows, rows with
    `target == ignore_index` REMOVED first, softmax over the classes iff any remaining value is outside [0,1]."""
    q = np.moveaxis(np.asarray(preds).astype(np.float32), 1, -1).reshape(-1, num_classes)
    t = np.asarray(target).astype(np.int64).reshape(-1)
    if ignore_index is not None:
        keep = t != ignore_index
        q, t = q[keep], t[keep]
    if q.size and _needs_sigmoid(q):
        q = softmax_fn(np.ascontiguousarray(q))
    return q, t


def multiclass_curve_binned(preds: Array, target: Array, num_classes: int, thresholds: Array,
                            ignore_index: Optional[int] = None, softmax_fn: SigmoidFn = torch_cpu_softmax,
                            micro: bool = False) -> Array:
    """`_multiclass_precision_recall_curve_update` (:461-527) -> `[T, C, 2, 2]`: class c against the one-hot target
    `target == c`; `micro` -> the flattened binary task, `[T, 2, 2]` (:453-455, :473-474)."""
    q, t = multiclass_rows(preds, target, num_classes, ignore_index, softmax_fn)
    hot = (t[:, None] == np.arange(num_classes)[None, :]).astype(np.int64)
    if micro:
        return _bin_counts(q.reshape(-1), hot.reshape(-1), thresholds)
    T = np.asarray(thresholds).size
    out = np.zeros((T, num_classes, 2, 2), dtype=np.int64)
    for c in range(num_classes):
        out[:, c] = _bin_counts(q[:, c], hot[:, c], thresholds)
    return out


def multiclass_clf_curves(probs: Array, target: Array) -> List[Tuple[Array, Array, Array]]:
    """The per-class loop of `_multiclass_precision_recall_curve_compute` (:548-555): column c of the stored
    probabilities against `pos_label = c`."""
    probs = np.asarray(probs)
    target = np.asarray(target)
    return [binary_clf_curve(probs[:, c], target, pos_label=c) for c in range(probs.shape[1])]


# ------------------------------------------------------------------------------------------ K3: exact curves
def exact_state(preds: Array, sigmoid_fn: SigmoidFn = torch_cpu_sigmoid) -> Array:
    """The scores `update` stores in exact mode: sigmoid over the whole batch iff any value is outside
    [0,1] (precision_recall_curve.py:183-184, :756-757)."""
    q = np.asarray(preds).astype(np.float32)
    return sigmoid_fn(q) if _needs_sigmoid(q) else q


def binary_clf_curve(preds: Array, target: Array, pos_label: int = 1) -> Tuple[Array, Array, Array]:
    """`_binary_clf_curve` (precision_recall_curve.py:28-80) -> (fps, tps, thresholds), all float32.

    Scores descending with NaN first (torch.argsort(descending=True)); a boundary sits wherever
    `preds[i+1] - preds[i]` is non-zero — NaN and inf-inf differences count as non-zero — plus the last
    index; tps = cumsum(target == pos_label) at the boundaries as float32 (the reference's cumsum runs in
    float32 because the targets are multiplied by the Python float 1.0, H6); fps = 1 + idx - tps.
    """
    q = np.asarray(preds).reshape(-1).astype(np.float32)
    t = np.asarray(target).reshape(-1)
    n = q.size
    nan = np.isnan(q)
    order = np.concatenate([np.flatnonzero(nan), np.flatnonzero(~nan)[np.argsort(-q[~nan], kind="stable")]])
    q, t = q[order], t[order]
    with np.errstate(invalid="ignore"):
        diff = q[1:] - q[:-1]
        distinct = np.flatnonzero(~(diff == 0))
    idx = np.concatenate([distinct, [n - 1]]).astype(np.int64)
    pos = (t == pos_label).astype(np.int64)
    tps_int = np.cumsum(pos)[idx]
    tps = tps_int.astype(np.float32)  # exact below 2**24, where the reference's float32 cumsum is exact too
    fps = (1 + idx).astype(np.float32) - tps
    return fps.astype(np.float32), tps, q[idx]


def multilabel_clf_curves(
    preds: Array, target: Array, ignore_index: Optional[int] = None
) -> List[Tuple[Array, Array, Array]]:
    """The per-label loop of `_multilabel_roc_compute` / `_multilabel_precision_recall_curve_compute`
    (roc.py:345-355, precision_recall_curve.py:819-829): column i, ignored rows removed, pos_label 1."""
    preds = np.asarray(preds)
    target = np.asarray(target)
    out = []
    for i in range(preds.shape[1]):
        p, t = preds[:, i], target[:, i]
        if ignore_index is not None:
            keep = t != ignore_index
            p, t = p[keep], t[keep]
        out.append(binary_clf_curve(p, t, 1))
    return out


# ------------------------------------------------------------------------------------------ reducers (float64, compared at 1e-6)
def _safe_divide(num: Array, denom: Array) -> Array:
    """utilities/compute.py:46-55."""
    num = np.asarray(num, dtype=np.float64)
    denom = np.asarray(denom, dtype=np.float64).copy()
    denom[denom == 0] = 1
    return num / denom


def accuracy_reduce(tp, fp, tn, fn, average: Optional[str], multilabel: bool = False) -> Array:
    """`_accuracy_reduce` (functional/classification/accuracy.py:37-86), global."""
    tp, fp, tn, fn = (np.asarray(v, dtype=np.float64) for v in (tp, fp, tn, fn))
    if average == "binary":
        return _safe_divide(tp + tn, tp + tn + fp + fn)
    if average == "micro":
        tp, fn = tp.sum(), fn.sum()
        if multilabel:
            fp, tn = fp.sum(), tn.sum()
            return _safe_divide(tp + tn, tp + tn + fp + fn)
        return _safe_divide(tp, tp + fn)
    score = _safe_divide(tp + tn, tp + tn + fp + fn) if multilabel else _safe_divide(tp, tp + fn)
    return _adjust_weights(score, average, multilabel, tp, fp, fn)


def fbeta_reduce(tp, fp, tn, fn, beta: float, average: Optional[str], multilabel: bool = False) -> Array:
    """`_fbeta_reduce` (functional/classification/f_beta.py:37-57), global."""
    tp, fp, tn, fn = (np.asarray(v, dtype=np.float64) for v in (tp, fp, tn, fn))
    b2 = beta**2
    if average == "micro":
        tp, fn, fp = tp.sum(), fn.sum(), fp.sum()
    score = _safe_divide((1 + b2) * tp, (1 + b2) * tp + b2 * fn + fp)
    if average in ("binary", "micro"):
        return score
    return _adjust_weights(score, average, multilabel, tp, fp, fn)


def _adjust_weights(score, average, multilabel, tp, fp, fn) -> Array:
    """`_adjust_weights_safe_divide` (utilities/compute.py:58-69)."""
    if average is None or average == "none":
        return score
    if average == "weighted":
        weights = tp + fn
    else:
        weights = np.ones_like(score)
        if not multilabel:
            weights[tp + fp + fn == 0] = 0.0
    return _safe_divide(weights * score, weights.sum(-1, keepdims=True)).sum(-1)


def jaccard_reduce(confmat: Array, average: Optional[str], ignore_index: Optional[int] = None) -> Array:
    """`_jaccard_index_reduce` (functional/classification/jaccard.py:38-94)."""
    confmat = np.asarray(confmat, dtype=np.float64)
    if average == "binary":
        return confmat[1, 1] / (confmat[0, 1] + confmat[1, 0] + confmat[1, 1])
    ignore_cond = ignore_index is not None and 0 <= ignore_index < confmat.shape[0]
    multilabel = confmat.ndim == 3
    if multilabel:
        num = confmat[:, 1, 1]
        denom = confmat[:, 1, 1] + confmat[:, 0, 1] + confmat[:, 1, 0]
    else:
        num = np.diag(confmat)
        denom = confmat.sum(0) + confmat.sum(1) - num
    if average == "micro":
        num = num.sum()
        denom = denom.sum() - (denom[ignore_index] if ignore_cond else 0.0)
    jac = _safe_divide(num, denom)
    if average in (None, "none", "micro"):
        return jac
    if average == "weighted":
        weights = confmat[:, 1, 1] + confmat[:, 1, 0] if multilabel else confmat.sum(1)
    else:
        weights = np.ones_like(jac)
        if ignore_cond:
            weights[ignore_index] = 0.0
        if not multilabel:
            weights[confmat.sum(1) + confmat.sum(0) == 0] = 0.0
    return ((weights * jac) / weights.sum()).sum()


def _trapz(y: Array, x: Array) -> float:
    y, x = np.asarray(y, dtype=np.float64), np.asarray(x, dtype=np.float64)
    return float(((x[1:] - x[:-1]) * (y[1:] + y[:-1]) / 2).sum())


def binary_auroc_binned(confmat: Array) -> float:
    """`_binary_roc_compute` + `_binary_auroc_compute`, binned branch, max_fpr=None (roc.py:45-52,
    auroc.py:88-90)."""
    c = np.asarray(confmat, dtype=np.float64)
    tps, fps, fns, tns = c[:, 1, 1], c[:, 0, 1], c[:, 1, 0], c[:, 0, 0]
    tpr = _safe_divide(tps, tps + fns)[::-1]
    fpr = _safe_divide(fps, fps + tns)[::-1]
    return _trapz(tpr, fpr)


def binary_auroc_exact(fps: Array, tps: Array) -> float:
    """Exact branch (roc.py:53-78, auroc.py:88-90): prepend (0,0), normalise by the totals, trapz."""
    tps = np.concatenate([[0.0], np.asarray(tps, dtype=np.float64)])
    fps = np.concatenate([[0.0], np.asarray(fps, dtype=np.float64)])
    fpr = fps / fps[-1] if fps[-1] > 0 else np.zeros_like(fps)
    tpr = tps / tps[-1] if tps[-1] > 0 else np.zeros_like(tps)
    return _trapz(tpr, fpr)


def binary_average_precision_binned(confmat: Array) -> float:
    """`_binary_precision_recall_curve_compute` binned + `_binary_average_precision_compute`
    (precision_recall_curve.py:264-272, average_precision.py:70-75)."""
    c = np.asarray(confmat, dtype=np.float64)
    tps, fps, fns = c[:, 1, 1], c[:, 0, 1], c[:, 1, 0]
    precision = np.concatenate([_safe_divide(tps, tps + fps), [1.0]])
    recall = np.concatenate([_safe_divide(tps, tps + fns), [0.0]])
    return float(-((recall[1:] - recall[:-1]) * precision[:-1]).sum())


def binary_average_precision_exact(fps: Array, tps: Array) -> float:
    """Exact branch (precision_recall_curve.py:274-283, average_precision.py:70-75)."""
    tps = np.asarray(tps, dtype=np.float64)
    fps = np.asarray(fps, dtype=np.float64)
    precision = np.concatenate([(tps / (tps + fps))[::-1], [1.0]])
    recall = np.concatenate([(tps / tps[-1])[::-1], [0.0]])
    return float(-((recall[1:] - recall[:-1]) * precision[:-1]).sum())


# ------------------------------------------------------------------------------------------ K4: sync
def sync_sum(states: Sequence[Array]) -> Array:
    """`_sync_dist` with `dim_zero_sum` (metric.py:423-453): gather every rank's tensor, stack, sum."""
    return np.stack([np.asarray(s) for s in states]).sum(0)


def sync_cat(states: Sequence[Optional[Array]]) -> Optional[Array]:
    """`_sync_dist` with `dim_zero_cat`: rank-ordered concatenation along dim 0
    (utilities/distributed.py:144-147); ranks with an empty list contribute nothing."""
    parts = [np.asarray(s) for s in states if s is not None]
    return np.concatenate(parts, 0) if parts else None
