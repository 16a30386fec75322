"""Precision-recall curves: binned (threshold histogram) and exact (sort + scan) state updates.

Reference: functional/classification/precision_recall_curve.py.

* binned mode (``thresholds`` given): the reference compares every sample with every threshold
  (:210-250, :785-793).  Here one pass computes each sample's bin against the metric's own threshold
  buffer and a finalize step turns the histogram into the ``[T, (L,) 2, 2]`` state, added in place.
* exact mode (``thresholds=None``): ``update`` stores the (conditionally sigmoided) scores; ``compute``
  replaces ``_binary_clf_curve`` (:28-80: argsort, gather, diff/where, cumsum) by a batched radix sort
  and a segmented scan that emit ``(fps, tps, thresholds)`` for every label at once.

The curve post-processing (``precision = tps / (tps + fps)`` ...) is kept as the reference's torch
expressions on the kernels' outputs.
"""
from typing import List, Optional, Sequence, Tuple, Union

import torch
from torch import Tensor
from typing_extensions import Literal

from torchmetrics_forked_b200 import _native as nat
from torchmetrics_forked_b200 import kernels
from torchmetrics_forked_b200.utilities.checks import _check_same_shape
from torchmetrics_forked_b200.utilities.compute import _safe_divide, interp
from torchmetrics_forked_b200.utilities.enums import ClassificationTask

_LOOP_SWITCH = 50_000  # above this many samples the reference compares in the dtype of `preds` (:203-207)


# ------------------------------------------------------------------------------------------ shared helpers
def _adjust_threshold_arg(
    thresholds: Optional[Union[int, List[float], Tensor]] = None, device: Optional[torch.device] = None
) -> Optional[Tensor]:
    """int -> ``linspace(0, 1, n)``, list -> tensor, tensor / None unchanged (reference :83-91)."""
    if isinstance(thresholds, int):
        return torch.linspace(0, 1, thresholds, device=device)
    if isinstance(thresholds, list):
        return torch.tensor(thresholds, device=device)
    return thresholds


def _threshold_plan(thresholds: Tensor, device: torch.device) -> Tuple[Tensor, Optional[Tensor]]:
    """Ascending float32 copy of ``thresholds`` on ``device`` plus the int32 map sorted position -> original row.

    The kernels bin against an ascending buffer; users may pass any order (each threshold is evaluated
    independently in the reference), so the finalize step scatters through the permutation.  The
    permutation is ``None`` for thresholds that are already ascending (the usual linspace).
    """
    thr = thresholds.detach().to(dtype=torch.float32, device="cpu").contiguous()
    if thr.numel() < 2 or bool((thr[1:] >= thr[:-1]).all()):
        return thr.to(device), None
    ordered, perm = torch.sort(thr, stable=True)
    return ordered.to(device), perm.to(torch.int32).to(device)


def _curve_thresholds_for(preds: Tensor, thr_sorted: Tensor, dimensioned_compare: bool) -> Tensor:
    """Thresholds as the reference's comparison sees them.

    ``preds >= thresholds[i]`` with a 0-d threshold is evaluated in the dtype of ``preds``, so for
    half / bfloat16 scores the loop variant (:244) rounds each threshold first; the vectorised variants
    compare two dimensioned tensors and promote to float32.
    """
    if preds.dtype in (torch.float32, torch.float64) or dimensioned_compare:
        return thr_sorted
    return thr_sorted.to(preds.dtype).to(torch.float32)


def _dimensioned_compare(
    preds: Tensor, target: Tensor, ignore_index: Optional[int], per_kept: int, switch: int
) -> bool:
    """Does the reference take its vectorised variant (two dimensioned tensors -> float32 compare) or its threshold
    loop (0-d threshold -> compare in the dtype of ``preds``)?  It decides from the number of samples / rows LEFT
    after the ``ignore_index`` filter (``precision_recall_curve.py:203, 473``; ``per_kept`` comparisons each).  The
    answer only matters for 16-bit scores.  Without ``ignore_index``, or when even the unfiltered count is below the
    switch, it is known on the host; otherwise the kept count is read back — the reference synchronises in the same
    place (its boolean-mask filter, ``:177-179, 445-447``)."""
    if preds.dtype in (torch.float32, torch.float64):
        return True  # float32 thresholds either way
    total = target.numel()
    if total * per_kept <= switch:
        return True
    if ignore_index is None:
        return False
    kept = int((target != ignore_index).sum())
    return kept * per_kept <= switch


def _binary_clf_curve(
    preds: Tensor,
    target: Tensor,
    sample_weights: Optional[Union[Sequence, Tensor]] = None,
    pos_label: int = 1,
) -> Tuple[Tensor, Tensor, Tensor]:
    """``(fps, tps, thresholds)`` at every distinct score, scores descending (reference :28-80).

    ``tps`` / ``fps`` are float32 like the reference's (it multiplies the targets by the float ``1.0``,
    :63-78); they are exact integer counts cast to float32 (H6).
    """
    if sample_weights is not None:
        raise NotImplementedError("`sample_weights` is not used on the metric path and is not supported here")
    if isinstance(preds, Tensor) and preds.ndim > target.ndim:
        preds = preds[:, 0]
    fps, tps, thr = kernels.clf_curves(kernels.as_columns(preds), kernels.as_columns(target), None, pos_label)
    return fps[0], tps[0], thr[0]


def _curve_from_counts(tps: Tensor, fps: Tensor, fns: Tensor, n_extra: Tuple[int, ...]) -> Tuple[Tensor, Tensor]:
    """Binned precision / recall with the closing (1, 0) point appended (reference :264-272, :544-553)."""
    precision = _safe_divide(tps, tps + fps)
    recall = _safe_divide(tps, tps + fns)
    precision = torch.cat([precision, torch.ones(*n_extra, dtype=precision.dtype, device=precision.device)])
    recall = torch.cat([recall, torch.zeros(*n_extra, dtype=recall.dtype, device=recall.device)])
    return precision, recall


# ------------------------------------------------------------------------------------------ binary
def _binary_precision_recall_curve_arg_validation(
    thresholds: Optional[Union[int, List[float], Tensor]] = None, ignore_index: Optional[int] = None
) -> None:
    """Reference: precision_recall_curve.py:94-122."""
    if thresholds is not None and not isinstance(thresholds, (list, int, Tensor)):
        raise ValueError(
            "Expected argument `thresholds` to either be an integer, list of floats or"
            f" tensor of floats, but got {thresholds}"
        )
    if isinstance(thresholds, int) and thresholds < 2:
        raise ValueError(
            f"If argument `thresholds` is an integer, expected it to be larger than 1, but got {thresholds}"
        )
    if isinstance(thresholds, list) and not all(isinstance(t, float) and 0 <= t <= 1 for t in thresholds):
        raise ValueError(
            "If argument `thresholds` is a list, expected all elements to be floats in the [0,1] range,"
            f" but got {thresholds}"
        )
    if isinstance(thresholds, Tensor) and not thresholds.ndim == 1:
        raise ValueError("If argument `thresholds` is an tensor, expected the tensor to be 1d")
    if ignore_index is not None and not isinstance(ignore_index, int):
        raise ValueError(f"Expected argument `ignore_index` to either be `None` or an integer, but got {ignore_index}")


def _binary_precision_recall_curve_tensor_validation(
    preds: Tensor, target: Tensor, ignore_index: Optional[int] = None
) -> None:
    """Shape / dtype checks of :125-147; the {0,1} check of the targets (:149-159) runs in the kernel."""
    _check_same_shape(preds, target)
    if target.is_floating_point():
        raise ValueError(
            "Expected argument `target` to be an int or long tensor with ground truth labels"
            f" but got tensor with dtype {target.dtype}"
        )
    if not preds.is_floating_point():
        raise ValueError(
            "Expected argument `preds` to be an floating tensor with probability/logit scores,"
            f" but got tensor with dtype {preds.dtype}"
        )


def _binary_precision_recall_curve_update_(
    preds: Tensor, target: Tensor, thr_sorted: Tensor, perm: Optional[Tensor], ignore_index: Optional[int],
    confmat: Tensor, scratch: Tensor, err: Tensor,
) -> None:  # fmt: skip
    """Binned mode: ``confmat[t, target, preds >= thr_t] += 1`` in place (reference :162-250).

    Ignored samples take no part in the sigmoid decision (:176-184, H2).
    """
    thr = _curve_thresholds_for(preds, thr_sorted, _dimensioned_compare(preds, target, ignore_index, 1, _LOOP_SWITCH))
    kernels.thresh_bin_update(preds, target, 1, True, thr, False, ignore_index, False, scratch, err)
    kernels.thresh_bin_finalize(scratch, 1, thr.numel(), perm, kernels.TB_CURVE, (confmat,))


def _flag_non_binary_targets(target: Tensor, ignore_index: Optional[int], err: Tensor) -> None:
    """Exact mode stores the targets as they come, no kernel looks at them during `update`.  With ``validate_args``
    the reference's {0, 1} check (:149-159, a ``torch.unique``) becomes one elementwise expression whose verdict is
    ORed into the device error word — read where the other data errors are read, no synchronisation here."""
    bad = (target != 0) & (target != 1)
    if ignore_index is not None:
        bad &= target != ignore_index
    err.bitwise_or_(bad.any().to(torch.int32) * nat.ERR_TARGET_BINARY)


def _binary_precision_recall_curve_format_exact(
    preds: Tensor, target: Tensor, ignore_index: Optional[int], err: Tensor
) -> Tuple[Tensor, Tensor]:
    """Exact mode: flattened, ignore-filtered, conditionally sigmoided scores to store (reference :176-186)."""
    preds = preds.flatten()
    target = target.flatten()
    if ignore_index is not None:
        preds, target = kernels.compact_not_ignored(preds, target, ignore_index)
    return kernels.sigmoid_normalize(preds, None, None, err), target


def _binary_precision_recall_curve_compute(
    state: Union[Tensor, Tuple[Tensor, Tensor]], thresholds: Optional[Tensor], pos_label: int = 1
) -> Tuple[Tensor, Tensor, Tensor]:
    """Reference: precision_recall_curve.py:253-283."""
    if isinstance(state, Tensor) and thresholds is not None:
        precision, recall = _curve_from_counts(state[:, 1, 1], state[:, 0, 1], state[:, 1, 0], (1,))
        return precision, recall, thresholds
    fps, tps, thresholds = _binary_clf_curve(state[0], state[1], pos_label=pos_label)
    return _exact_pr_from_curve(fps, tps, thresholds)


def _exact_pr_from_curve(fps: Tensor, tps: Tensor, thresholds: Tensor) -> Tuple[Tensor, Tensor, Tensor]:
    precision = tps / (tps + fps)
    recall = tps / tps[-1]
    precision = torch.cat([precision.flip(0), torch.ones(1, dtype=precision.dtype, device=precision.device)])
    recall = torch.cat([recall.flip(0), torch.zeros(1, dtype=recall.dtype, device=recall.device)])
    return precision, recall, thresholds.flip(0).detach().clone()


def _binary_state_functional(
    preds: Tensor, target: Tensor, thresholds: Optional[Tensor], ignore_index: Optional[int], validate_args: bool,
    where: str,
) -> Union[Tensor, Tuple[Tensor, Tensor]]:  # fmt: skip
    dev = preds.device
    err = torch.zeros(1, dtype=torch.int32, device=dev)
    if thresholds is None:
        if validate_args:
            _flag_non_binary_targets(target, ignore_index, err)
        state = _binary_precision_recall_curve_format_exact(preds, target, ignore_index, err)
    else:
        thr_sorted, perm = _threshold_plan(thresholds, dev)
        state = torch.zeros(len(thresholds), 2, 2, dtype=torch.long, device=dev)
        scratch = torch.zeros(kernels.thresh_bin_scratch_numel(1, len(thresholds)), dtype=torch.long, device=dev)
        _binary_precision_recall_curve_update_(preds, target, thr_sorted, perm, ignore_index, state, scratch, err)
    if validate_args:
        kernels.raise_on_error_word(err, where)
    return state


def binary_precision_recall_curve(
    preds: Tensor,
    target: Tensor,
    thresholds: Optional[Union[int, List[float], Tensor]] = None,
    ignore_index: Optional[int] = None,
    validate_args: bool = True,
) -> Tuple[Tensor, Tensor, Tensor]:
    """``(precision, recall, thresholds)`` of a binary task (reference :286-357)."""
    if validate_args:
        _binary_precision_recall_curve_arg_validation(thresholds, ignore_index)
        _binary_precision_recall_curve_tensor_validation(preds, target, ignore_index)
    thresholds = _adjust_threshold_arg(thresholds, preds.device)
    state = _binary_state_functional(preds, target, thresholds, ignore_index, validate_args,
                                     "binary_precision_recall_curve")
    return _binary_precision_recall_curve_compute(state, thresholds)


# ------------------------------------------------------------------------------------------ multilabel
def _multilabel_precision_recall_curve_arg_validation(
    num_labels: int,
    thresholds: Optional[Union[int, List[float], Tensor]] = None,
    ignore_index: Optional[int] = None,
) -> None:
    """Reference: precision_recall_curve.py:692-706 (delegates to the multiclass check, :360-377)."""
    if not isinstance(num_labels, int) or num_labels < 2:
        raise ValueError(f"Expected argument `num_classes` to be an integer larger than 1, but got {num_labels}")
    _binary_precision_recall_curve_arg_validation(thresholds, ignore_index)


def _multilabel_precision_recall_curve_tensor_validation(
    preds: Tensor, target: Tensor, num_labels: int, ignore_index: Optional[int] = None
) -> None:
    """Reference: precision_recall_curve.py:709-726."""
    _binary_precision_recall_curve_tensor_validation(preds, target, ignore_index)
    if preds.shape[1] != num_labels:
        raise ValueError(
            "Expected both `target.shape[1]` and `preds.shape[1]` to be equal to the number of labels"
            f" but got {preds.shape[1]} and expected {num_labels}"
        )


def _multilabel_precision_recall_curve_update_(
    preds: Tensor, target: Tensor, num_labels: int, thr_sorted: Tensor, perm: Optional[Tensor],
    ignore_index: Optional[int], confmat: Tensor, scratch: Tensor, err: Tensor,
) -> None:  # fmt: skip
    """Binned mode: ``confmat[t, label, target, preds >= thr_t] += 1`` in place (reference :739-793).

    The sigmoid decision looks at every element, ignored or not (:756-757); ignored elements are then
    dropped from all four cells (:760-766, :791).
    """
    kernels.thresh_bin_update(preds, target, num_labels, False, thr_sorted, False, ignore_index, True, scratch, err)
    kernels.thresh_bin_finalize(scratch, num_labels, thr_sorted.numel(), perm, kernels.TB_CURVE, (confmat,))


def _multilabel_precision_recall_curve_format_exact(
    preds: Tensor, target: Tensor, num_labels: int, err: Tensor
) -> Tuple[Tensor, Tensor]:
    """Exact mode: ``[N*..., L]`` views of the conditionally sigmoided scores and the targets (:752-757)."""
    preds = kernels.sigmoid_normalize(preds, None, None, err)
    preds = preds.transpose(0, 1).reshape(num_labels, -1).T
    target = target.transpose(0, 1).reshape(num_labels, -1).T
    return preds, target


def _multilabel_clf_curves(
    state: Tuple[Tensor, Tensor], num_labels: int, ignore_index: Optional[int]
) -> Tuple[List[Tensor], List[Tensor], List[Tensor]]:
    """Per-label ``(fps, tps, thresholds)``: the reference's Python loop over ``_binary_clf_curve``
    (:819-829, roc.py:345-355) as one batched sort + scan."""
    return kernels.clf_curves(state[0], state[1], ignore_index, 1)


def _multilabel_precision_recall_curve_compute(
    state: Union[Tensor, Tuple[Tensor, Tensor]],
    num_labels: int,
    thresholds: Optional[Tensor],
    ignore_index: Optional[int] = None,
) -> Union[Tuple[Tensor, Tensor, Tensor], Tuple[List[Tensor], List[Tensor], List[Tensor]]]:
    """Reference: precision_recall_curve.py:796-830."""
    if isinstance(state, Tensor) and thresholds is not None:
        precision, recall = _curve_from_counts(
            state[:, :, 1, 1], state[:, :, 0, 1], state[:, :, 1, 0], (1, num_labels)
        )
        return precision.T, recall.T, thresholds
    fps, tps, thr = _multilabel_clf_curves(state, num_labels, ignore_index)
    precision_list, recall_list, thres_list = [], [], []
    for i in range(num_labels):
        p, r, t = _exact_pr_from_curve(fps[i], tps[i], thr[i])
        precision_list.append(p)
        recall_list.append(r)
        thres_list.append(t)
    return precision_list, recall_list, thres_list


def _multilabel_state_functional(
    preds: Tensor, target: Tensor, num_labels: int, thresholds: Optional[Tensor], ignore_index: Optional[int],
    validate_args: bool, where: str,
) -> Union[Tensor, Tuple[Tensor, Tensor]]:  # fmt: skip
    dev = preds.device
    err = torch.zeros(1, dtype=torch.int32, device=dev)
    if thresholds is None:
        if validate_args:
            _flag_non_binary_targets(target, ignore_index, err)
        state = _multilabel_precision_recall_curve_format_exact(preds, target, num_labels, err)
    else:
        thr_sorted, perm = _threshold_plan(thresholds, dev)
        state = torch.zeros(len(thresholds), num_labels, 2, 2, dtype=torch.long, device=dev)
        scratch = torch.zeros(
            kernels.thresh_bin_scratch_numel(num_labels, len(thresholds)), dtype=torch.long, device=dev
        )
        _multilabel_precision_recall_curve_update_(
            preds, target, num_labels, thr_sorted, perm, ignore_index, state, scratch, err
        )
    if validate_args:
        kernels.raise_on_error_word(err, where)
    return state


def multilabel_precision_recall_curve(
    preds: Tensor,
    target: Tensor,
    num_labels: int,
    thresholds: Optional[Union[int, List[float], Tensor]] = None,
    ignore_index: Optional[int] = None,
    validate_args: bool = True,
) -> Union[Tuple[Tensor, Tensor, Tensor], Tuple[List[Tensor], List[Tensor], List[Tensor]]]:
    """Per-label ``(precision, recall, thresholds)`` (reference :833-931)."""
    if validate_args:
        _multilabel_precision_recall_curve_arg_validation(num_labels, thresholds, ignore_index)
        _multilabel_precision_recall_curve_tensor_validation(preds, target, num_labels, ignore_index)
    thresholds = _adjust_threshold_arg(thresholds, preds.device)
    state = _multilabel_state_functional(preds, target, num_labels, thresholds, ignore_index, validate_args,
                                         "multilabel_precision_recall_curve")
    return _multilabel_precision_recall_curve_compute(state, num_labels, thresholds, ignore_index)


# ------------------------------------------------------------------------------------------ multiclass (row f2)
_MC_VECTOR_SWITCH = 1_000_000  # reference :472: above this many comparisons it loops over thresholds (0-d compare)


def _multiclass_precision_recall_curve_arg_validation(
    num_classes: int,
    thresholds: Optional[Union[int, List[float], Tensor]] = None,
    ignore_index: Optional[int] = None,
    average: Optional[str] = None,
) -> None:
    """Reference: precision_recall_curve.py:360-377."""
    if not isinstance(num_classes, int) or num_classes < 2:
        raise ValueError(f"Expected argument `num_classes` to be an integer larger than 1, but got {num_classes}")
    if average not in (None, "micro", "macro"):
        raise ValueError(f"Expected argument `average` to be one of None, 'micro' or 'macro', but got {average}")
    _binary_precision_recall_curve_arg_validation(thresholds, ignore_index)


def _multiclass_precision_recall_curve_tensor_validation(
    preds: Tensor, target: Tensor, num_classes: int, ignore_index: Optional[int] = None
) -> None:
    """Shape / dtype checks of :380-420; the range check of the targets (the reference counts unique values, a
    host synchronisation) is done by the kernels through the error word."""
    if not preds.ndim == target.ndim + 1:
        raise ValueError(
            f"Expected `preds` to have one more dimension than `target` but got {preds.ndim} and {target.ndim}"
        )
    if target.is_floating_point():
        raise ValueError(
            f"Expected argument `target` to be an int or long tensor, but got tensor with dtype {target.dtype}"
        )
    if not preds.is_floating_point():
        raise ValueError(f"Expected `preds` to be a float tensor, but got {preds.dtype}")
    if preds.shape[1] != num_classes:
        raise ValueError(
            "Expected `preds.shape[1]` to be equal to the number of classes but"
            f" got {preds.shape[1]} and {num_classes}."
        )
    if preds.shape[0] != target.shape[0] or preds.shape[2:] != target.shape[1:]:
        raise ValueError(
            "Expected the shape of `preds` should be (N, C, ...) and the shape of `target` should be (N, ...)"
            f" but got {preds.shape} and {target.shape}"
        )


def _multiclass_rows(preds: Tensor, target: Tensor, num_classes: int) -> Tuple[Tensor, Tensor]:
    """``[M, C]`` scores and ``[M]`` class indices, extra dimensions folded into the rows (reference :439-440).
    For ``[N, C]`` inputs this is a view; inputs with extra dimensions are transposed by torch."""
    return preds.transpose(0, 1).reshape(num_classes, -1).T, target.flatten()


def _multiclass_precision_recall_curve_update_(
    preds: Tensor, target: Tensor, num_classes: int, thr_sorted: Tensor, perm: Optional[Tensor],
    ignore_index: Optional[int], average: Optional[str], confmat: Tensor, scratch: Tensor, err: Tensor,
) -> None:  # fmt: skip
    """Binned mode: ``confmat[t, class, class == target, softmax(preds)[class] >= thr_t] += 1`` in place
    (reference :423-527); ``average="micro"`` folds the classes into one binary ``[T, 2, 2]`` state.

    Rows with ``target == ignore_index`` are dropped and take no part in the softmax decision (:442-448).
    """
    rows, tgt = _multiclass_rows(preds, target, num_classes)
    micro = average == "micro"
    per_row = num_classes if micro else num_classes * num_classes  # the reference counts preds.numel() (* C)
    thr = _curve_thresholds_for(
        rows, thr_sorted,
        _dimensioned_compare(rows, tgt, ignore_index, per_row, _LOOP_SWITCH if micro else _MC_VECTOR_SWITCH),
    )
    labels = 1 if micro else num_classes
    if num_classes <= kernels.MC_CURVE_MAX_CLASSES:
        kernels.mc_curve_update(rows, tgt, thr, ignore_index, micro, scratch, err)
    else:
        # rows wider than the register-resident kernel: torch's own softmax (its wide-row kernel has another
        # reduction order) behind the device-side range decision, then the multilabel binning with one-hot targets
        probs = kernels.softmax_normalize(rows, tgt, ignore_index)
        keep = None if ignore_index is None else tgt != ignore_index
        safe = tgt if keep is None else torch.where(keep, tgt, torch.zeros_like(tgt))
        hot = torch.nn.functional.one_hot(safe.long().clamp(0, num_classes - 1), num_classes)
        if keep is not None:
            hot = torch.where(keep.unsqueeze(1), hot, torch.full_like(hot, ignore_index))
        if micro:
            kernels.thresh_bin_update(probs.flatten(), hot.flatten(), 1, True, thr, False, ignore_index, False, scratch, err)
        else:
            kernels.thresh_bin_update(probs, hot, num_classes, False, thr, False, ignore_index, False, scratch, err)
    kernels.thresh_bin_finalize(scratch, labels, thr.numel(), perm, kernels.TB_CURVE, (confmat,))


def _multiclass_precision_recall_curve_format_exact(
    preds: Tensor, target: Tensor, num_classes: int, ignore_index: Optional[int], average: Optional[str],
    validate_args: bool = False,
) -> Tuple[Tensor, Tensor]:  # fmt: skip
    """Exact mode: the ``[M', C]`` (conditionally softmaxed) scores and ``[M']`` targets to store; ``micro``
    stores them flattened with one-hot targets (reference :439-456)."""
    rows, tgt = _multiclass_rows(preds, target, num_classes)
    if ignore_index is not None:
        keep = tgt != ignore_index  # boolean-mask indexing: data-dependent size, synchronises like the reference
        rows, tgt = rows[keep], tgt[keep]
    if validate_args and bool(((tgt < 0) | (tgt >= num_classes)).any()):
        raise RuntimeError(
            "Detected more unique values in `target` than `num_classes`: `target` contains values outside"
            f" [0, {num_classes}) that are not `ignore_index`."
        )
    probs = kernels.softmax_normalize(rows)
    if average == "micro":
        return probs.flatten(), torch.nn.functional.one_hot(tgt.long(), num_classes=num_classes).flatten()
    return probs, tgt


def _multiclass_state_functional(
    preds: Tensor, target: Tensor, num_classes: int, thresholds: Optional[Tensor], ignore_index: Optional[int],
    average: Optional[str], validate_args: bool, where: str,
) -> Union[Tensor, Tuple[Tensor, Tensor]]:  # fmt: skip
    dev = preds.device
    if thresholds is None:
        return _multiclass_precision_recall_curve_format_exact(
            preds, target, num_classes, ignore_index, average, validate_args
        )
    err = torch.zeros(1, dtype=torch.int32, device=dev)
    thr_sorted, perm = _threshold_plan(thresholds, dev)
    labels = 1 if average == "micro" else num_classes
    shape = (len(thresholds), 2, 2) if average == "micro" else (len(thresholds), num_classes, 2, 2)
    state = torch.zeros(*shape, dtype=torch.long, device=dev)
    scratch = torch.zeros(kernels.thresh_bin_scratch_numel(labels, len(thresholds)), dtype=torch.long, device=dev)
    _multiclass_precision_recall_curve_update_(
        preds, target, num_classes, thr_sorted, perm, ignore_index, average, state, scratch, err
    )
    if validate_args:
        kernels.raise_on_error_word(err, where)
    return state


def _macro_curve(
    xs: Sequence[Tensor], ys: Sequence[Tensor], thresholds: Tensor, descending_thresholds: bool
) -> Tuple[Tensor, Tensor, Tensor]:
    """``average="macro"``: all classes' x values merged and sorted, y = mean of the per-class curves interpolated
    there (reference :557-571, roc.py:189-200)."""
    merged = torch.cat(list(xs), 0).sort().values
    mean_y = torch.zeros_like(merged)
    for x, y in zip(xs, ys):
        mean_y += interp(merged, x, y)
    mean_y /= len(xs)
    return merged, mean_y, thresholds.sort(descending=descending_thresholds).values


def _multiclass_precision_recall_curve_compute(
    state: Union[Tensor, Tuple[Tensor, Tensor]], num_classes: int, thresholds: Optional[Tensor],
    average: Optional[str] = None,
) -> Union[Tuple[Tensor, Tensor, Tensor], Tuple[List[Tensor], List[Tensor], List[Tensor]]]:  # fmt: skip
    """Reference: precision_recall_curve.py:530-575."""
    if average == "micro":
        return _binary_precision_recall_curve_compute(state, thresholds)
    binned = isinstance(state, Tensor) and thresholds is not None
    if binned:
        precision, recall = _curve_from_counts(
            state[:, :, 1, 1], state[:, :, 0, 1], state[:, :, 1, 0], (1, num_classes)
        )
        precision, recall = precision.T, recall.T
        if average == "macro":
            return _macro_curve(list(precision), list(recall), thresholds.repeat(num_classes), False)
        return precision, recall, thresholds
    fps, tps, thr = kernels.clf_curves(state[0], state[1], None, class_targets=True)
    curves = [_exact_pr_from_curve(fps[i], tps[i], thr[i]) for i in range(num_classes)]
    precision_list, recall_list, thres_list = ([c[k] for c in curves] for k in range(3))
    if average == "macro":
        return _macro_curve(precision_list, recall_list, torch.cat(thres_list, 0), False)
    return precision_list, recall_list, thres_list


def multiclass_precision_recall_curve(
    preds: Tensor,
    target: Tensor,
    num_classes: int,
    thresholds: Optional[Union[int, List[float], Tensor]] = None,
    average: Optional[Literal["micro", "macro"]] = None,
    ignore_index: Optional[int] = None,
    validate_args: bool = True,
) -> Union[Tuple[Tensor, Tensor, Tensor], Tuple[List[Tensor], List[Tensor], List[Tensor]]]:
    """One-vs-rest ``(precision, recall, thresholds)`` per class (reference :578-689)."""
    if validate_args:
        _multiclass_precision_recall_curve_arg_validation(num_classes, thresholds, ignore_index, average)
        _multiclass_precision_recall_curve_tensor_validation(preds, target, num_classes, ignore_index)
    thresholds = _adjust_threshold_arg(thresholds, preds.device)
    state = _multiclass_state_functional(preds, target, num_classes, thresholds, ignore_index, average,
                                         validate_args, "multiclass_precision_recall_curve")
    return _multiclass_precision_recall_curve_compute(state, num_classes, thresholds, average)


def precision_recall_curve(
    preds: Tensor,
    target: Tensor,
    task: Literal["binary", "multiclass", "multilabel"],
    thresholds: Optional[Union[int, List[float], Tensor]] = None,
    num_classes: Optional[int] = None,
    num_labels: Optional[int] = None,
    ignore_index: Optional[int] = None,
    validate_args: bool = True,
) -> Union[Tuple[Tensor, Tensor, Tensor], Tuple[List[Tensor], List[Tensor], List[Tensor]]]:
    """Task dispatcher (reference :934-1001); the multiclass (softmax) curves are a later row of SURVEY §8."""
    task = ClassificationTask.from_str(task)
    if task == ClassificationTask.BINARY:
        return binary_precision_recall_curve(preds, target, thresholds, ignore_index, validate_args)
    if task == ClassificationTask.MULTICLASS:
        if not isinstance(num_classes, int):
            raise ValueError(f"`num_classes` is expected to be `int` but `{type(num_classes)} was passed.`")
        return multiclass_precision_recall_curve(
            preds, target, num_classes, thresholds, None, ignore_index, validate_args
        )
    if task == ClassificationTask.MULTILABEL:
        if not isinstance(num_labels, int):
            raise ValueError(f"`num_labels` is expected to be `int` but `{type(num_labels)} was passed.`")
        return multilabel_precision_recall_curve(preds, target, num_labels, thresholds, ignore_index, validate_args)
    raise ValueError(f"Task {task} not supported.")
