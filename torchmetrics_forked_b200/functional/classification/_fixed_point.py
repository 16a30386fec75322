"""Operating-point metrics on top of the curve states: recall at fixed precision, precision
at fixed recall, specificity at fixed sensitivity.

The state and its update are the precision-recall-curve metric's (K2 binned / K3 exact); `compute` builds the
curve and picks one point of it.  Reference: functional/classification/recall_fixed_precision.py,
precision_fixed_recall.py, specificity_sensitivity.py.  The reference selects the point with Python-level loops
over the curve (``max`` over zipped 0-d tensors, repeated ``torch.where``); an exact-mode curve has up to N points,
so the selection here is a handful of masked reductions on the device with the same ordering rules.
"""
from typing import Callable, List, Optional, Sequence, Tuple, Union

import torch
from torch import Tensor
from typing_extensions import Literal

from torchmetrics_forked_b200.functional.classification.precision_recall_curve import (
    _adjust_threshold_arg,
    _binary_precision_recall_curve_arg_validation,
    _binary_precision_recall_curve_compute,
    _binary_precision_recall_curve_tensor_validation,
    _binary_state_functional,
    _multiclass_precision_recall_curve_arg_validation,
    _multiclass_precision_recall_curve_compute,
    _multiclass_precision_recall_curve_tensor_validation,
    _multiclass_state_functional,
    _multilabel_precision_recall_curve_arg_validation,
    _multilabel_precision_recall_curve_compute,
    _multilabel_precision_recall_curve_tensor_validation,
    _multilabel_state_functional,
)
from torchmetrics_forked_b200.functional.classification.roc import (
    _binary_roc_compute,
    _multiclass_roc_compute,
    _multilabel_roc_compute,
)
from torchmetrics_forked_b200.utilities.enums import ClassificationTask

State = Union[Tensor, Tuple[Tensor, Tensor]]
Selector = Callable[[Tensor, Tensor, Tensor, float], Tuple[Tensor, Tensor]]


def _lexicographic_best(keys: Sequence[Tensor], keep: Tensor) -> Optional[Tensor]:
    """Index of the lexicographically largest ``(keys[0][i], keys[1][i], ...)`` among ``keep``; ``None`` if
    nothing is kept.  Equal tuples resolve to the first index."""
    if not bool(keep.any()):
        return None
    for k in keys:
        best = k[keep].max()
        keep = keep & (k == best)
    return torch.nonzero(keep)[0, 0]


def _no_point(value_like: Tensor, thresholds: Tensor) -> Tuple[Tensor, Tensor]:
    """The reference's answer when no curve point qualifies (or the best value is 0): ``(0, 1e6)``."""
    return (
        torch.tensor(0.0, device=value_like.device, dtype=value_like.dtype),
        torch.tensor(1e6, device=thresholds.device, dtype=thresholds.dtype),
    )


def _recall_at_precision(
    precision: Tensor, recall: Tensor, thresholds: Tensor, min_precision: float
) -> Tuple[Tensor, Tensor]:
    """Highest recall among the points with ``precision >= min_precision``; ties by precision, then by threshold
    (reference recall_fixed_precision.py:40-76).  The curve's closing (1, 0) point has no threshold and is left
    out, as in the reference's ``zip``."""
    n = min(precision.shape[0], recall.shape[0], thresholds.shape[0])
    precision, recall, thresholds = precision[:n], recall[:n], thresholds[:n]
    i = _lexicographic_best((recall, precision, thresholds), precision >= min_precision)
    if i is None or recall[i] == 0.0:
        return _no_point(recall, thresholds) if i is None else (recall[i], _no_point(recall, thresholds)[1])
    return recall[i], thresholds[i]


def _precision_at_recall(
    precision: Tensor, recall: Tensor, thresholds: Tensor, min_recall: float
) -> Tuple[Tensor, Tensor]:
    """Highest precision among the points with ``recall >= min_recall``; ties by recall, then by threshold
    (reference precision_fixed_recall.py:42-60)."""
    n = min(precision.shape[0], recall.shape[0], thresholds.shape[0])
    precision, recall, thresholds = precision[:n], recall[:n], thresholds[:n]
    i = _lexicographic_best((precision, recall, thresholds), recall >= min_recall)
    if i is None or precision[i] == 0.0:
        return _no_point(precision, thresholds) if i is None else (precision[i], _no_point(precision, thresholds)[1])
    return precision[i], thresholds[i]


def _specificity_at_sensitivity(
    specificity: Tensor, sensitivity: Tensor, thresholds: Tensor, min_sensitivity: float
) -> Tuple[Tensor, Tensor]:
    """Highest specificity among the points with ``sensitivity >= min_sensitivity``, first such point on ties
    (reference specificity_sensitivity.py:47-70)."""
    i = _lexicographic_best((specificity,), sensitivity >= min_sensitivity)
    if i is None:
        return _no_point(specificity, thresholds)
    return specificity[i], thresholds[i]


def _check_level(name: str, value: float) -> None:
    # the reference's test (`not isinstance(float) and not (0 <= v <= 1)`) only fires for non-floats out of range;
    # kept as is so that the same arguments are accepted
    if not isinstance(value, float) and not (0 <= value <= 1):
        raise ValueError(f"Expected argument `{name}` to be an float in the [0,1] range, but got {value}")


def _binary_pr_point(state: State, thresholds: Optional[Tensor], level: float, select: Selector) -> Tuple[Tensor, Tensor]:
    precision, recall, thr = _binary_precision_recall_curve_compute(state, thresholds)
    return select(precision, recall, thr, level)


def _binary_roc_point(state: State, thresholds: Optional[Tensor], level: float) -> Tuple[Tensor, Tensor]:
    fpr, sensitivity, thr = _binary_roc_compute(state, thresholds)
    return _specificity_at_sensitivity(1 - fpr, sensitivity, thr, level)


def _stack_points(points: List[Tuple[Tensor, Tensor]]) -> Tuple[Tensor, Tensor]:
    return torch.stack([p[0] for p in points]), torch.stack([p[1] for p in points])


def _multilabel_pr_point(
    state: State, num_labels: int, thresholds: Optional[Tensor], ignore_index: Optional[int], level: float,
    select: Selector,
) -> Tuple[Tensor, Tensor]:  # fmt: skip
    precision, recall, thr = _multilabel_precision_recall_curve_compute(state, num_labels, thresholds, ignore_index)
    shared = isinstance(state, Tensor)  # binned: one threshold vector for all labels
    return _stack_points(
        [select(precision[i], recall[i], thr if shared else thr[i], level) for i in range(num_labels)]
    )


def _multilabel_roc_point(
    state: State, num_labels: int, thresholds: Optional[Tensor], ignore_index: Optional[int], level: float
) -> Tuple[Tensor, Tensor]:
    fpr, sensitivity, thr = _multilabel_roc_compute(state, num_labels, thresholds, ignore_index)
    shared = isinstance(state, Tensor)
    return _stack_points(
        [
            _specificity_at_sensitivity(1 - fpr[i], sensitivity[i], thr if shared else thr[i], level)
            for i in range(num_labels)
        ]
    )


def _multiclass_pr_point(
    state: State, num_classes: int, thresholds: Optional[Tensor], level: float, select: Selector
) -> Tuple[Tensor, Tensor]:
    precision, recall, thr = _multiclass_precision_recall_curve_compute(state, num_classes, thresholds)
    shared = isinstance(state, Tensor)
    return _stack_points(
        [select(precision[i], recall[i], thr if shared else thr[i], level) for i in range(num_classes)]
    )


def _multiclass_roc_point(
    state: State, num_classes: int, thresholds: Optional[Tensor], level: float
) -> Tuple[Tensor, Tensor]:
    fpr, sensitivity, thr = _multiclass_roc_compute(state, num_classes, thresholds)
    shared = isinstance(state, Tensor)
    return _stack_points(
        [
            _specificity_at_sensitivity(1 - fpr[i], sensitivity[i], thr if shared else thr[i], level)
            for i in range(num_classes)
        ]
    )


def fixed_point_family(stem: str, level_name: str, select: Optional[Selector]):
    """``(binary_<stem>, multiclass_<stem>, multilabel_<stem>, <stem>)``; ``select=None`` means the ROC based
    specificity metric."""

    def binary(
        preds: Tensor,
        target: Tensor,
        level: float,
        thresholds: Optional[Union[int, List[float], Tensor]] = None,
        ignore_index: Optional[int] = None,
        validate_args: bool = True,
    ) -> Tuple[Tensor, Tensor]:
        if validate_args:
            _binary_precision_recall_curve_arg_validation(thresholds, ignore_index)
            _check_level(level_name, level)
            _binary_precision_recall_curve_tensor_validation(preds, target, ignore_index)
        thresholds = _adjust_threshold_arg(thresholds, preds.device)
        state = _binary_state_functional(preds, target, thresholds, ignore_index, validate_args, f"binary_{stem}")
        if select is None:
            return _binary_roc_point(state, thresholds, level)
        return _binary_pr_point(state, thresholds, level, select)

    def multiclass(
        preds: Tensor,
        target: Tensor,
        num_classes: int,
        level: float,
        thresholds: Optional[Union[int, List[float], Tensor]] = None,
        ignore_index: Optional[int] = None,
        validate_args: bool = True,
    ) -> Tuple[Tensor, Tensor]:
        if validate_args:
            _multiclass_precision_recall_curve_arg_validation(num_classes, thresholds, ignore_index)
            _check_level(level_name, level)
            _multiclass_precision_recall_curve_tensor_validation(preds, target, num_classes, ignore_index)
        thresholds = _adjust_threshold_arg(thresholds, preds.device)
        state = _multiclass_state_functional(
            preds, target, num_classes, thresholds, ignore_index, None, validate_args, f"multiclass_{stem}"
        )
        if select is None:
            return _multiclass_roc_point(state, num_classes, thresholds, level)
        return _multiclass_pr_point(state, num_classes, thresholds, level, select)

    def multilabel(
        preds: Tensor,
        target: Tensor,
        num_labels: int,
        level: float,
        thresholds: Optional[Union[int, List[float], Tensor]] = None,
        ignore_index: Optional[int] = None,
        validate_args: bool = True,
    ) -> Tuple[Tensor, Tensor]:
        if validate_args:
            _multilabel_precision_recall_curve_arg_validation(num_labels, thresholds, ignore_index)
            _check_level(level_name, level)
            _multilabel_precision_recall_curve_tensor_validation(preds, target, num_labels, ignore_index)
        thresholds = _adjust_threshold_arg(thresholds, preds.device)
        state = _multilabel_state_functional(
            preds, target, num_labels, thresholds, ignore_index, validate_args, f"multilabel_{stem}"
        )
        if select is None:
            return _multilabel_roc_point(state, num_labels, thresholds, ignore_index, level)
        return _multilabel_pr_point(state, num_labels, thresholds, ignore_index, level, select)

    def by_task(
        preds: Tensor,
        target: Tensor,
        task: Literal["binary", "multiclass", "multilabel"],
        level: float,
        thresholds: Optional[Union[int, List[float], Tensor]] = None,
        num_classes: Optional[int] = None,
        num_labels: Optional[int] = None,
        ignore_index: Optional[int] = None,
        validate_args: bool = True,
    ) -> Optional[Tuple[Tensor, Tensor]]:
        task = ClassificationTask.from_str(task)
        if task == ClassificationTask.BINARY:
            return binary(preds, target, level, thresholds, ignore_index, validate_args)
        if task == ClassificationTask.MULTICLASS:
            if not isinstance(num_classes, int):
                raise ValueError(f"`num_classes` is expected to be `int` but `{type(num_classes)} was passed.`")
            return multiclass(preds, target, num_classes, level, thresholds, ignore_index, validate_args)
        if task == ClassificationTask.MULTILABEL:
            if not isinstance(num_labels, int):
                raise ValueError(f"`num_labels` is expected to be `int` but `{type(num_labels)} was passed.`")
            return multilabel(preds, target, num_labels, level, thresholds, ignore_index, validate_args)
        return None

    # the reference names the level argument after the metric (min_precision / min_recall / min_sensitivity)
    import functools
    import inspect

    def rename(fn: Callable, name: str) -> Callable:
        sig = inspect.signature(fn)
        params = [p.replace(name=level_name) if p.name == "level" else p for p in sig.parameters.values()]

        @functools.wraps(fn)
        def wrapper(*args, **kwargs):
            if level_name in kwargs:
                kwargs["level"] = kwargs.pop(level_name)
            return fn(*args, **kwargs)

        wrapper.__signature__ = sig.replace(parameters=params)
        wrapper.__name__ = wrapper.__qualname__ = name
        wrapper.__doc__ = f"``{name}``: curve state (K2 / K3 kernels), then one operating point of the curve."
        return wrapper

    return (
        rename(binary, f"binary_{stem}"),
        rename(multiclass, f"multiclass_{stem}"),
        rename(multilabel, f"multilabel_{stem}"),
        rename(by_task, stem),
    )
