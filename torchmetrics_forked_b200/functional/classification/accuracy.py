"""Accuracy from the fused stat-score states (reference: functional/classification/accuracy.py).

The update is the stat-scores kernel; only the reduction below is accuracy specific.  It is evaluated
with the same fp32 torch expressions as the reference (:37-86) on a handful of numbers.
"""
from typing import Optional

from torch import Tensor
from typing_extensions import Literal

from torchmetrics_forked_b200.functional.classification.stat_scores import (
    _binary_stat_scores_arg_validation,
    _binary_stat_scores_functional_state,
    _binary_stat_scores_tensor_validation,
    _multiclass_stat_scores_arg_validation,
    _multiclass_stat_scores_functional_state,
    _multiclass_stat_scores_tensor_validation,
    _multilabel_stat_scores_arg_validation,
    _multilabel_stat_scores_functional_state,
    _multilabel_stat_scores_tensor_validation,
    _require_global,
)
from torchmetrics_forked_b200.utilities.compute import _adjust_weights_safe_divide, _safe_divide
from torchmetrics_forked_b200.utilities.enums import ClassificationTask


def _accuracy_reduce(
    tp: Tensor,
    fp: Tensor,
    tn: Tensor,
    fn: Tensor,
    average: Optional[str],
    multidim_average: str = "global",
    multilabel: bool = False,
) -> Tensor:
    """``(tp + tn) / all`` (binary, multilabel) or ``tp / (tp + fn)`` (multiclass), then class averaging."""
    if average == "binary":
        return _safe_divide(tp + tn, tp + tn + fp + fn)
    if average == "micro":
        dim = 0 if multidim_average == "global" else 1
        tp, fn = tp.sum(dim=dim), fn.sum(dim=dim)
        if multilabel:
            fp, tn = fp.sum(dim=dim), tn.sum(dim=dim)
            return _safe_divide(tp + tn, tp + tn + fp + fn)
        return _safe_divide(tp, tp + fn)
    score = _safe_divide(tp + tn, tp + tn + fp + fn) if multilabel else _safe_divide(tp, tp + fn)
    return _adjust_weights_safe_divide(score, average, multilabel, tp, fp, fn)


def binary_accuracy(
    preds: Tensor,
    target: Tensor,
    threshold: float = 0.5,
    multidim_average: Literal["global", "samplewise"] = "global",
    ignore_index: Optional[int] = None,
    validate_args: bool = True,
) -> Tensor:
    """Reference: accuracy.py:89-161."""
    if validate_args:
        _binary_stat_scores_arg_validation(threshold, multidim_average, ignore_index)
        _binary_stat_scores_tensor_validation(preds, target, multidim_average, ignore_index)
    _require_global(multidim_average)
    tp, fp, tn, fn = _binary_stat_scores_functional_state(
        preds, target, threshold, ignore_index, validate_args, "binary_accuracy",
        multidim_average=multidim_average,
    )
    return _accuracy_reduce(tp, fp, tn, fn, average="binary", multidim_average=multidim_average)


def multiclass_accuracy(
    preds: Tensor,
    target: Tensor,
    num_classes: int,
    average: Optional[Literal["micro", "macro", "weighted", "none"]] = "macro",
    top_k: int = 1,
    multidim_average: Literal["global", "samplewise"] = "global",
    ignore_index: Optional[int] = None,
    validate_args: bool = True,
) -> Tensor:
    """Reference: accuracy.py:164-269."""
    if validate_args:
        _multiclass_stat_scores_arg_validation(num_classes, top_k, average, multidim_average, ignore_index)
        _multiclass_stat_scores_tensor_validation(preds, target, num_classes, multidim_average, ignore_index)
    _require_global(multidim_average, top_k)
    tp, fp, tn, fn = _multiclass_stat_scores_functional_state(
        preds, target, num_classes, average, ignore_index, validate_args, "multiclass_accuracy", top_k,
        multidim_average=multidim_average,
    )
    return _accuracy_reduce(tp, fp, tn, fn, average=average, multidim_average=multidim_average)


def multilabel_accuracy(
    preds: Tensor,
    target: Tensor,
    num_labels: int,
    threshold: float = 0.5,
    average: Optional[Literal["micro", "macro", "weighted", "none"]] = "macro",
    multidim_average: Literal["global", "samplewise"] = "global",
    ignore_index: Optional[int] = None,
    validate_args: bool = True,
) -> Tensor:
    """Reference: accuracy.py:272-371."""
    if validate_args:
        _multilabel_stat_scores_arg_validation(num_labels, threshold, average, multidim_average, ignore_index)
        _multilabel_stat_scores_tensor_validation(preds, target, num_labels, multidim_average, ignore_index)
    _require_global(multidim_average)
    tp, fp, tn, fn = _multilabel_stat_scores_functional_state(
        preds, target, num_labels, threshold, ignore_index, validate_args, "multilabel_accuracy",
        multidim_average=multidim_average,
    )
    return _accuracy_reduce(tp, fp, tn, fn, average=average, multidim_average=multidim_average, multilabel=True)


def accuracy(
    preds: Tensor,
    target: Tensor,
    task: Literal["binary", "multiclass", "multilabel"],
    threshold: float = 0.5,
    num_classes: Optional[int] = None,
    num_labels: Optional[int] = None,
    average: Optional[Literal["micro", "macro", "weighted", "none"]] = "micro",
    multidim_average: Literal["global", "samplewise"] = "global",
    top_k: Optional[int] = 1,
    ignore_index: Optional[int] = None,
    validate_args: bool = True,
) -> Tensor:
    """Task dispatcher (reference :374-436)."""
    task = ClassificationTask.from_str(task)
    if task == ClassificationTask.BINARY:
        return binary_accuracy(preds, target, threshold, multidim_average, ignore_index, validate_args)
    if task == ClassificationTask.MULTICLASS:
        if not isinstance(num_classes, int):
            raise ValueError(
                f"Optional arg `num_classes` must be type `int` when task is {task}. Got {type(num_classes)}"
            )
        if not isinstance(top_k, int):
            raise ValueError(f"Optional arg `top_k` must be type `int` when task is {task}. Got {type(top_k)}")
        return multiclass_accuracy(
            preds, target, num_classes, average, top_k, multidim_average, ignore_index, validate_args
        )
    if task == ClassificationTask.MULTILABEL:
        if not isinstance(num_labels, int):
            raise ValueError(
                f"Optional arg `num_labels` must be type `int` when task is {task}. Got {type(num_labels)}"
            )
        return multilabel_accuracy(
            preds, target, num_labels, threshold, average, multidim_average, ignore_index, validate_args
        )
    raise ValueError(f"Not handled value: {task}")
