"""Average precision (reference: functional/classification/average_precision.py)."""
from typing import List, Optional, Tuple, Union

import torch
from torch import Tensor
from typing_extensions import Literal

from torchmetrics_forked_b200.functional.classification.auroc import _flatten_valid
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
from torchmetrics_forked_b200.utilities.compute import _safe_divide
from torchmetrics_forked_b200.utilities.enums import ClassificationTask
from torchmetrics_forked_b200.utilities.prints import rank_zero_warn


def _reduce_average_precision(
    precision: Union[Tensor, List[Tensor]],
    recall: Union[Tensor, List[Tensor]],
    average: Optional[str] = "macro",
    weights: Optional[Tensor] = None,
) -> Tensor:
    """Step-wise area ``-sum((r[1:] - r[:-1]) * p[:-1])`` per label, nan-aware averaging (reference :43-67)."""
    if isinstance(precision, Tensor) and isinstance(recall, Tensor):
        res = -torch.sum((recall[:, 1:] - recall[:, :-1]) * precision[:, :-1], 1)
    else:
        res = torch.stack([-torch.sum((r[1:] - r[:-1]) * p[:-1]) for p, r in zip(precision, recall)])
    if average is None or average == "none":
        return res
    if torch.isnan(res).any():
        rank_zero_warn(
            f"Average precision score for one or more classes was `nan`. Ignoring these classes in {average}-average",
            UserWarning,
        )
    keep = ~torch.isnan(res)
    if average == "macro":
        return res[keep].mean()
    if average == "weighted" and weights is not None:
        weights = _safe_divide(weights[keep], weights[keep].sum())
        return (res[keep] * weights).sum()
    raise ValueError("Received an incompatible combinations of inputs to make reduction.")


def _exact_ap_from_scores(scores: Tensor) -> Tensor:
    """Per-label average precision (float32) from :func:`kernels.clf_curve_scores`: ``sum (r_b - r_{b-1}) p_b``
    with ``r = tps / tps_last`` — the reference's flipped step sum (:43-53, :70-75) — NaN without positives."""
    _, ap_sum, _, pos = scores.unbind(1)
    return (ap_sum / pos).to(torch.float32)


def _binary_average_precision_compute(
    state: Union[Tensor, Tuple[Tensor, Tensor]], thresholds: Optional[Tensor]
) -> Tensor:
    if not (isinstance(state, Tensor) and thresholds is not None):
        from torchmetrics_forked_b200 import kernels

        scores = kernels.clf_curve_scores(kernels.as_columns(state[0]), kernels.as_columns(state[1]), None, 1)
        return _exact_ap_from_scores(scores)[0]
    precision, recall, _ = _binary_precision_recall_curve_compute(state, thresholds)
    return -torch.sum((recall[1:] - recall[:-1]) * precision[:-1])


def binary_average_precision(
    preds: Tensor,
    target: Tensor,
    thresholds: Optional[Union[int, List[float], Tensor]] = None,
    ignore_index: Optional[int] = None,
    validate_args: bool = True,
) -> Tensor:
    """Reference: average_precision.py:78-149."""
    if validate_args:
        _binary_precision_recall_curve_arg_validation(thresholds, ignore_index)
        _binary_precision_recall_curve_tensor_validation(preds, target, ignore_index)
    thresholds = _adjust_threshold_arg(thresholds, preds.device)
    state = _binary_state_functional(preds, target, thresholds, ignore_index, validate_args,
                                     "binary_average_precision")
    return _binary_average_precision_compute(state, thresholds)


def _multiclass_average_precision_arg_validation(
    num_classes: int,
    average: Optional[str] = "macro",
    thresholds: Optional[Union[int, List[float], Tensor]] = None,
    ignore_index: Optional[int] = None,
) -> None:
    _multiclass_precision_recall_curve_arg_validation(num_classes, thresholds, ignore_index)
    allowed_average = ("macro", "weighted", "none", None)
    if average not in allowed_average:
        raise ValueError(f"Expected argument `average` to be one of {allowed_average} but got {average}")


def _multiclass_average_precision_compute(
    state: Union[Tensor, Tuple[Tensor, Tensor]],
    num_classes: int,
    average: Optional[str] = "macro",
    thresholds: Optional[Tensor] = None,
) -> Tensor:
    """One-vs-rest average precision per class, then averaging (reference average_precision.py:164-176)."""
    if not (isinstance(state, Tensor) and thresholds is not None):
        from torchmetrics_forked_b200 import kernels
        from torchmetrics_forked_b200.functional.classification.auroc import _reduce_scores

        scores = kernels.clf_curve_scores(state[0], state[1], None, class_targets=True)
        return _reduce_scores(_exact_ap_from_scores(scores), average, scores[:, 3].float())
    precision, recall, _ = _multiclass_precision_recall_curve_compute(state, num_classes, thresholds)
    return _reduce_average_precision(precision, recall, average, weights=state[0][:, 1, :].sum(-1))


def multiclass_average_precision(
    preds: Tensor,
    target: Tensor,
    num_classes: int,
    average: Optional[Literal["macro", "weighted", "none"]] = "macro",
    thresholds: Optional[Union[int, List[float], Tensor]] = None,
    ignore_index: Optional[int] = None,
    validate_args: bool = True,
) -> Tensor:
    """Reference: average_precision.py:179-269."""
    if validate_args:
        _multiclass_average_precision_arg_validation(num_classes, average, thresholds, ignore_index)
        _multiclass_precision_recall_curve_tensor_validation(preds, target, num_classes, ignore_index)
    thresholds = _adjust_threshold_arg(thresholds, preds.device)
    state = _multiclass_state_functional(preds, target, num_classes, thresholds, ignore_index, None, validate_args,
                                         "multiclass_average_precision")
    return _multiclass_average_precision_compute(state, num_classes, average, thresholds)


def _multilabel_average_precision_arg_validation(
    num_labels: int,
    average: Optional[str],
    thresholds: Optional[Union[int, List[float], Tensor]] = None,
    ignore_index: Optional[int] = None,
) -> None:
    _multilabel_precision_recall_curve_arg_validation(num_labels, thresholds, ignore_index)
    allowed_average = ("micro", "macro", "weighted", "none", None)
    if average not in allowed_average:
        raise ValueError(f"Expected argument `average` to be one of {allowed_average} but got {average}")


def _multilabel_average_precision_compute(
    state: Union[Tensor, Tuple[Tensor, Tensor]],
    num_labels: int,
    average: Optional[str],
    thresholds: Optional[Tensor],
    ignore_index: Optional[int] = None,
) -> Tensor:
    """Reference: average_precision.py:284-309."""
    if average == "micro":
        if isinstance(state, Tensor) and thresholds is not None:
            state = state.sum(1)
        else:
            state = _flatten_valid(state, ignore_index)
        return _binary_average_precision_compute(state, thresholds)
    if not (isinstance(state, Tensor) and thresholds is not None):
        from torchmetrics_forked_b200 import kernels
        from torchmetrics_forked_b200.functional.classification.auroc import _reduce_scores

        scores = kernels.clf_curve_scores(state[0], state[1], ignore_index, 1)
        # weights: the reference counts (target == 1) per label (:308); column 3 is that count already
        return _reduce_scores(_exact_ap_from_scores(scores), average, scores[:, 3].float())
    precision, recall, _ = _multilabel_precision_recall_curve_compute(state, num_labels, thresholds, ignore_index)
    return _reduce_average_precision(precision, recall, average, weights=state[0][:, 1, :].sum(-1))


def multilabel_average_precision(
    preds: Tensor,
    target: Tensor,
    num_labels: int,
    average: Optional[Literal["micro", "macro", "weighted", "none"]] = "macro",
    thresholds: Optional[Union[int, List[float], Tensor]] = None,
    ignore_index: Optional[int] = None,
    validate_args: bool = True,
) -> Tensor:
    """Reference: average_precision.py:312-406."""
    if validate_args:
        _multilabel_average_precision_arg_validation(num_labels, average, thresholds, ignore_index)
        _multilabel_precision_recall_curve_tensor_validation(preds, target, num_labels, ignore_index)
    thresholds = _adjust_threshold_arg(thresholds, preds.device)
    state = _multilabel_state_functional(preds, target, num_labels, thresholds, ignore_index, validate_args,
                                         "multilabel_average_precision")
    return _multilabel_average_precision_compute(state, num_labels, average, thresholds, ignore_index)


def average_precision(
    preds: Tensor,
    target: Tensor,
    task: Literal["binary", "multiclass", "multilabel"],
    thresholds: Optional[Union[int, List[float], Tensor]] = None,
    num_classes: Optional[int] = None,
    num_labels: Optional[int] = None,
    average: Optional[Literal["macro", "weighted", "none"]] = "macro",
    ignore_index: Optional[int] = None,
    validate_args: bool = True,
) -> Optional[Tensor]:
    """Task dispatcher (reference :409-467)."""
    task = ClassificationTask.from_str(task)
    if task == ClassificationTask.BINARY:
        return binary_average_precision(preds, target, thresholds, ignore_index, validate_args)
    if task == ClassificationTask.MULTICLASS:
        if not isinstance(num_classes, int):
            raise ValueError(f"`num_classes` is expected to be `int` but `{type(num_classes)} was passed.`")
        return multiclass_average_precision(preds, target, num_classes, average, thresholds, ignore_index, validate_args)
    if task == ClassificationTask.MULTILABEL:
        if not isinstance(num_labels, int):
            raise ValueError(f"`num_labels` is expected to be `int` but `{type(num_labels)} was passed.`")
        return multilabel_average_precision(preds, target, num_labels, average, thresholds, ignore_index, validate_args)
    return None
