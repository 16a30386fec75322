"""ROC curves from the binned state or from the exact (sort + scan) curve.

Reference: functional/classification/roc.py:40-80 (binary), :329-356 (multilabel).
"""
from typing import List, Optional, Tuple, Union

import torch
from torch import Tensor
from typing_extensions import Literal

from torchmetrics_forked_b200.functional.classification.precision_recall_curve import (
    _adjust_threshold_arg,
    _binary_clf_curve,
    _binary_precision_recall_curve_arg_validation,
    _binary_precision_recall_curve_tensor_validation,
    _binary_state_functional,
    _macro_curve,
    _multiclass_precision_recall_curve_arg_validation,
    _multiclass_precision_recall_curve_tensor_validation,
    _multiclass_state_functional,
    _multilabel_clf_curves,
    _multilabel_precision_recall_curve_arg_validation,
    _multilabel_precision_recall_curve_tensor_validation,
    _multilabel_state_functional,
)
from torchmetrics_forked_b200 import kernels
from torchmetrics_forked_b200.utilities.compute import _safe_divide
from torchmetrics_forked_b200.utilities.enums import ClassificationTask
from torchmetrics_forked_b200.utilities.prints import rank_zero_warn


def _roc_from_curve(fps: Tensor, tps: Tensor, thres: Tensor) -> Tuple[Tensor, Tensor, Tensor]:
    """Exact ROC: prepend the (0, 0) point, normalise by the totals (reference :53-78)."""
    tps = torch.cat([torch.zeros(1, dtype=tps.dtype, device=tps.device), tps])
    fps = torch.cat([torch.zeros(1, dtype=fps.dtype, device=fps.device), fps])
    thres = torch.cat([torch.ones(1, dtype=thres.dtype, device=thres.device), thres])
    if fps[-1] <= 0:
        rank_zero_warn(
            "No negative samples in targets, false positive value should be meaningless."
            " Returning zero tensor in false positive score",
            UserWarning,
        )
        fpr = torch.zeros_like(thres)
    else:
        fpr = fps / fps[-1]
    if tps[-1] <= 0:
        rank_zero_warn(
            "No positive samples in targets, true positive value should be meaningless."
            " Returning zero tensor in true positive score",
            UserWarning,
        )
        tpr = torch.zeros_like(thres)
    else:
        tpr = tps / tps[-1]
    return fpr, tpr, thres


def _binary_roc_compute(
    state: Union[Tensor, Tuple[Tensor, Tensor]], thresholds: Optional[Tensor], pos_label: int = 1
) -> Tuple[Tensor, Tensor, Tensor]:
    if isinstance(state, Tensor) and thresholds is not None:
        tps, fps, fns, tns = state[:, 1, 1], state[:, 0, 1], state[:, 1, 0], state[:, 0, 0]
        tpr = _safe_divide(tps, tps + fns).flip(0)
        fpr = _safe_divide(fps, fps + tns).flip(0)
        return fpr, tpr, thresholds.flip(0)
    fps, tps, thres = _binary_clf_curve(preds=state[0], target=state[1], pos_label=pos_label)
    return _roc_from_curve(fps, tps, thres)


def binary_roc(
    preds: Tensor,
    target: Tensor,
    thresholds: Optional[Union[int, List[float], Tensor]] = None,
    ignore_index: Optional[int] = None,
    validate_args: bool = True,
) -> Tuple[Tensor, Tensor, Tensor]:
    """``(fpr, tpr, thresholds)`` of a binary task (reference :83-159)."""
    if validate_args:
        _binary_precision_recall_curve_arg_validation(thresholds, ignore_index)
        _binary_precision_recall_curve_tensor_validation(preds, target, ignore_index)
    thresholds = _adjust_threshold_arg(thresholds, preds.device)
    state = _binary_state_functional(preds, target, thresholds, ignore_index, validate_args, "binary_roc")
    return _binary_roc_compute(state, thresholds)


def _multiclass_roc_compute(
    state: Union[Tensor, Tuple[Tensor, Tensor]], num_classes: int, thresholds: Optional[Tensor],
    average: Optional[str] = None,
) -> Union[Tuple[Tensor, Tensor, Tensor], Tuple[List[Tensor], List[Tensor], List[Tensor]]]:  # fmt: skip
    """One-vs-rest ROC per class (reference :162-204)."""
    if average == "micro":
        return _binary_roc_compute(state, thresholds, pos_label=1)
    if isinstance(state, Tensor) and thresholds is not None:
        tps, fps, fns, tns = state[:, :, 1, 1], state[:, :, 0, 1], state[:, :, 1, 0], state[:, :, 0, 0]
        tpr = _safe_divide(tps, tps + fns).flip(0).T
        fpr = _safe_divide(fps, fps + tns).flip(0).T
        thres = thresholds.flip(0)
        if average == "macro":
            return _macro_curve(list(fpr), list(tpr), thres.repeat(num_classes), True)
        return fpr, tpr, thres
    fps_l, tps_l, thr_l = kernels.clf_curves(state[0], state[1], None, class_targets=True)
    curves = [_roc_from_curve(fps_l[i], tps_l[i], thr_l[i]) for i in range(num_classes)]
    fpr, tpr, thres = ([c[k] for c in curves] for k in range(3))
    if average == "macro":
        return _macro_curve(fpr, tpr, torch.cat(thres, 0), True)
    return fpr, tpr, thres


def multiclass_roc(
    preds: Tensor,
    target: Tensor,
    num_classes: int,
    thresholds: Optional[Union[int, List[float], Tensor]] = None,
    average: Optional[Literal["micro", "macro"]] = None,
    ignore_index: Optional[int] = None,
    validate_args: bool = True,
) -> Union[Tuple[Tensor, Tensor, Tensor], Tuple[List[Tensor], List[Tensor], List[Tensor]]]:
    """Per-class ``(fpr, tpr, thresholds)`` (reference :207-326)."""
    if validate_args:
        _multiclass_precision_recall_curve_arg_validation(num_classes, thresholds, ignore_index, average)
        _multiclass_precision_recall_curve_tensor_validation(preds, target, num_classes, ignore_index)
    thresholds = _adjust_threshold_arg(thresholds, preds.device)
    state = _multiclass_state_functional(preds, target, num_classes, thresholds, ignore_index, average,
                                         validate_args, "multiclass_roc")
    return _multiclass_roc_compute(state, num_classes, thresholds, average)


def _multilabel_roc_compute(
    state: Union[Tensor, Tuple[Tensor, Tensor]],
    num_labels: int,
    thresholds: Optional[Tensor],
    ignore_index: Optional[int] = None,
) -> Union[Tuple[Tensor, Tensor, Tensor], Tuple[List[Tensor], List[Tensor], List[Tensor]]]:
    if isinstance(state, Tensor) and thresholds is not None:
        tps, fps, fns, tns = state[:, :, 1, 1], state[:, :, 0, 1], state[:, :, 1, 0], state[:, :, 0, 0]
        tpr = _safe_divide(tps, tps + fns).flip(0).T
        fpr = _safe_divide(fps, fps + tns).flip(0).T
        return fpr, tpr, thresholds.flip(0)
    fps_l, tps_l, thr_l = _multilabel_clf_curves(state, num_labels, ignore_index)
    fpr, tpr, thres = [], [], []
    for i in range(num_labels):
        f, t, th = _roc_from_curve(fps_l[i], tps_l[i], thr_l[i])
        fpr.append(f)
        tpr.append(t)
        thres.append(th)
    return fpr, tpr, thres


def multilabel_roc(
    preds: Tensor,
    target: Tensor,
    num_labels: int,
    thresholds: Optional[Union[int, List[float], Tensor]] = None,
    ignore_index: Optional[int] = None,
    validate_args: bool = True,
) -> Union[Tuple[Tensor, Tensor, Tensor], Tuple[List[Tensor], List[Tensor], List[Tensor]]]:
    """Per-label ``(fpr, tpr, thresholds)`` (reference :359-468)."""
    if validate_args:
        _multilabel_precision_recall_curve_arg_validation(num_labels, thresholds, ignore_index)
        _multilabel_precision_recall_curve_tensor_validation(preds, target, num_labels, ignore_index)
    thresholds = _adjust_threshold_arg(thresholds, preds.device)
    state = _multilabel_state_functional(preds, target, num_labels, thresholds, ignore_index, validate_args,
                                         "multilabel_roc")
    return _multilabel_roc_compute(state, num_labels, thresholds, ignore_index)


def roc(
    preds: Tensor,
    target: Tensor,
    task: Literal["binary", "multiclass", "multilabel"],
    thresholds: Optional[Union[int, List[float], Tensor]] = None,
    num_classes: Optional[int] = None,
    num_labels: Optional[int] = None,
    ignore_index: Optional[int] = None,
    validate_args: bool = True,
) -> Union[Tuple[Tensor, Tensor, Tensor], Tuple[List[Tensor], List[Tensor], List[Tensor]]]:
    """Task dispatcher (reference :471-550)."""
    task = ClassificationTask.from_str(task)
    if task == ClassificationTask.BINARY:
        return binary_roc(preds, target, thresholds, ignore_index, validate_args)
    if task == ClassificationTask.MULTICLASS:
        if not isinstance(num_classes, int):
            raise ValueError(f"`num_classes` is expected to be `int` but `{type(num_classes)} was passed.`")
        return multiclass_roc(preds, target, num_classes, thresholds, None, ignore_index, validate_args)
    if task == ClassificationTask.MULTILABEL:
        if not isinstance(num_labels, int):
            raise ValueError(f"`num_labels` is expected to be `int` but `{type(num_labels)} was passed.`")
        return multilabel_roc(preds, target, num_labels, thresholds, ignore_index, validate_args)
    raise ValueError(f"Task {task} not supported, expected one of {ClassificationTask}.")
