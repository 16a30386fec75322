"""Matthews correlation coefficient from the fused confusion-matrix state
(reference: functional/classification/matthews_corrcoef.py).  SURVEY §8 row f1: only the reduction is new.
"""
from typing import Optional

import torch
from torch import Tensor
from typing_extensions import Literal

from torchmetrics_forked_b200.functional.classification.confusion_matrix import (
    _binary_confusion_matrix_arg_validation,
    _binary_confusion_matrix_tensor_validation,
    _multiclass_confusion_matrix_arg_validation,
    _multiclass_confusion_matrix_tensor_validation,
    _multilabel_confusion_matrix_arg_validation,
    _multilabel_confusion_matrix_tensor_validation,
    binary_confusion_matrix,
    multiclass_confusion_matrix,
    multilabel_confusion_matrix,
)
from torchmetrics_forked_b200.utilities.enums import ClassificationTask


def _left_to_right_sum(x: Tensor) -> Tensor:
    """float32 sum in index order.  The reference adds these products with Python's ``sum`` (one float32 add per
    class, left to right); the covariances below subtract numbers of size N^2 from each other, so the order of
    the additions decides low-order bits that the cancellation then magnifies.  On the device a one-CTA kernel adds
    in index order (``tmb_sum_sequential_f32``); the state never leaves the GPU."""
    from torchmetrics_forked_b200 import _native as nat

    x = x.detach().to(torch.float32).contiguous()
    if not x.is_cuda:  # a state that was moved to the host: the reference's own expression (sequential by nature)
        return sum(x, torch.zeros((), dtype=torch.float32))
    out = torch.empty((), dtype=torch.float32, device=x.device)
    with nat.on_device(x.device):
        rc = nat.lib().tmb_sum_sequential_f32(nat.ptr(x), x.numel(), nat.ptr(out), nat.stream_ptr(x.device))
    nat.check(rc, "tmb_sum_sequential_f32")
    return out


def _matthews_corrcoef_reduce(confmat: Tensor) -> Tensor:
    """Multiclass MCC ``(c s - t.p) / sqrt((s^2 - p.p)(s^2 - t.t))`` with the reference's special cases (:37-78).

    A multilabel ``[L, 2, 2]`` state is summed over labels into one binary matrix first.  Binary matrices: a
    perfect / perfectly wrong prediction returns +1 / -1 (in the state's integer dtype, like the reference), and
    a zero denominator is replaced by the epsilon-regularised expression of the BMC Genomics paper; other
    matrices with a zero denominator give 0.
    """
    if confmat.ndim == 3:
        confmat = confmat.sum(0)
    binary = confmat.numel() == 4
    if binary:
        tn, fp, fn, tp = confmat.reshape(-1)
        right, wrong = tp + tn, fp + fn
        if right != 0 and wrong == 0:
            return torch.tensor(1.0, dtype=confmat.dtype, device=confmat.device)
        if right == 0 and wrong != 0:
            return torch.tensor(-1.0, dtype=confmat.dtype, device=confmat.device)
    per_target = confmat.sum(dim=-1).float()
    per_pred = confmat.sum(dim=-2).float()
    correct = torch.trace(confmat).float()
    total = confmat.sum().float()
    numerator = correct * total - _left_to_right_sum(per_target * per_pred)
    denom = (total**2 - _left_to_right_sum(per_pred * per_pred)) * (total**2 - _left_to_right_sum(per_target * per_target))
    if denom == 0:
        if not binary:
            return torch.tensor(0, dtype=confmat.dtype, device=confmat.device)
        # a zero denominator means one row or one column of the 2x2 matrix is empty, so one of tp / tn and one of
        # fp / fn is zero: the reference's conditional assignments of `a` and `b` both happen
        eps = torch.tensor(torch.finfo(torch.float32).eps, dtype=torch.float32, device=confmat.device)
        numerator = torch.sqrt(eps) * (right - wrong)
        denom = (tp + fp + eps) * (tp + fn + eps) * (tn + fp + eps) * (tn + fn + eps)
    return numerator / torch.sqrt(denom)


def binary_matthews_corrcoef(
    preds: Tensor,
    target: Tensor,
    threshold: float = 0.5,
    ignore_index: Optional[int] = None,
    validate_args: bool = True,
) -> Tensor:
    """Reference: matthews_corrcoef.py:81-132."""
    if validate_args:
        _binary_confusion_matrix_arg_validation(threshold, ignore_index, normalize=None)
        _binary_confusion_matrix_tensor_validation(preds, target, ignore_index)
    return _matthews_corrcoef_reduce(binary_confusion_matrix(preds, target, threshold, None, ignore_index, validate_args))


def multiclass_matthews_corrcoef(
    preds: Tensor,
    target: Tensor,
    num_classes: int,
    ignore_index: Optional[int] = None,
    validate_args: bool = True,
) -> Tensor:
    """Reference: matthews_corrcoef.py:135-189."""
    if validate_args:
        _multiclass_confusion_matrix_arg_validation(num_classes, ignore_index, normalize=None)
        _multiclass_confusion_matrix_tensor_validation(preds, target, num_classes, ignore_index)
    return _matthews_corrcoef_reduce(
        multiclass_confusion_matrix(preds, target, num_classes, None, ignore_index, validate_args)
    )


def multilabel_matthews_corrcoef(
    preds: Tensor,
    target: Tensor,
    num_labels: int,
    threshold: float = 0.5,
    ignore_index: Optional[int] = None,
    validate_args: bool = True,
) -> Tensor:
    """Reference: matthews_corrcoef.py:192-244."""
    if validate_args:
        _multilabel_confusion_matrix_arg_validation(num_labels, threshold, ignore_index, normalize=None)
        _multilabel_confusion_matrix_tensor_validation(preds, target, num_labels, ignore_index)
    return _matthews_corrcoef_reduce(
        multilabel_confusion_matrix(preds, target, num_labels, threshold, None, ignore_index, validate_args)
    )


def matthews_corrcoef(
    preds: Tensor,
    target: Tensor,
    task: Literal["binary", "multiclass", "multilabel"],
    threshold: float = 0.5,
    num_classes: Optional[int] = None,
    num_labels: Optional[int] = None,
    ignore_index: Optional[int] = None,
    validate_args: bool = True,
) -> Tensor:
    """Task dispatcher (reference :247-287)."""
    task = ClassificationTask.from_str(task)
    if task == ClassificationTask.BINARY:
        return binary_matthews_corrcoef(preds, target, threshold, ignore_index, validate_args)
    if task == ClassificationTask.MULTICLASS:
        if not isinstance(num_classes, int):
            raise ValueError(f"`num_classes` is expected to be `int` but `{type(num_classes)} was passed.`")
        return multiclass_matthews_corrcoef(preds, target, num_classes, ignore_index, validate_args)
    if task == ClassificationTask.MULTILABEL:
        if not isinstance(num_labels, int):
            raise ValueError(f"`num_labels` is expected to be `int` but `{type(num_labels)} was passed.`")
        return multilabel_matthews_corrcoef(preds, target, num_labels, threshold, ignore_index, validate_args)
    raise ValueError(f"Not handled value: {task}")
