"""Cohen's kappa from the fused confusion-matrix state (reference: functional/classification/cohen_kappa.py).

SURVEY §8 row f1: the update is the confusion-matrix kernel (K1 / K1b); only the reduction is new.
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
    binary_confusion_matrix,
    multiclass_confusion_matrix,
)
from torchmetrics_forked_b200.utilities.enums import ClassificationTaskNoMultilabel

_WEIGHTS = ("linear", "quadratic", "none", None)


def _cohen_kappa_reduce(confmat: Tensor, weights: Optional[str] = None) -> Tensor:
    """``1 - sum(W * observed) / sum(W * expected)`` with the disagreement weights ``W`` (reference :33-54).

    ``expected`` is the outer product of the row and column totals over the grand total; ``W`` is 0 on the
    diagonal and 1 elsewhere (unweighted), ``|i - j|`` (linear) or ``(i - j)^2`` (quadratic).
    """
    if weights not in _WEIGHTS:
        raise ValueError(
            f"Received {weights} for argument ``weights`` but should be either None, 'linear' or 'quadratic'"
        )
    observed = confmat if confmat.is_floating_point() else confmat.float()
    n = observed.shape[0]
    per_pred = observed.sum(dim=0, keepdim=True)
    per_target = observed.sum(dim=1, keepdim=True)
    expected = per_target @ per_pred / per_pred.sum()
    idx = torch.arange(n, dtype=observed.dtype, device=observed.device)
    gap = idx.unsqueeze(0) - idx.unsqueeze(1)  # gap[i, j] = j - i
    if weights in (None, "none"):
        w = (gap != 0).to(observed.dtype)
    elif weights == "linear":
        w = gap.abs()
    else:
        w = torch.pow(gap, 2.0)
    return 1 - torch.sum(w * observed) / torch.sum(w * expected)


def _check_weights(weights: Optional[str]) -> None:
    if weights not in _WEIGHTS:
        raise ValueError(f"Expected argument `weight` to be one of {_WEIGHTS}, but got {weights}.")


def _binary_cohen_kappa_arg_validation(
    threshold: float = 0.5, ignore_index: Optional[int] = None, weights: Optional[str] = None
) -> None:
    _binary_confusion_matrix_arg_validation(threshold, ignore_index, normalize=None)
    _check_weights(weights)


def _multiclass_cohen_kappa_arg_validation(
    num_classes: int, ignore_index: Optional[int] = None, weights: Optional[str] = None
) -> None:
    _multiclass_confusion_matrix_arg_validation(num_classes, ignore_index, normalize=None)
    _check_weights(weights)


def binary_cohen_kappa(
    preds: Tensor,
    target: Tensor,
    threshold: float = 0.5,
    weights: Optional[Literal["linear", "quadratic", "none"]] = None,
    ignore_index: Optional[int] = None,
    validate_args: bool = True,
) -> Tensor:
    """Reference: cohen_kappa.py:75-138."""
    if validate_args:
        _binary_cohen_kappa_arg_validation(threshold, ignore_index, weights)
        _binary_confusion_matrix_tensor_validation(preds, target, ignore_index)
    confmat = binary_confusion_matrix(preds, target, threshold, None, ignore_index, validate_args)
    return _cohen_kappa_reduce(confmat, weights)


def multiclass_cohen_kappa(
    preds: Tensor,
    target: Tensor,
    num_classes: int,
    weights: Optional[Literal["linear", "quadratic", "none"]] = None,
    ignore_index: Optional[int] = None,
    validate_args: bool = True,
) -> Tensor:
    """Reference: cohen_kappa.py:159-227."""
    if validate_args:
        _multiclass_cohen_kappa_arg_validation(num_classes, ignore_index, weights)
        _multiclass_confusion_matrix_tensor_validation(preds, target, num_classes, ignore_index)
    confmat = multiclass_confusion_matrix(preds, target, num_classes, None, ignore_index, validate_args)
    return _cohen_kappa_reduce(confmat, weights)


def cohen_kappa(
    preds: Tensor,
    target: Tensor,
    task: Literal["binary", "multiclass"],
    threshold: float = 0.5,
    num_classes: Optional[int] = None,
    weights: Optional[Literal["linear", "quadratic", "none"]] = None,
    ignore_index: Optional[int] = None,
    validate_args: bool = True,
) -> Tensor:
    """Task dispatcher (reference :230-271); there is no multilabel kappa."""
    task = ClassificationTaskNoMultilabel.from_str(task)
    if task == ClassificationTaskNoMultilabel.BINARY:
        return binary_cohen_kappa(preds, target, threshold, weights, ignore_index, validate_args)
    if task == ClassificationTaskNoMultilabel.MULTICLASS:
        if not isinstance(num_classes, int):
            raise ValueError(f"`num_classes` is expected to be `int` but `{type(num_classes)} was passed.`")
        return multiclass_cohen_kappa(preds, target, num_classes, weights, ignore_index, validate_args)
    raise ValueError(f"Not handled value: {task}")
