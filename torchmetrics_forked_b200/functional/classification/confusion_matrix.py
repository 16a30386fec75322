"""Confusion matrices: argument checks, fused in-place state update, normalisation.

Reference: functional/classification/confusion_matrix.py.  The reference's ``_format`` + ``_update``
pair (argmax / sigmoid+threshold, flatten, boolean-mask, ``target*C+pred``, ``bincount``) is ONE kernel
here that adds straight into the int64 state; the ``_compute`` stage (normalisation of a tiny matrix)
is kept as torch expressions.
"""
from typing import Optional

import torch
from torch import Tensor
from typing_extensions import Literal

from torchmetrics_forked_b200 import kernels
from torchmetrics_forked_b200.utilities.checks import _check_same_shape
from torchmetrics_forked_b200.utilities.enums import ClassificationTask
from torchmetrics_forked_b200.utilities.prints import rank_zero_warn

_NORMALIZE = ("true", "pred", "all", "none", None)


def _confusion_matrix_reduce(confmat: Tensor, normalize: Optional[str] = None) -> Tensor:
    """Normalise over targets (``"true"``), predictions (``"pred"``) or everything (``"all"``); NaN -> 0.

    Reference: functional/classification/confusion_matrix.py:26-59.
    """
    if normalize not in _NORMALIZE:
        raise ValueError(f"Argument `normalize` needs to one of the following: {_NORMALIZE}")
    if normalize is None or normalize == "none":
        return confmat
    confmat = confmat if confmat.is_floating_point() else confmat.float()
    dims = {"true": -1, "pred": -2, "all": [-2, -1]}[normalize]
    confmat = confmat / confmat.sum(dim=dims, keepdim=True)
    bad = torch.isnan(confmat)
    n_bad = int(bad.sum())
    if n_bad:
        confmat[bad] = 0
        rank_zero_warn(f"{n_bad} NaN values found in confusion matrix have been replaced with zeros.")
    return confmat


def _check_ignore_index(ignore_index: Optional[int]) -> None:
    if ignore_index is not None and not isinstance(ignore_index, int):
        raise ValueError(f"Expected argument `ignore_index` to either be `None` or an integer, but got {ignore_index}")


def _check_normalize(normalize: Optional[str]) -> None:
    if normalize not in _NORMALIZE:
        raise ValueError(f"Expected argument `normalize` to be one of {_NORMALIZE}, but got {normalize}.")


# ------------------------------------------------------------------------------------------ binary
def _binary_confusion_matrix_arg_validation(
    threshold: float = 0.5, ignore_index: Optional[int] = None, normalize: Optional[str] = None
) -> None:
    """Reference: confusion_matrix.py:62-80."""
    if not (isinstance(threshold, float) and (0 <= threshold <= 1)):
        raise ValueError(f"Expected argument `threshold` to be a float in the [0,1] range, but got {threshold}.")
    _check_ignore_index(ignore_index)
    _check_normalize(normalize)


def _binary_confusion_matrix_tensor_validation(preds: Tensor, target: Tensor, ignore_index: Optional[int] = None) -> None:
    """Shape check on the host; the {0,1} content check of confusion_matrix.py:96-115 happens in the kernel."""
    _check_same_shape(preds, target)


def _binary_confusion_matrix_update_(
    preds: Tensor, target: Tensor, threshold: float, ignore_index: Optional[int], confmat: Tensor, scratch: Tensor,
    err: Tensor,
) -> None:
    """``confmat[target, pred > threshold] += 1`` in place (reference ``_format`` + ``_update``, :118-153).

    The sigmoid decision (any non-ignored pred outside [0,1]) is taken on the device, H1/H2 of SURVEY.md.
    """
    kernels.thresh_count(
        preds, target, num_labels=1, threshold=threshold, ignore_index=ignore_index,
        flag_counts_ignored=False, layout=kernels.TC_CONFMAT, scratch=scratch, out=(confmat,), err=err,
    )  # fmt: skip


def binary_confusion_matrix(
    preds: Tensor,
    target: Tensor,
    threshold: float = 0.5,
    normalize: Optional[Literal["true", "pred", "all", "none"]] = None,
    ignore_index: Optional[int] = None,
    validate_args: bool = True,
) -> Tensor:
    """``[2, 2]`` confusion matrix of a binary task (reference :167-227)."""
    if validate_args:
        _binary_confusion_matrix_arg_validation(threshold, ignore_index, normalize)
        _binary_confusion_matrix_tensor_validation(preds, target, ignore_index)
    confmat = torch.zeros(2, 2, dtype=torch.long, device=preds.device)
    err = torch.zeros(1, dtype=torch.int32, device=preds.device)
    scratch = torch.zeros(kernels.thresh_count_scratch_numel(1), dtype=torch.long, device=preds.device)
    _binary_confusion_matrix_update_(preds, target, threshold, ignore_index, confmat, scratch, err)
    if validate_args:
        kernels.raise_on_error_word(err, "binary_confusion_matrix")
    return _confusion_matrix_reduce(confmat, normalize)


# ------------------------------------------------------------------------------------------ multiclass
def _multiclass_confusion_matrix_arg_validation(
    num_classes: int, ignore_index: Optional[int] = None, normalize: Optional[str] = None
) -> None:
    """Reference: confusion_matrix.py:230-248."""
    if not isinstance(num_classes, int) or num_classes < 2:
        raise ValueError(f"Expected argument `num_classes` to be an integer larger than 1, but got {num_classes}")
    _check_ignore_index(ignore_index)
    _check_normalize(normalize)


def _multiclass_shape_validation(preds: Tensor, target: Tensor, num_classes: int) -> None:
    """The shape part of the reference's tensor validation (:263-286); contents are checked on the device."""
    if preds.ndim == target.ndim + 1:
        if not preds.is_floating_point():
            raise ValueError("If `preds` have one dimension more than `target`, `preds` should be a float tensor.")
        if preds.shape[1] != num_classes:
            raise ValueError(
                "If `preds` have one dimension more than `target`, `preds.shape[1]` should be"
                " equal to number of classes."
            )
        if preds.shape[2:] != target.shape[1:]:
            raise ValueError(
                "If `preds` have one dimension more than `target`, the shape of `preds` should be"
                " (N, C, ...), and the shape of `target` should be (N, ...)."
            )
    elif preds.ndim == target.ndim:
        if preds.shape != target.shape:
            raise ValueError(
                "The `preds` and `target` should have the same shape,",
                f" got `preds` with shape={preds.shape} and `target` with shape={target.shape}.",
            )
    else:
        raise ValueError(
            "Either `preds` and `target` both should have the (same) shape (N, ...), or `target` should be (N, ...)"
            " and `preds` should be (N, C, ...)."
        )


def _multiclass_confusion_matrix_tensor_validation(
    preds: Tensor, target: Tensor, num_classes: int, ignore_index: Optional[int] = None
) -> None:
    _multiclass_shape_validation(preds, target, num_classes)


def _multiclass_hist_update_(
    preds: Tensor, target: Tensor, num_classes: int, ignore_index: Optional[int], mode: int, out: Tensor, err: Tensor
) -> None:
    """Dispatch float logits (argmax fused) vs integer labels to the matching K1 entry point."""
    if preds.ndim == target.ndim + 1:
        if not preds.is_floating_point():
            raise ValueError("If `preds` have one dimension more than `target`, `preds` should be a float tensor.")
        kernels.mc_argmax_hist(preds, target, num_classes, ignore_index, mode, out, err)
    elif preds.is_floating_point():
        raise ValueError("`preds` with the same shape as `target` must hold integer class labels")
    else:
        kernels.mc_label_hist(preds, target, num_classes, ignore_index, mode, out, err)


def _multiclass_confusion_matrix_update_(
    preds: Tensor, target: Tensor, num_classes: int, ignore_index: Optional[int], confmat: Tensor, err: Tensor
) -> None:
    """``confmat[target, argmax(preds)] += 1`` in place (reference ``_format`` + ``_update``, :306-337)."""
    _multiclass_hist_update_(preds, target, num_classes, ignore_index, kernels.MC_CONFMAT, confmat, err)


def multiclass_confusion_matrix(
    preds: Tensor,
    target: Tensor,
    num_classes: int,
    normalize: Optional[Literal["true", "pred", "all", "none"]] = None,
    ignore_index: Optional[int] = None,
    validate_args: bool = True,
) -> Tensor:
    """``[C, C]`` confusion matrix, rows = target, columns = prediction (reference :351-416)."""
    if validate_args:
        _multiclass_confusion_matrix_arg_validation(num_classes, ignore_index, normalize)
        _multiclass_confusion_matrix_tensor_validation(preds, target, num_classes, ignore_index)
    confmat = torch.zeros(num_classes, num_classes, dtype=torch.long, device=preds.device)
    err = torch.zeros(1, dtype=torch.int32, device=preds.device)
    _multiclass_confusion_matrix_update_(preds, target, num_classes, ignore_index, confmat, err)
    if validate_args:
        kernels.raise_on_error_word(err, "multiclass_confusion_matrix")
    return _confusion_matrix_reduce(confmat, normalize)


# ------------------------------------------------------------------------------------------ multilabel
def _multilabel_confusion_matrix_arg_validation(
    num_labels: int, threshold: float = 0.5, ignore_index: Optional[int] = None, normalize: Optional[str] = None
) -> None:
    """Reference: confusion_matrix.py:419-441."""
    if not isinstance(num_labels, int) or num_labels < 2:
        raise ValueError(f"Expected argument `num_labels` to be an integer larger than 1, but got {num_labels}")
    if not (isinstance(threshold, float) and (0 <= threshold <= 1)):
        raise ValueError(f"Expected argument `threshold` to be a float, but got {threshold}.")
    _check_ignore_index(ignore_index)
    _check_normalize(normalize)


def _multilabel_confusion_matrix_tensor_validation(
    preds: Tensor, target: Tensor, num_labels: int, ignore_index: Optional[int] = None
) -> None:
    """Shape checks of confusion_matrix.py:455-462; the {0,1} content checks run in the kernel."""
    _check_same_shape(preds, target)
    if preds.shape[1] != num_labels:
        raise ValueError(
            "Expected both `target.shape[1]` and `preds.shape[1]` to be equal to the number of labels"
            f" but got {preds.shape[1]} and expected {num_labels}"
        )


def _multilabel_confusion_matrix_update_(
    preds: Tensor, target: Tensor, num_labels: int, threshold: float, ignore_index: Optional[int], confmat: Tensor,
    scratch: Tensor, err: Tensor,
) -> None:
    """``confmat[label, target, pred > threshold] += 1`` in place (reference :486-526)."""
    kernels.thresh_count(
        preds, target, num_labels=num_labels, threshold=threshold, ignore_index=ignore_index,
        flag_counts_ignored=True, layout=kernels.TC_CONFMAT, scratch=scratch, out=(confmat,), err=err,
    )  # fmt: skip


def multilabel_confusion_matrix(
    preds: Tensor,
    target: Tensor,
    num_labels: int,
    threshold: float = 0.5,
    normalize: Optional[Literal["true", "pred", "all", "none"]] = None,
    ignore_index: Optional[int] = None,
    validate_args: bool = True,
) -> Tensor:
    """``[L, 2, 2]`` confusion matrices, one per label (reference :540-604)."""
    if validate_args:
        _multilabel_confusion_matrix_arg_validation(num_labels, threshold, ignore_index, normalize)
        _multilabel_confusion_matrix_tensor_validation(preds, target, num_labels, ignore_index)
    confmat = torch.zeros(num_labels, 2, 2, dtype=torch.long, device=preds.device)
    err = torch.zeros(1, dtype=torch.int32, device=preds.device)
    scratch = torch.zeros(kernels.thresh_count_scratch_numel(num_labels), dtype=torch.long, device=preds.device)
    _multilabel_confusion_matrix_update_(preds, target, num_labels, threshold, ignore_index, confmat, scratch, err)
    if validate_args:
        kernels.raise_on_error_word(err, "multilabel_confusion_matrix")
    return _confusion_matrix_reduce(confmat, normalize)


def confusion_matrix(
    preds: Tensor,
    target: Tensor,
    task: Literal["binary", "multiclass", "multilabel"],
    threshold: float = 0.5,
    num_classes: Optional[int] = None,
    num_labels: Optional[int] = None,
    normalize: Optional[Literal["true", "pred", "all", "none"]] = None,
    ignore_index: Optional[int] = None,
    validate_args: bool = True,
) -> Tensor:
    """Task dispatcher (reference :607-665)."""
    task = ClassificationTask.from_str(task)
    if task == ClassificationTask.BINARY:
        return binary_confusion_matrix(preds, target, threshold, normalize, ignore_index, validate_args)
    if task == ClassificationTask.MULTICLASS:
        if not isinstance(num_classes, int):
            raise ValueError(f"`num_classes` is expected to be `int` but `{type(num_classes)} was passed.`")
        return multiclass_confusion_matrix(preds, target, num_classes, normalize, ignore_index, validate_args)
    if task == ClassificationTask.MULTILABEL:
        if not isinstance(num_labels, int):
            raise ValueError(f"`num_labels` is expected to be `int` but `{type(num_labels)} was passed.`")
        return multilabel_confusion_matrix(preds, target, num_labels, threshold, normalize, ignore_index, validate_args)
    raise ValueError(f"Task {task} not supported.")
