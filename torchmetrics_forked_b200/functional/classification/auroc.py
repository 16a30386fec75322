"""Area under the ROC curve (reference: functional/classification/auroc.py)."""
from typing import List, Optional, Tuple, Union

import torch
from torch import Tensor, tensor
from typing_extensions import Literal

from torchmetrics_forked_b200.functional.classification.precision_recall_curve import (
    _adjust_threshold_arg,
    _binary_precision_recall_curve_arg_validation,
    _binary_precision_recall_curve_tensor_validation,
    _binary_state_functional,
    _multiclass_precision_recall_curve_arg_validation,
    _multiclass_precision_recall_curve_tensor_validation,
    _multiclass_state_functional,
    _multilabel_precision_recall_curve_arg_validation,
    _multilabel_precision_recall_curve_tensor_validation,
    _multilabel_state_functional,
)
from torchmetrics_forked_b200.functional.classification.roc import (
    _binary_roc_compute,
    _multiclass_roc_compute,
    _multilabel_roc_compute,
)
from torchmetrics_forked_b200.utilities.compute import _auc_compute_without_check, _safe_divide
from torchmetrics_forked_b200.utilities.enums import ClassificationTask
from torchmetrics_forked_b200.utilities.prints import rank_zero_warn


def _reduce_scores(res: Tensor, average: Optional[str], weights: Optional[Tensor]) -> Tensor:
    """nan-aware macro / weighted averaging of per-label scores (the tail of the reference's ``_reduce_auroc``)."""
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


def _reduce_auroc(
    fpr: Union[Tensor, List[Tensor]],
    tpr: Union[Tensor, List[Tensor]],
    average: Optional[str] = "macro",
    weights: Optional[Tensor] = None,
) -> Tensor:
    """Trapezoidal area per class / label, then nan-aware averaging (reference :45-69)."""
    if isinstance(fpr, Tensor) and isinstance(tpr, Tensor):
        res = _auc_compute_without_check(fpr, tpr, 1.0, axis=1)
    else:
        res = torch.stack([_auc_compute_without_check(x, y, 1.0) for x, y in zip(fpr, tpr)])
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


def _exact_auroc_from_scores(scores: Tensor) -> Tensor:
    """Per-label AUROC (float32) from the float64 ``[L, 4]`` reductions of :func:`kernels.clf_curve_scores`.

    Same result as normalising the exact ROC curve and integrating it (reference roc.py:53-78 +
    utilities/compute.py:88-96) — 0 with a warning when a label has no negative or no positive sample — but
    from one fused pass over the curve instead of ~12 torch kernels per label.
    """
    if bool(torch.isnan(scores).any()):
        raise RuntimeError(
            "exact curves need non-negative scores (probabilities / sigmoid outputs): negative scores were found"
        )
    area, _, neg, pos = scores.unbind(1)
    if bool((neg <= 0).any()):
        rank_zero_warn(
            "No negative samples in targets, false positive value should be meaningless."
            " Returning zero tensor in false positive score",
            UserWarning,
        )
    if bool((pos <= 0).any()):
        rank_zero_warn(
            "No positive samples in targets, true positive value should be meaningless."
            " Returning zero tensor in true positive score",
            UserWarning,
        )
    ok = (neg > 0) & (pos > 0)
    denom = torch.where(ok, neg * pos, torch.ones_like(area))
    return torch.where(ok, area / denom, torch.zeros_like(area)).to(torch.float32)


def _binary_auroc_arg_validation(
    max_fpr: Optional[float] = None,
    thresholds: Optional[Union[int, List[float], Tensor]] = None,
    ignore_index: Optional[int] = None,
) -> None:
    _binary_precision_recall_curve_arg_validation(thresholds, ignore_index)
    if max_fpr is not None and not isinstance(max_fpr, float) and 0 < max_fpr <= 1:
        raise ValueError(f"Arguments `max_fpr` should be a float in range (0, 1], but got: {max_fpr}")


def _binary_auroc_compute(
    state: Union[Tensor, Tuple[Tensor, Tensor]],
    thresholds: Optional[Tensor],
    max_fpr: Optional[float] = None,
    pos_label: int = 1,
) -> Tensor:
    """AUROC, optionally the standardised partial area up to ``max_fpr`` (reference :82-106)."""
    if not (isinstance(state, Tensor) and thresholds is not None) and (max_fpr is None or max_fpr == 1):
        # exact mode, full area: sort + scan + fused area reduction, no curve is materialised on the host side
        from torchmetrics_forked_b200 import kernels

        preds, target = state[0], state[1]
        if isinstance(preds, Tensor) and preds.ndim > target.ndim:
            preds = preds[:, 0]
        scores = kernels.clf_curve_scores(
            kernels.as_columns(preds), kernels.as_columns(target), None, pos_label, want_ap=False
        )
        return _exact_auroc_from_scores(scores)[0]
    fpr, tpr, _ = _binary_roc_compute(state, thresholds, pos_label)
    if max_fpr is None or max_fpr == 1 or fpr.sum() == 0 or tpr.sum() == 0:
        return _auc_compute_without_check(fpr, tpr, 1.0)
    max_area: Tensor = tensor(max_fpr, device=fpr.device)
    # add one point at max_fpr by linear interpolation
    stop = torch.bucketize(max_area, fpr, out_int32=True, right=True)
    weight = (max_area - fpr[stop - 1]) / (fpr[stop] - fpr[stop - 1])
    interp_tpr: Tensor = torch.lerp(tpr[stop - 1], tpr[stop], weight)
    tpr = torch.cat([tpr[:stop], interp_tpr.view(1)])
    fpr = torch.cat([fpr[:stop], max_area.view(1)])
    partial_auc = _auc_compute_without_check(fpr, tpr, 1.0)
    # McClish correction: 0.5 if non-discriminant, 1 if maximal
    min_area: Tensor = 0.5 * max_area**2
    return 0.5 * (1 + (partial_auc - min_area) / (max_area - min_area))


def binary_auroc(
    preds: Tensor,
    target: Tensor,
    max_fpr: Optional[float] = None,
    thresholds: Optional[Union[int, List[float], Tensor]] = None,
    ignore_index: Optional[int] = None,
    validate_args: bool = True,
) -> Tensor:
    """Reference: auroc.py:109-177."""
    if validate_args:
        _binary_auroc_arg_validation(max_fpr, thresholds, ignore_index)
        _binary_precision_recall_curve_tensor_validation(preds, target, ignore_index)
    thresholds = _adjust_threshold_arg(thresholds, preds.device)
    state = _binary_state_functional(preds, target, thresholds, ignore_index, validate_args, "binary_auroc")
    return _binary_auroc_compute(state, thresholds, max_fpr)


def _multilabel_auroc_arg_validation(
    num_labels: int,
    average: Optional[str],
    thresholds: Optional[Union[int, List[float], Tensor]] = None,
    ignore_index: Optional[int] = None,
) -> None:
    _multilabel_precision_recall_curve_arg_validation(num_labels, thresholds, ignore_index)
    allowed_average = ("micro", "macro", "weighted", "none", None)
    if average not in allowed_average:
        raise ValueError(f"Expected argument `average` to be one of {allowed_average} but got {average}")


def _multiclass_auroc_arg_validation(
    num_classes: int,
    average: Optional[str] = "macro",
    thresholds: Optional[Union[int, List[float], Tensor]] = None,
    ignore_index: Optional[int] = None,
) -> None:
    _multiclass_precision_recall_curve_arg_validation(num_classes, thresholds, ignore_index)
    allowed_average = ("macro", "weighted", "none", None)
    if average not in allowed_average:
        raise ValueError(f"Expected argument `average` to be one of {allowed_average} but got {average}")


def _multiclass_auroc_compute(
    state: Union[Tensor, Tuple[Tensor, Tensor]],
    num_classes: int,
    average: Optional[str] = "macro",
    thresholds: Optional[Tensor] = None,
) -> Tensor:
    """One-vs-rest AUROC per class, then averaging (reference auroc.py:192-204)."""
    if not (isinstance(state, Tensor) and thresholds is not None):
        # exact mode: all classes in one batched sort + scan + area reduction; column 3 (positives per class) is
        # the reference's `_bincount(target, minlength=num_classes)` weight
        from torchmetrics_forked_b200 import kernels

        scores = kernels.clf_curve_scores(state[0], state[1], None, want_ap=False, class_targets=True)
        return _reduce_scores(_exact_auroc_from_scores(scores), average, scores[:, 3].float())
    fpr, tpr, _ = _multiclass_roc_compute(state, num_classes, thresholds)
    return _reduce_auroc(fpr, tpr, average, weights=state[0][:, 1, :].sum(-1))


def multiclass_auroc(
    preds: Tensor,
    target: Tensor,
    num_classes: int,
    average: Optional[Literal["macro", "weighted", "none"]] = "macro",
    thresholds: Optional[Union[int, List[float], Tensor]] = None,
    ignore_index: Optional[int] = None,
    validate_args: bool = True,
) -> Tensor:
    """Reference: auroc.py:207-292."""
    if validate_args:
        _multiclass_auroc_arg_validation(num_classes, average, thresholds, ignore_index)
        _multiclass_precision_recall_curve_tensor_validation(preds, target, num_classes, ignore_index)
    thresholds = _adjust_threshold_arg(thresholds, preds.device)
    state = _multiclass_state_functional(preds, target, num_classes, thresholds, ignore_index, None, validate_args,
                                         "multiclass_auroc")
    return _multiclass_auroc_compute(state, num_classes, average, thresholds)


def _flatten_valid(state: Tuple[Tensor, Tensor], ignore_index: Optional[int]) -> Tuple[Tensor, Tensor]:
    """Micro averaging in exact mode: every (sample, label) pair becomes one binary sample."""
    if not isinstance(state[0], Tensor):  # per-update chunks: flatten (and filter) each one where it lies
        pairs = [_flatten_valid((p, t), ignore_index) for p, t in zip(state[0], state[1])]
        return [p for p, _ in pairs], [t for _, t in pairs]
    preds, target = state[0].flatten(), state[1].flatten()
    if ignore_index is not None:
        from torchmetrics_forked_b200 import kernels

        preds, target = kernels.compact_not_ignored(preds.contiguous(), target.contiguous(), ignore_index)
    return preds, target


def _multilabel_auroc_compute(
    state: Union[Tensor, Tuple[Tensor, Tensor]],
    num_labels: int,
    average: Optional[str],
    thresholds: Optional[Tensor],
    ignore_index: Optional[int] = None,
) -> Tensor:
    """Reference: auroc.py:307-332."""
    if average == "micro":
        if isinstance(state, Tensor) and thresholds is not None:
            return _binary_auroc_compute(state.sum(1), thresholds, max_fpr=None)
        return _binary_auroc_compute(_flatten_valid(state, ignore_index), thresholds, max_fpr=None)
    if not (isinstance(state, Tensor) and thresholds is not None):
        # exact mode: all labels in one batched sort + scan + area reduction
        from torchmetrics_forked_b200 import kernels

        scores = kernels.clf_curve_scores(state[0], state[1], ignore_index, 1, want_ap=False)
        # weights: the reference counts (target == 1) per label (:331); column 3 is that count already
        return _reduce_scores(_exact_auroc_from_scores(scores), average, scores[:, 3].float())
    fpr, tpr, _ = _multilabel_roc_compute(state, num_labels, thresholds, ignore_index)
    return _reduce_auroc(fpr, tpr, average, weights=state[0][:, 1, :].sum(-1))


def multilabel_auroc(
    preds: Tensor,
    target: Tensor,
    num_labels: int,
    average: Optional[Literal["micro", "macro", "weighted", "none"]] = "macro",
    thresholds: Optional[Union[int, List[float], Tensor]] = None,
    ignore_index: Optional[int] = None,
    validate_args: bool = True,
) -> Tensor:
    """Reference: auroc.py:335-424."""
    if validate_args:
        _multilabel_auroc_arg_validation(num_labels, average, thresholds, ignore_index)
        _multilabel_precision_recall_curve_tensor_validation(preds, target, num_labels, ignore_index)
    thresholds = _adjust_threshold_arg(thresholds, preds.device)
    state = _multilabel_state_functional(preds, target, num_labels, thresholds, ignore_index, validate_args,
                                         "multilabel_auroc")
    return _multilabel_auroc_compute(state, num_labels, average, thresholds, ignore_index)


def auroc(
    preds: Tensor,
    target: Tensor,
    task: Literal["binary", "multiclass", "multilabel"],
    thresholds: Optional[Union[int, List[float], Tensor]] = None,
    num_classes: Optional[int] = None,
    num_labels: Optional[int] = None,
    average: Optional[Literal["macro", "weighted", "none"]] = "macro",
    max_fpr: Optional[float] = None,
    ignore_index: Optional[int] = None,
    validate_args: bool = True,
) -> Optional[Tensor]:
    """Task dispatcher (reference :427-479)."""
    task = ClassificationTask.from_str(task)
    if task == ClassificationTask.BINARY:
        return binary_auroc(preds, target, max_fpr, thresholds, ignore_index, validate_args)
    if task == ClassificationTask.MULTICLASS:
        if not isinstance(num_classes, int):
            raise ValueError(f"`num_classes` is expected to be `int` but `{type(num_classes)} was passed.`")
        return multiclass_auroc(preds, target, num_classes, average, thresholds, ignore_index, validate_args)
    if task == ClassificationTask.MULTILABEL:
        if not isinstance(num_labels, int):
            raise ValueError(f"`num_labels` is expected to be `int` but `{type(num_labels)} was passed.`")
        return multilabel_auroc(preds, target, num_labels, average, thresholds, ignore_index, validate_args)
    return None
