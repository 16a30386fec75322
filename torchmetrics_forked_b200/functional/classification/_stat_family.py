"""Metrics that are "the stat-scores update + a reducer" (SURVEY §8 row f1).

Precision, recall, specificity and Hamming distance share the update with StatScores — the fused K1 / K1b
kernels — and differ only in a few fp32 torch expressions over the ``tp / fp / tn / fn`` states.  Instead of
four copies of the three task functions plus dispatcher, :func:`functional_family` builds them from the
reducer; the signatures, defaults and error messages are the reference's
(functional/classification/precision_recall.py, specificity.py, hamming.py).
"""
from typing import Callable, Optional, Tuple

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

Reducer = Callable[..., Tensor]  # (tp, fp, tn, fn, average, multidim_average="global", multilabel=False)


def _class_sum(x: Tensor, multidim_average: str) -> Tensor:
    return x.sum(dim=0 if multidim_average == "global" else 1)


def _precision_recall_reduce(
    stat: str, tp: Tensor, fp: Tensor, tn: Tensor, fn: Tensor, average: Optional[str],
    multidim_average: str = "global", multilabel: bool = False,
) -> Tensor:  # fmt: skip
    """``tp / (tp + fp)`` (precision) or ``tp / (tp + fn)`` (recall); reference precision_recall.py:37-57."""
    other = fp if stat == "precision" else fn
    if average == "binary":
        return _safe_divide(tp, tp + other)
    if average == "micro":
        tp, other = _class_sum(tp, multidim_average), _class_sum(other, multidim_average)
        return _safe_divide(tp, tp + other)
    return _adjust_weights_safe_divide(_safe_divide(tp, tp + other), average, multilabel, tp, fp, fn)


def _precision_reduce(tp, fp, tn, fn, average, multidim_average="global", multilabel=False) -> Tensor:
    return _precision_recall_reduce("precision", tp, fp, tn, fn, average, multidim_average, multilabel)


def _recall_reduce(tp, fp, tn, fn, average, multidim_average="global", multilabel=False) -> Tensor:
    return _precision_recall_reduce("recall", tp, fp, tn, fn, average, multidim_average, multilabel)


def _specificity_reduce(tp, fp, tn, fn, average, multidim_average="global", multilabel=False) -> Tensor:
    """``tn / (tn + fp)``; reference specificity.py:37-54."""
    if average == "binary":
        return _safe_divide(tn, tn + fp)
    if average == "micro":
        tn_s, fp_s = _class_sum(tn, multidim_average), _class_sum(fp, multidim_average)
        return _safe_divide(tn_s, tn_s + fp_s)
    return _adjust_weights_safe_divide(_safe_divide(tn, tn + fp), average, multilabel, tp, fp, fn)


def _hamming_distance_reduce(tp, fp, tn, fn, average, multidim_average="global", multilabel=False) -> Tensor:
    """One minus accuracy in each task's own sense (multiclass: ``tp / (tp + fn)``); reference hamming.py:37-83."""
    if average == "binary":
        return 1 - _safe_divide(tp + tn, tp + fp + tn + fn)
    if average == "micro":
        tp_s, fn_s = _class_sum(tp, multidim_average), _class_sum(fn, multidim_average)
        if multilabel:
            fp_s, tn_s = _class_sum(fp, multidim_average), _class_sum(tn, multidim_average)
            return 1 - _safe_divide(tp_s + tn_s, tp_s + tn_s + fp_s + fn_s)
        return 1 - _safe_divide(tp_s, tp_s + fn_s)
    hit = _safe_divide(tp + tn, tp + tn + fp + fn) if multilabel else _safe_divide(tp, tp + fn)
    return _adjust_weights_safe_divide(1 - hit, average, multilabel, tp, fp, fn)


def functional_family(stem: str, reducer: Reducer) -> Tuple[Callable, Callable, Callable, Callable]:
    """``(binary_<stem>, multiclass_<stem>, multilabel_<stem>, <stem>)`` built around ``reducer``."""

    def binary(
        preds: Tensor,
        target: Tensor,
        threshold: float = 0.5,
        multidim_average: Literal["global", "samplewise"] = "global",
        ignore_index: Optional[int] = None,
        validate_args: bool = True,
    ) -> Tensor:
        if validate_args:
            _binary_stat_scores_arg_validation(threshold, multidim_average, ignore_index)
            _binary_stat_scores_tensor_validation(preds, target, multidim_average, ignore_index)
        _require_global(multidim_average)
        tp, fp, tn, fn = _binary_stat_scores_functional_state(
            preds, target, threshold, ignore_index, validate_args, f"binary_{stem}",
            multidim_average=multidim_average,
        )
        return reducer(tp, fp, tn, fn, average="binary", multidim_average=multidim_average)

    def multiclass(
        preds: Tensor,
        target: Tensor,
        num_classes: int,
        average: Optional[Literal["micro", "macro", "weighted", "none"]] = "macro",
        top_k: int = 1,
        multidim_average: Literal["global", "samplewise"] = "global",
        ignore_index: Optional[int] = None,
        validate_args: bool = True,
    ) -> Tensor:
        if validate_args:
            _multiclass_stat_scores_arg_validation(num_classes, top_k, average, multidim_average, ignore_index)
            _multiclass_stat_scores_tensor_validation(preds, target, num_classes, multidim_average, ignore_index)
        _require_global(multidim_average, top_k)
        tp, fp, tn, fn = _multiclass_stat_scores_functional_state(
            preds, target, num_classes, average, ignore_index, validate_args, f"multiclass_{stem}", top_k,
            multidim_average=multidim_average,
        )
        return reducer(tp, fp, tn, fn, average=average, multidim_average=multidim_average)

    def multilabel(
        preds: Tensor,
        target: Tensor,
        num_labels: int,
        threshold: float = 0.5,
        average: Optional[Literal["micro", "macro", "weighted", "none"]] = "macro",
        multidim_average: Literal["global", "samplewise"] = "global",
        ignore_index: Optional[int] = None,
        validate_args: bool = True,
    ) -> Tensor:
        if validate_args:
            _multilabel_stat_scores_arg_validation(num_labels, threshold, average, multidim_average, ignore_index)
            _multilabel_stat_scores_tensor_validation(preds, target, num_labels, multidim_average, ignore_index)
        _require_global(multidim_average)
        tp, fp, tn, fn = _multilabel_stat_scores_functional_state(
            preds, target, num_labels, threshold, ignore_index, validate_args, f"multilabel_{stem}",
            multidim_average=multidim_average,
        )
        return reducer(tp, fp, tn, fn, average=average, multidim_average=multidim_average, multilabel=True)

    def by_task(
        preds: Tensor,
        target: Tensor,
        task: Literal["binary", "multiclass", "multilabel"],
        threshold: float = 0.5,
        num_classes: Optional[int] = None,
        num_labels: Optional[int] = None,
        average: Optional[Literal["micro", "macro", "weighted", "none"]] = "micro",
        multidim_average: Optional[Literal["global", "samplewise"]] = "global",
        top_k: Optional[int] = 1,
        ignore_index: Optional[int] = None,
        validate_args: bool = True,
    ) -> Tensor:
        task = ClassificationTask.from_str(task)
        if task == ClassificationTask.BINARY:
            return binary(preds, target, threshold, multidim_average, ignore_index, validate_args)
        if task == ClassificationTask.MULTICLASS:
            if not isinstance(num_classes, int):
                raise ValueError(f"`num_classes` is expected to be `int` but `{type(num_classes)} was passed.`")
            if not isinstance(top_k, int):
                raise ValueError(f"`top_k` is expected to be `int` but `{type(top_k)} was passed.`")
            return multiclass(preds, target, num_classes, average, top_k, multidim_average, ignore_index, validate_args)
        if task == ClassificationTask.MULTILABEL:
            if not isinstance(num_labels, int):
                raise ValueError(f"`num_labels` is expected to be `int` but `{type(num_labels)} was passed.`")
            return multilabel(preds, target, num_labels, threshold, average, multidim_average, ignore_index, validate_args)
        raise ValueError(f"Not handled value: {task}")

    for fn, name in ((binary, f"binary_{stem}"), (multiclass, f"multiclass_{stem}"),
                     (multilabel, f"multilabel_{stem}"), (by_task, stem)):
        fn.__name__ = fn.__qualname__ = name
        fn.__doc__ = f"``{name}``: the stat-scores kernels followed by ``{reducer.__name__}``."
    return binary, multiclass, multilabel, by_task
