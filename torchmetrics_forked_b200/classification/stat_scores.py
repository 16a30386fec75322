"""StatScores metric classes: int64 ``tp/fp/tn/fn`` states updated in place by the fused kernels.

Reference: classification/stat_scores.py (``_AbstractStatScores`` :43-88, ``BinaryStatScores`` :91-190,
``MulticlassStatScores`` :193-342, ``MultilabelStatScores`` :345-487, ``StatScores`` :490-546).
"""
from typing import Any, Optional, Tuple

import torch
from torch import Tensor
from typing_extensions import Literal

from torchmetrics_forked_b200 import kernels
from torchmetrics_forked_b200.classification.base import _ClassificationTaskWrapper
from torchmetrics_forked_b200.functional.classification.stat_scores import (
    _binary_stat_scores_arg_validation,
    _binary_stat_scores_compute,
    _binary_stat_scores_tensor_validation,
    _binary_stat_scores_update_,
    _multiclass_stat_scores_arg_validation,
    _multiclass_stat_scores_compute,
    _multiclass_stat_scores_tensor_validation,
    _binary_stat_scores_functional_state,
    _multiclass_stat_scores_functional_state,
    _multilabel_stat_scores_functional_state,
    _multiclass_stat_scores_update_,
    _multiclass_stats_mode,
    _multilabel_stat_scores_arg_validation,
    _multilabel_stat_scores_compute,
    _multilabel_stat_scores_tensor_validation,
    _multilabel_stat_scores_update_,
    _require_global,
)
from torchmetrics_forked_b200.metric import Metric
from torchmetrics_forked_b200.utilities.data import dim_zero_cat
from torchmetrics_forked_b200.utilities.enums import ClassificationTask


class _AbstractStatScores(Metric):
    """Four int64 sum-states of ``size`` entries each."""

    tp: Tensor
    fp: Tensor
    tn: Tensor
    fn: Tensor

    def _create_state(self, size: int, multidim_average: str = "global") -> None:
        """int64 sum-states, or — ``samplewise`` — list states that collect the per-sample counts of every update
        (reference :52-70)."""
        for name in ("tp", "fp", "tn", "fn"):
            if multidim_average == "samplewise":
                self.add_state(name, [], dist_reduce_fx="cat")
            else:
                self.add_state(name, torch.zeros(size, dtype=torch.long), dist_reduce_fx="sum")

    def _append_samplewise(self, counts: Tuple[Tensor, Tensor, Tensor, Tensor]) -> None:
        for name, value in zip(("tp", "fp", "tn", "fn"), counts):
            getattr(self, name).append(value)

    def _final_state(self) -> Tuple[Tensor, Tensor, Tensor, Tensor]:
        return dim_zero_cat(self.tp), dim_zero_cat(self.fp), dim_zero_cat(self.tn), dim_zero_cat(self.fn)

    def _states_on(self, device: torch.device) -> Tuple[Tensor, Tensor, Tensor, Tensor]:
        """The state tensors the kernel adds into; they must already live on the input's device."""
        if self.tp.device != device:
            raise RuntimeError(
                f"Expected all tensors to be on the same device, but the metric states are on {self.tp.device} and the"
                f" inputs on {device}"
            )
        return self.tp, self.fp, self.tn, self.fn


class BinaryStatScores(_AbstractStatScores):
    """tp / fp / tn / fn / support of a binary task."""

    is_differentiable: bool = False
    higher_is_better: Optional[bool] = None
    full_state_update: bool = False

    def __init__(
        self,
        threshold: float = 0.5,
        multidim_average: Literal["global", "samplewise"] = "global",
        ignore_index: Optional[int] = None,
        validate_args: bool = True,
        **kwargs: Any,
    ) -> None:
        super(_AbstractStatScores, self).__init__(**kwargs)
        if validate_args:
            _binary_stat_scores_arg_validation(threshold, multidim_average, ignore_index)
        self.threshold = threshold
        self.multidim_average = multidim_average
        self.ignore_index = ignore_index
        self.validate_args = validate_args
        self._create_state(size=1, multidim_average=multidim_average)

    def update(self, preds: Tensor, target: Tensor) -> None:
        if self.validate_args:
            _binary_stat_scores_tensor_validation(preds, target, self.multidim_average, self.ignore_index)
        dev = preds.device
        if self.multidim_average == "samplewise":
            self._append_samplewise(
                _binary_stat_scores_functional_state(
                    preds, target, self.threshold, self.ignore_index, self.validate_args, "BinaryStatScores.update",
                    "samplewise",
                )
            )
            return
        tp, fp, tn, fn = self._states_on(dev)
        scratch = self._scratch_buffer("tb", kernels.thresh_count_scratch_numel(1), dev)
        _binary_stat_scores_update_(
            preds, target, self.threshold, self.ignore_index, tp, fp, tn, fn, scratch, self._error_word(dev)
        )
        if self.validate_args:
            self._raise_on_data_errors()

    def compute(self) -> Tensor:
        tp, fp, tn, fn = self._final_state()
        return _binary_stat_scores_compute(tp, fp, tn, fn, self.multidim_average)


class MulticlassStatScores(_AbstractStatScores):
    """tp / fp / tn / fn / support per class (or micro-summed) of a multiclass task."""

    is_differentiable: bool = False
    higher_is_better: Optional[bool] = None
    full_state_update: bool = False

    def __init__(
        self,
        num_classes: int,
        top_k: int = 1,
        average: Optional[Literal["micro", "macro", "weighted", "none"]] = "macro",
        multidim_average: Literal["global", "samplewise"] = "global",
        ignore_index: Optional[int] = None,
        validate_args: bool = True,
        **kwargs: Any,
    ) -> None:
        super(_AbstractStatScores, self).__init__(**kwargs)
        if validate_args:
            _multiclass_stat_scores_arg_validation(num_classes, top_k, average, multidim_average, ignore_index)
        _require_global(multidim_average, top_k)
        self.num_classes = num_classes
        self.top_k = top_k
        self.average = average
        self.multidim_average = multidim_average
        self.ignore_index = ignore_index
        self.validate_args = validate_args
        self._create_state(
            size=1 if (average == "micro" and top_k == 1) else num_classes, multidim_average=multidim_average
        )

    def update(self, preds: Tensor, target: Tensor) -> None:
        if self.validate_args:
            _multiclass_stat_scores_tensor_validation(
                preds, target, self.num_classes, self.multidim_average, self.ignore_index
            )
        dev = preds.device
        if self.multidim_average == "samplewise":
            self._append_samplewise(
                _multiclass_stat_scores_functional_state(
                    preds, target, self.num_classes, self.average, self.ignore_index, self.validate_args,
                    "MulticlassStatScores.update", self.top_k, "samplewise",
                )
            )
            return
        tp, fp, tn, fn = self._states_on(dev)
        mode = _multiclass_stats_mode(self.average, self.top_k)
        scratch = self._scratch_buffer(f"mc{mode}", kernels.mc_scratch_numel(self.num_classes, mode), dev)
        _multiclass_stat_scores_update_(
            preds, target, self.num_classes, self.average, self.ignore_index, tp, fp, tn, fn, scratch,
            self._error_word(dev), self.top_k,
        )  # fmt: skip
        if self.validate_args:
            self._raise_on_data_errors()

    def compute(self) -> Tensor:
        tp, fp, tn, fn = self._final_state()
        return _multiclass_stat_scores_compute(tp, fp, tn, fn, self.average, self.multidim_average)


class MultilabelStatScores(_AbstractStatScores):
    """tp / fp / tn / fn / support per label of a multilabel task."""

    is_differentiable: bool = False
    higher_is_better: Optional[bool] = None
    full_state_update: bool = False

    def __init__(
        self,
        num_labels: int,
        threshold: float = 0.5,
        average: Optional[Literal["micro", "macro", "weighted", "none"]] = "macro",
        multidim_average: Literal["global", "samplewise"] = "global",
        ignore_index: Optional[int] = None,
        validate_args: bool = True,
        **kwargs: Any,
    ) -> None:
        super(_AbstractStatScores, self).__init__(**kwargs)
        if validate_args:
            _multilabel_stat_scores_arg_validation(num_labels, threshold, average, multidim_average, ignore_index)
        self.num_labels = num_labels
        self.threshold = threshold
        self.average = average
        self.multidim_average = multidim_average
        self.ignore_index = ignore_index
        self.validate_args = validate_args
        self._create_state(size=num_labels, multidim_average=multidim_average)

    def update(self, preds: Tensor, target: Tensor) -> None:
        if self.validate_args:
            _multilabel_stat_scores_tensor_validation(
                preds, target, self.num_labels, self.multidim_average, self.ignore_index
            )
        dev = preds.device
        if self.multidim_average == "samplewise":
            self._append_samplewise(
                _multilabel_stat_scores_functional_state(
                    preds, target, self.num_labels, self.threshold, self.ignore_index, self.validate_args,
                    "MultilabelStatScores.update", "samplewise",
                )
            )
            return
        tp, fp, tn, fn = self._states_on(dev)
        scratch = self._scratch_buffer("tb", kernels.thresh_count_scratch_numel(self.num_labels), dev)
        _multilabel_stat_scores_update_(
            preds, target, self.num_labels, self.threshold, self.ignore_index, tp, fp, tn, fn, scratch,
            self._error_word(dev),
        )  # fmt: skip
        if self.validate_args:
            self._raise_on_data_errors()

    def compute(self) -> Tensor:
        tp, fp, tn, fn = self._final_state()
        return _multilabel_stat_scores_compute(tp, fp, tn, fn, self.average, self.multidim_average)


def _task_dispatch(
    task: str, binary_cls: type, multiclass_cls: type, multilabel_cls: type, threshold: float,
    num_classes: Optional[int], num_labels: Optional[int], average: Optional[str], top_k: Optional[int],
    kwargs: dict, strict_messages: bool = False,
) -> Metric:  # fmt: skip
    """Shared body of the stat-score based ``task=`` wrappers (reference e.g. classification/accuracy.py:486-522)."""
    task = ClassificationTask.from_str(task)
    if task == ClassificationTask.BINARY:
        return binary_cls(threshold, **kwargs)
    if task == ClassificationTask.MULTICLASS:
        if not isinstance(num_classes, int):
            raise ValueError(
                f"Optional arg `num_classes` must be type `int` when task is {task}. Got {type(num_classes)}"
                if strict_messages
                else f"`num_classes` is expected to be `int` but `{type(num_classes)} was passed.`"
            )
        if not isinstance(top_k, int):
            raise ValueError(
                f"Optional arg `top_k` must be type `int` when task is {task}. Got {type(top_k)}"
                if strict_messages
                else f"`top_k` is expected to be `int` but `{type(top_k)} was passed.`"
            )
        return multiclass_cls(num_classes, top_k, average, **kwargs)
    if task == ClassificationTask.MULTILABEL:
        if not isinstance(num_labels, int):
            raise ValueError(
                f"Optional arg `num_labels` must be type `int` when task is {task}. Got {type(num_labels)}"
                if strict_messages
                else f"`num_labels` is expected to be `int` but `{type(num_labels)} was passed.`"
            )
        return multilabel_cls(num_labels, threshold, average, **kwargs)
    raise ValueError(f"Task {task} not supported!")


class StatScores(_ClassificationTaskWrapper):
    """``StatScores(task=...)`` returns the task specific class (reference :490-546)."""

    def __new__(  # type: ignore[misc]
        cls,
        task: Literal["binary", "multiclass", "multilabel"],
        threshold: float = 0.5,
        num_classes: Optional[int] = None,
        num_labels: Optional[int] = None,
        average: Optional[Literal["micro", "macro", "weighted", "none"]] = "micro",
        multidim_average: Optional[Literal["global", "samplewise"]] = "global",
        top_k: Optional[int] = 1,
        ignore_index: Optional[int] = None,
        validate_args: bool = True,
        **kwargs: Any,
    ) -> Metric:
        assert multidim_average is not None  # noqa: S101
        kwargs.update(
            {"multidim_average": multidim_average, "ignore_index": ignore_index, "validate_args": validate_args}
        )
        return _task_dispatch(
            task, BinaryStatScores, MulticlassStatScores, MultilabelStatScores, threshold, num_classes, num_labels,
            average, top_k, kwargs,
        )  # fmt: skip
