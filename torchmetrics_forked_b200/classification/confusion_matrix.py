"""ConfusionMatrix metric classes: one int64 ``confmat`` sum-state, updated in place by the fused kernels.

Reference: classification/confusion_matrix.py (binary :48-176, multiclass :179-330, multilabel :333-486,
wrapper :489-560).
"""
from typing import Any, Optional

import torch
from torch import Tensor
from typing_extensions import Literal

from torchmetrics_forked_b200 import kernels
from torchmetrics_forked_b200.classification.base import _ClassificationTaskWrapper
from torchmetrics_forked_b200.functional.classification.confusion_matrix import (
    _binary_confusion_matrix_arg_validation,
    _binary_confusion_matrix_tensor_validation,
    _binary_confusion_matrix_update_,
    _confusion_matrix_reduce,
    _multiclass_confusion_matrix_arg_validation,
    _multiclass_confusion_matrix_tensor_validation,
    _multiclass_confusion_matrix_update_,
    _multilabel_confusion_matrix_arg_validation,
    _multilabel_confusion_matrix_tensor_validation,
    _multilabel_confusion_matrix_update_,
)
from torchmetrics_forked_b200.metric import Metric
from torchmetrics_forked_b200.utilities.enums import ClassificationTask


def _state_on(state: Tensor, device: torch.device) -> Tensor:
    if state.device != device:
        raise RuntimeError(
            f"Expected all tensors to be on the same device, but the metric state is on {state.device} and the"
            f" inputs on {device}"
        )
    return state


class BinaryConfusionMatrix(Metric):
    """``[2, 2]`` confusion matrix, ``confmat[target, pred]``."""

    is_differentiable: bool = False
    higher_is_better: Optional[bool] = None
    full_state_update: bool = False
    confmat: Tensor

    def __init__(
        self,
        threshold: float = 0.5,
        ignore_index: Optional[int] = None,
        normalize: Optional[Literal["true", "pred", "all", "none"]] = None,
        validate_args: bool = True,
        **kwargs: Any,
    ) -> None:
        super().__init__(**kwargs)
        if validate_args:
            _binary_confusion_matrix_arg_validation(threshold, ignore_index, normalize)
        self.threshold = threshold
        self.ignore_index = ignore_index
        self.normalize = normalize
        self.validate_args = validate_args
        self.add_state("confmat", torch.zeros(2, 2, dtype=torch.long), dist_reduce_fx="sum")

    def update(self, preds: Tensor, target: Tensor) -> None:
        if self.validate_args:
            _binary_confusion_matrix_tensor_validation(preds, target, self.ignore_index)
        dev = preds.device
        scratch = self._scratch_buffer("tb", kernels.thresh_count_scratch_numel(1), dev)
        _binary_confusion_matrix_update_(
            preds, target, self.threshold, self.ignore_index, _state_on(self.confmat, dev), scratch,
            self._error_word(dev),
        )  # fmt: skip
        if self.validate_args:
            self._raise_on_data_errors()

    def compute(self) -> Tensor:
        return _confusion_matrix_reduce(self.confmat, self.normalize)


class MulticlassConfusionMatrix(Metric):
    """``[C, C]`` confusion matrix, rows = target, columns = prediction."""

    is_differentiable: bool = False
    higher_is_better: Optional[bool] = None
    full_state_update: bool = False
    confmat: Tensor

    def __init__(
        self,
        num_classes: int,
        ignore_index: Optional[int] = None,
        normalize: Optional[Literal["none", "true", "pred", "all"]] = None,
        validate_args: bool = True,
        **kwargs: Any,
    ) -> None:
        super().__init__(**kwargs)
        if validate_args:
            _multiclass_confusion_matrix_arg_validation(num_classes, ignore_index, normalize)
        self.num_classes = num_classes
        self.ignore_index = ignore_index
        self.normalize = normalize
        self.validate_args = validate_args
        self.add_state("confmat", torch.zeros(num_classes, num_classes, dtype=torch.long), dist_reduce_fx="sum")

    def update(self, preds: Tensor, target: Tensor) -> None:
        if self.validate_args:
            _multiclass_confusion_matrix_tensor_validation(preds, target, self.num_classes, self.ignore_index)
        dev = preds.device
        _multiclass_confusion_matrix_update_(
            preds, target, self.num_classes, self.ignore_index, _state_on(self.confmat, dev), self._error_word(dev)
        )
        if self.validate_args:
            self._raise_on_data_errors()

    def compute(self) -> Tensor:
        return _confusion_matrix_reduce(self.confmat, self.normalize)


class MultilabelConfusionMatrix(Metric):
    """``[L, 2, 2]`` confusion matrices, one per label."""

    is_differentiable: bool = False
    higher_is_better: Optional[bool] = None
    full_state_update: bool = False
    confmat: Tensor

    def __init__(
        self,
        num_labels: int,
        threshold: float = 0.5,
        ignore_index: Optional[int] = None,
        normalize: Optional[Literal["none", "true", "pred", "all"]] = None,
        validate_args: bool = True,
        **kwargs: Any,
    ) -> None:
        super().__init__(**kwargs)
        if validate_args:
            _multilabel_confusion_matrix_arg_validation(num_labels, threshold, ignore_index, normalize)
        self.num_labels = num_labels
        self.threshold = threshold
        self.ignore_index = ignore_index
        self.normalize = normalize
        self.validate_args = validate_args
        self.add_state("confmat", torch.zeros(num_labels, 2, 2, dtype=torch.long), dist_reduce_fx="sum")

    def update(self, preds: Tensor, target: Tensor) -> None:
        if self.validate_args:
            _multilabel_confusion_matrix_tensor_validation(preds, target, self.num_labels, self.ignore_index)
        dev = preds.device
        scratch = self._scratch_buffer("tb", kernels.thresh_count_scratch_numel(self.num_labels), dev)
        _multilabel_confusion_matrix_update_(
            preds, target, self.num_labels, self.threshold, self.ignore_index, _state_on(self.confmat, dev), scratch,
            self._error_word(dev),
        )  # fmt: skip
        if self.validate_args:
            self._raise_on_data_errors()

    def compute(self) -> Tensor:
        return _confusion_matrix_reduce(self.confmat, self.normalize)


class ConfusionMatrix(_ClassificationTaskWrapper):
    """``ConfusionMatrix(task=...)`` (reference :489-560)."""

    def __new__(  # type: ignore[misc]
        cls,
        task: Literal["binary", "multiclass", "multilabel"],
        threshold: float = 0.5,
        num_classes: Optional[int] = None,
        num_labels: Optional[int] = None,
        normalize: Optional[Literal["true", "pred", "all", "none"]] = None,
        ignore_index: Optional[int] = None,
        validate_args: bool = True,
        **kwargs: Any,
    ) -> Metric:
        task = ClassificationTask.from_str(task)
        kwargs.update({"normalize": normalize, "ignore_index": ignore_index, "validate_args": validate_args})
        if task == ClassificationTask.BINARY:
            return BinaryConfusionMatrix(threshold, **kwargs)
        if task == ClassificationTask.MULTICLASS:
            if not isinstance(num_classes, int):
                raise ValueError(f"`num_classes` is expected to be `int` but `{type(num_classes)} was passed.`")
            return MulticlassConfusionMatrix(num_classes, **kwargs)
        if task == ClassificationTask.MULTILABEL:
            if not isinstance(num_labels, int):
                raise ValueError(f"`num_labels` is expected to be `int` but `{type(num_labels)} was passed.`")
            return MultilabelConfusionMatrix(num_labels, threshold, **kwargs)
        raise ValueError(f"Task {task} not supported!")
