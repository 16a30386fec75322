"""Metric classes that are a StatScores update plus a reducer: see
:mod:`torchmetrics_forked_b200.functional.classification._stat_family`.

:func:`metric_family` returns ``(Binary<Name>, Multiclass<Name>, Multilabel<Name>, <Name>)``; the classes inherit
``__init__`` / ``update`` / the int64 states from the StatScores classes (i.e. the fused kernels) and only add
``compute``.  They are given their public names and the caller's module so that pickling and ``repr`` behave
as for hand-written classes.
"""
from typing import Any, Callable, Optional, Tuple

from torch import Tensor
from typing_extensions import Literal

from torchmetrics_forked_b200.classification.base import _ClassificationTaskWrapper
from torchmetrics_forked_b200.classification.stat_scores import (
    BinaryStatScores,
    MulticlassStatScores,
    MultilabelStatScores,
    _task_dispatch,
)
from torchmetrics_forked_b200.metric import Metric


def metric_family(
    name: str, reducer: Callable[..., Tensor], higher_is_better: bool, module: str
) -> Tuple[type, type, type, type]:
    class _Binary(BinaryStatScores):
        def compute(self) -> Tensor:
            tp, fp, tn, fn = self._final_state()
            return reducer(tp, fp, tn, fn, average="binary", multidim_average=self.multidim_average)

    class _Multiclass(MulticlassStatScores):
        plot_legend_name: str = "Class"

        def compute(self) -> Tensor:
            tp, fp, tn, fn = self._final_state()
            return reducer(tp, fp, tn, fn, average=self.average, multidim_average=self.multidim_average)

    class _Multilabel(MultilabelStatScores):
        plot_legend_name: str = "Label"

        def compute(self) -> Tensor:
            tp, fp, tn, fn = self._final_state()
            return reducer(
                tp, fp, tn, fn, average=self.average, multidim_average=self.multidim_average, multilabel=True
            )

    for cls, prefix in ((_Binary, "Binary"), (_Multiclass, "Multiclass"), (_Multilabel, "Multilabel")):
        cls.is_differentiable = False
        cls.higher_is_better = higher_is_better
        cls.full_state_update = False
        cls.plot_lower_bound = 0.0
        cls.plot_upper_bound = 1.0
        cls.__name__ = cls.__qualname__ = prefix + name
        cls.__module__ = module
        cls.__doc__ = f"{prefix} {name}: StatScores update (fused kernels) + ``{reducer.__name__}``."

    class _Task(_ClassificationTaskWrapper):
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
            kwargs.update(
                {"multidim_average": multidim_average, "ignore_index": ignore_index, "validate_args": validate_args}
            )
            return _task_dispatch(
                task, _Binary, _Multiclass, _Multilabel, threshold, num_classes, num_labels, average, top_k, kwargs
            )

    _Task.__name__ = _Task.__qualname__ = name
    _Task.__module__ = module
    _Task.__doc__ = f"``{name}(task=...)`` returns the task specific class."
    return _Binary, _Multiclass, _Multilabel, _Task
