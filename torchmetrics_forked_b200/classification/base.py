"""Base of the ``task=...`` wrapper classes (reference: classification/base.py:19-32)."""
from typing import Any

from torchmetrics_forked_b200.metric import Metric


class _ClassificationTaskWrapper(Metric):
    """``__new__`` of a subclass returns the task specific metric; the wrapper itself is never updated."""

    def update(self, *args: Any, **kwargs: Any) -> None:
        raise NotImplementedError(
            f"{self.__class__.__name__} metric does not have a global `update` method. Use the task specific metric."
        )

    def compute(self) -> None:
        raise NotImplementedError(
            f"{self.__class__.__name__} metric does not have a global `compute` method. Use the task specific metric."
        )
