"""PrecisionRecallCurve metric classes.

Reference: classification/precision_recall_curve.py (binary :45-243, multilabel :430-640, wrapper :643-675).
Binned mode keeps an int64 ``confmat`` sum-state of shape ``[T, 2, 2]`` / ``[T, L, 2, 2]``; exact mode keeps
``preds`` / ``target`` list states ("cat").  ``thresholds`` is a non-persistent buffer created on the CPU at
construction exactly like the reference's (:150-156, H4) — the kernels read that buffer, never a recomputed
linspace.
"""
from typing import Any, List, Optional, Tuple, Union

import torch
from torch import Tensor
from typing_extensions import Literal

from torchmetrics_forked_b200 import kernels
from torchmetrics_forked_b200.classification.base import _ClassificationTaskWrapper
from torchmetrics_forked_b200.functional.classification.precision_recall_curve import (
    _adjust_threshold_arg,
    _binary_precision_recall_curve_arg_validation,
    _binary_precision_recall_curve_compute,
    _binary_precision_recall_curve_format_exact,
    _flag_non_binary_targets,
    _binary_precision_recall_curve_tensor_validation,
    _binary_precision_recall_curve_update_,
    _multiclass_precision_recall_curve_arg_validation,
    _multiclass_precision_recall_curve_compute,
    _multiclass_precision_recall_curve_format_exact,
    _multiclass_precision_recall_curve_tensor_validation,
    _multiclass_precision_recall_curve_update_,
    _multilabel_precision_recall_curve_arg_validation,
    _multilabel_precision_recall_curve_compute,
    _multilabel_precision_recall_curve_format_exact,
    _multilabel_precision_recall_curve_tensor_validation,
    _multilabel_precision_recall_curve_update_,
    _threshold_plan,
)
from torchmetrics_forked_b200.metric import Metric
from torchmetrics_forked_b200.utilities.data import dim_zero_cat
from torchmetrics_forked_b200.utilities.enums import ClassificationTask


_F64_SCORES = 1 << 40  # travels with the error bits of the label-sharded sync, never stored in them


class _CurveStateMixin:
    """Shared state set-up: binned (``confmat``) vs exact (``preds`` / ``target`` lists)."""

    def _init_curve_state(self, thresholds: Optional[Union[int, List[float], Tensor]], confmat_shape) -> None:
        thresholds = _adjust_threshold_arg(thresholds)
        if thresholds is None:
            self.thresholds = thresholds
            self.add_state("preds", default=[], dist_reduce_fx="cat")
            self.add_state("target", default=[], dist_reduce_fx="cat")
            return
        self.register_buffer("thresholds", thresholds, persistent=False)
        ordered, perm = _threshold_plan(thresholds, torch.device("cpu"))
        # the usual case (ascending float32) bins against the `thresholds` buffer itself
        self._thr_direct = perm is None and thresholds.dtype == torch.float32 and thresholds.is_contiguous()
        self.register_buffer("_thr_sorted", None if self._thr_direct else ordered, persistent=False)
        self.register_buffer("_thr_perm", perm, persistent=False)
        self.add_state(
            "confmat", default=torch.zeros(len(thresholds), *confmat_shape, dtype=torch.long), dist_reduce_fx="sum"
        )

    def _bin_plan(self, device: torch.device) -> Tuple[Tensor, Optional[Tensor]]:
        thr = self.thresholds if self._thr_direct else self._thr_sorted
        if thr.device != device:
            raise RuntimeError(
                f"Expected all tensors to be on the same device, but the metric is on {thr.device} and the inputs"
                f" on {device}"
            )
        return thr, self._thr_perm

    # ---- label-sharded compute of the scalar exact-mode metrics (AUROC / AveragePrecision) under a process group ----
    _sharded_scores: bool = False  # subclasses whose compute() is a reduction of kernels.clf_curve_scores set this

    def _label_sharded_ok(self, dist_sync_fn: Any) -> bool:
        """Decided from the metric's configuration alone, so that every rank decides alike."""
        from torchmetrics_forked_b200 import _sharded

        if not (self._sharded_scores and self.thresholds is None and dist_sync_fn is None and _sharded.enabled()):
            return False
        if not self.__dict__.get("_in_compute_sync", False) or self.compute_on_cpu or self.device.type != "cuda":
            return False
        if getattr(self, "max_fpr", None) not in (None, 1, 1.0) or getattr(self, "average", None) == "micro":
            return False
        group = self.process_group
        try:
            return torch.distributed.get_backend(group) == "nccl"
        except Exception:  # noqa: BLE001
            return False

    def _sync_dist(self, dist_sync_fn: Any = None, process_group: Any = None) -> None:
        """``compute()`` of an exact-mode AUROC / AP does not gather the scores of all ranks onto every rank
        (reference metric.py:423-453): the states stay local and the sort is dealt out by label
        (:mod:`torchmetrics_forked_b200._sharded`).  Only the row counts and the data-error bits are exchanged here.
        An explicit ``sync()`` (and every other metric) takes the packed path of the base class."""
        if not self._label_sharded_ok(dist_sync_fn):
            return super()._sync_dist(dist_sync_fn, process_group)
        from torchmetrics_forked_b200 import _sharded

        group = process_group or self.process_group
        n_local = sum(int(p.shape[0]) for p in self.preds) if isinstance(self.preds, list) else int(self.preds.shape[0])
        word = int(self._err_word.item()) if self._err_word is not None else 0
        chunks = self.preds if isinstance(self.preds, list) else [self.preds]
        if any(c.dtype == torch.float64 for c in chunks):
            word |= _F64_SCORES  # what a rank stores is data, not configuration: the ranks tell each other
        rows, err_or = _sharded.exchange_rows_and_errors(n_local, word, self.device, group)
        if err_or & _F64_SCORES:
            # float64 scores are ranked among ALL rows before the sort (see _f64): gathered like any other state
            return super()._sync_dist(dist_sync_fn, process_group)
        self.__dict__["_err_merged"] = torch.tensor(err_or)
        labels = getattr(self, "num_labels", None) or getattr(self, "num_classes", None) or 1
        self.__dict__["_shard_info"] = _sharded.ShardInfo(group, rows, int(labels), self.device)

    def unsync(self, should_unsync: bool = True) -> None:
        self.__dict__.pop("_shard_info", None)
        return super().unsync(should_unsync)

    def _curve_state(self) -> Union[Tensor, Tuple[Tensor, Tensor]]:
        info = self.__dict__.get("_shard_info")
        if info is not None and self.thresholds is None:
            from torchmetrics_forked_b200 import _sharded

            preds = self.preds if isinstance(self.preds, list) else [self.preds]
            target = self.target if isinstance(self.target, list) else [self.target]
            return _sharded.ShardedChunks(preds, info), _sharded.ShardedChunks(target, info)
        if self.thresholds is None:
            preds, target = self.preds, self.target
            if self.compute_on_cpu and self.device.type == "cuda":
                # `compute_on_cpu=True` parks the per-update buffers in host memory after every update
                # (metric.py:172-173); the sort runs on the metric's device, so they come back for `compute()`
                preds = [p.to(self.device) for p in preds] if isinstance(preds, list) else preds.to(self.device)
                target = [t.to(self.device) for t in target] if isinstance(target, list) else target.to(self.device)
            if isinstance(preds, list) and len(preds) > 1:
                # the per-update buffers go to the kernels as they are (chunk table): no concatenated copy
                return list(preds), list(target)
            return dim_zero_cat(preds), dim_zero_cat(target)
        return self.confmat


class BinaryPrecisionRecallCurve(_CurveStateMixin, Metric):
    is_differentiable: bool = False
    higher_is_better: Optional[bool] = None
    full_state_update: bool = False

    preds: List[Tensor]
    target: List[Tensor]
    confmat: Tensor

    def __init__(
        self,
        thresholds: Optional[Union[int, List[float], Tensor]] = None,
        ignore_index: Optional[int] = None,
        validate_args: bool = True,
        **kwargs: Any,
    ) -> None:
        super().__init__(**kwargs)
        if validate_args:
            _binary_precision_recall_curve_arg_validation(thresholds, ignore_index)
        self.ignore_index = ignore_index
        self.validate_args = validate_args
        self._init_curve_state(thresholds, (2, 2))

    def update(self, preds: Tensor, target: Tensor) -> None:
        if self.validate_args:
            _binary_precision_recall_curve_tensor_validation(preds, target, self.ignore_index)
        dev = preds.device
        err = self._error_word(dev)
        if self.thresholds is None:
            if self.validate_args:
                _flag_non_binary_targets(target, self.ignore_index, err)
            p, t = _binary_precision_recall_curve_format_exact(preds, target, self.ignore_index, err)
            self.preds.append(p)
            self.target.append(t)
        else:
            thr, perm = self._bin_plan(dev)
            scratch = self._scratch_buffer("tb", kernels.thresh_bin_scratch_numel(1, thr.numel()), dev)
            _binary_precision_recall_curve_update_(
                preds, target, thr, perm, self.ignore_index, self.confmat, scratch, err
            )
        if self.validate_args:
            self._raise_on_data_errors()

    def compute(self) -> Tuple[Tensor, Tensor, Tensor]:
        return _binary_precision_recall_curve_compute(self._curve_state(), self.thresholds)


class MulticlassPrecisionRecallCurve(_CurveStateMixin, Metric):
    """One-vs-rest curves of a multiclass task (reference :200-385).

    Binned mode keeps the int64 ``[T, C, 2, 2]`` state updated in place by the fused softmax + binning kernel;
    exact mode stores the (conditionally softmaxed) ``[M, C]`` scores and ``[M]`` targets per update.  With
    ``average="micro"`` the binned state is ``[T, 2, 2]`` (the reference declares ``[T, C, 2, 2]`` and then adds
    a ``[T, 2, 2]`` update to it, which only broadcasts when ``T == C``).
    """

    is_differentiable: bool = False
    higher_is_better: Optional[bool] = None
    full_state_update: bool = False

    preds: List[Tensor]
    target: List[Tensor]
    confmat: Tensor

    def __init__(
        self,
        num_classes: int,
        thresholds: Optional[Union[int, List[float], Tensor]] = None,
        average: Optional[Literal["micro", "macro"]] = None,
        ignore_index: Optional[int] = None,
        validate_args: bool = True,
        **kwargs: Any,
    ) -> None:
        super().__init__(**kwargs)
        if validate_args:
            _multiclass_precision_recall_curve_arg_validation(num_classes, thresholds, ignore_index, average)
        self.num_classes = num_classes
        self.average = average
        self.ignore_index = ignore_index
        self.validate_args = validate_args
        self._init_curve_state(thresholds, (2, 2) if average == "micro" else (num_classes, 2, 2))

    def update(self, preds: Tensor, target: Tensor) -> None:
        if self.validate_args:
            _multiclass_precision_recall_curve_tensor_validation(preds, target, self.num_classes, self.ignore_index)
        dev = preds.device
        if self.thresholds is None:
            p, t = _multiclass_precision_recall_curve_format_exact(
                preds, target, self.num_classes, self.ignore_index, self.average, self.validate_args
            )
            self.preds.append(p)
            self.target.append(t)
            return
        err = self._error_word(dev)
        thr, perm = self._bin_plan(dev)
        labels = 1 if self.average == "micro" else self.num_classes
        scratch = self._scratch_buffer("tb", kernels.thresh_bin_scratch_numel(labels, thr.numel()), dev)
        _multiclass_precision_recall_curve_update_(
            preds, target, self.num_classes, thr, perm, self.ignore_index, self.average, self.confmat, scratch, err
        )
        if self.validate_args:
            self._raise_on_data_errors()

    def compute(self) -> Union[Tuple[Tensor, Tensor, Tensor], Tuple[List[Tensor], List[Tensor], List[Tensor]]]:
        return _multiclass_precision_recall_curve_compute(
            self._curve_state(), self.num_classes, self.thresholds, self.average
        )


class MultilabelPrecisionRecallCurve(_CurveStateMixin, Metric):
    is_differentiable: bool = False
    higher_is_better: Optional[bool] = None
    full_state_update: bool = False

    preds: List[Tensor]
    target: List[Tensor]
    confmat: Tensor

    def __init__(
        self,
        num_labels: int,
        thresholds: Optional[Union[int, List[float], Tensor]] = None,
        ignore_index: Optional[int] = None,
        validate_args: bool = True,
        **kwargs: Any,
    ) -> None:
        super().__init__(**kwargs)
        if validate_args:
            _multilabel_precision_recall_curve_arg_validation(num_labels, thresholds, ignore_index)
        self.num_labels = num_labels
        self.ignore_index = ignore_index
        self.validate_args = validate_args
        self._init_curve_state(thresholds, (num_labels, 2, 2))

    def update(self, preds: Tensor, target: Tensor) -> None:
        if self.validate_args:
            _multilabel_precision_recall_curve_tensor_validation(preds, target, self.num_labels, self.ignore_index)
        dev = preds.device
        err = self._error_word(dev)
        if self.thresholds is None:
            if self.validate_args:
                _flag_non_binary_targets(target, self.ignore_index, err)
            p, t = _multilabel_precision_recall_curve_format_exact(preds, target, self.num_labels, err)
            self.preds.append(p)
            self.target.append(t)
        else:
            thr, perm = self._bin_plan(dev)
            scratch = self._scratch_buffer(
                "tb", kernels.thresh_bin_scratch_numel(self.num_labels, thr.numel()), dev
            )
            _multilabel_precision_recall_curve_update_(
                preds, target, self.num_labels, thr, perm, self.ignore_index, self.confmat, scratch, err
            )
        if self.validate_args:
            self._raise_on_data_errors()

    def compute(self) -> Union[Tuple[Tensor, Tensor, Tensor], Tuple[List[Tensor], List[Tensor], List[Tensor]]]:
        return _multilabel_precision_recall_curve_compute(
            self._curve_state(), self.num_labels, self.thresholds, self.ignore_index
        )


class PrecisionRecallCurve(_ClassificationTaskWrapper):
    """``PrecisionRecallCurve(task=...)`` (reference :643-675)."""

    def __new__(  # type: ignore[misc]
        cls,
        task: Literal["binary", "multiclass", "multilabel"],
        thresholds: Optional[Union[int, List[float], Tensor]] = None,
        num_classes: Optional[int] = None,
        num_labels: Optional[int] = None,
        ignore_index: Optional[int] = None,
        validate_args: bool = True,
        **kwargs: Any,
    ) -> Metric:
        task = ClassificationTask.from_str(task)
        kwargs.update({"thresholds": thresholds, "ignore_index": ignore_index, "validate_args": validate_args})
        if task == ClassificationTask.BINARY:
            return BinaryPrecisionRecallCurve(**kwargs)
        if task == ClassificationTask.MULTICLASS:
            if not isinstance(num_classes, int):
                raise ValueError(f"`num_classes` is expected to be `int` but `{type(num_classes)} was passed.`")
            return MulticlassPrecisionRecallCurve(num_classes, **kwargs)
        if task == ClassificationTask.MULTILABEL:
            if not isinstance(num_labels, int):
                raise ValueError(f"`num_labels` is expected to be `int` but `{type(num_labels)} was passed.`")
            return MultilabelPrecisionRecallCurve(num_labels, **kwargs)
        raise ValueError(f"Task {task} not supported!")
