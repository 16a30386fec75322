"""`MetricCollection` — several metrics behind one ``update / compute / forward / reset``.

Behavioural contract of the reference class (collections.py:34-660): metrics are stored under their
class name (sequence input) or the dict key; positional arguments go to every metric, keyword arguments
are filtered by each metric's ``update`` signature; after the first ``update`` metrics whose states are
identical form a *compute group* — only the first member is updated from then on and the others hold
references to its state tensors (which is why the kernels accumulate in place).
"""
from collections import OrderedDict
from copy import deepcopy
from typing import Any, Dict, Hashable, Iterable, Iterator, List, Optional, Sequence, Tuple, Union

import torch
from torch import Tensor
from torch.nn import ModuleDict

from torchmetrics_forked_b200.metric import Metric
from torchmetrics_forked_b200.utilities.data import allclose
from torchmetrics_forked_b200.utilities.prints import rank_zero_warn


def _find_duplicate_keys(results: Dict[str, Any]) -> bool:
    """True when two dict-valued results share an inner key (then keys get the metric name as prefix)."""
    seen = set()
    for key, value in results.items():
        for name in value if isinstance(value, dict) else (key,):
            if name in seen:
                return True
            seen.add(name)
    return False


_SHARED_FNS: List[Any] = []


def _shared_update_functions() -> List[Any]:
    """The kernels module and the three helpers of the shared update, imported once (the classification package
    imports this module, so the import cannot sit at the top; a function-level import on every update costs
    microseconds of a launch-bound call)."""
    if not _SHARED_FNS:
        from torchmetrics_forked_b200 import kernels
        from torchmetrics_forked_b200.functional.classification.confusion_matrix import (
            _multiclass_confusion_matrix_tensor_validation,
            _multiclass_hist_update_,
        )
        from torchmetrics_forked_b200.functional.classification.stat_scores import (
            _multiclass_stat_scores_tensor_validation,
        )

        _SHARED_FNS.extend([kernels, _multiclass_confusion_matrix_tensor_validation, _multiclass_hist_update_,
                            _multiclass_stat_scores_tensor_validation])
    return _SHARED_FNS


class MetricCollection(ModuleDict):
    """Chain metrics that share a call pattern.

    Args:
        metrics: a ``Metric``, a sequence of metrics / collections, or a dict name -> metric / collection.
        additional_metrics: further metrics when ``metrics`` is not a dict.
        prefix: string put in front of every output key.
        postfix: string appended to every output key.
        compute_groups: ``True`` (discover groups after the first update), ``False``, or an explicit list of
            lists of metric names.
    """

    _modules: Dict[str, Metric]  # type: ignore[assignment]
    _groups: Dict[int, List[str]]

    def __init__(
        self,
        metrics: Union[Metric, Sequence[Metric], Dict[str, Metric]],
        *additional_metrics: Metric,
        prefix: Optional[str] = None,
        postfix: Optional[str] = None,
        compute_groups: Union[bool, List[List[str]]] = True,
    ) -> None:
        super().__init__()
        self.prefix = self._check_arg(prefix, "prefix")
        self.postfix = self._check_arg(postfix, "postfix")
        self._enable_compute_groups = compute_groups
        self._groups_checked: bool = False
        self._state_is_copy: bool = False
        self._shared_plan_cache: Any = "unset"  # see _shared_argmax_plan; recomputed whenever the groups change
        self.add_metrics(metrics, *additional_metrics)

    # ------------------------------------------------------------------ the four verbs
    @torch.jit.unused
    def forward(self, *args: Any, **kwargs: Any) -> Dict[str, Any]:
        """``forward`` of every metric; returns name -> batch value."""
        return self._compute_and_reduce("forward", *args, **kwargs)

    def update(self, *args: Any, **kwargs: Any) -> None:
        """``update`` of every metric (of one member per compute group once groups are known)."""
        if self._groups_checked:
            shared = self._shared_argmax_update(args, kwargs)
            for members in self._groups.values():
                head = getattr(self, members[0])
                if members[0] not in shared:
                    head.update(*args, **head._filter_kwargs(**kwargs))
                # the followers see the new state through the shared tensors, so their cached results are
                # stale too (the 1.3.0dev reference keeps serving the old value here; deliberate deviation)
                for name in members[1:]:
                    getattr(self, name)._computed = None
            if self._state_is_copy:
                # states were handed out as copies in between (or `forward` un-linked them): link them again
                self._compute_groups_create_state_ref()
        else:
            for m in self.values(copy_state=False):
                m.update(*args, **m._filter_kwargs(**kwargs))
            if self._enable_compute_groups:
                self._merge_compute_groups()
                self._compute_groups_create_state_ref()
                self._groups_checked = True

    def compute(self) -> Dict[str, Any]:
        return self._compute_and_reduce("compute")

    def reset(self) -> None:
        for m in self._modules.values():
            m.reset()
        if self._enable_compute_groups and self._groups_checked:
            self._compute_groups_create_state_ref()

    # ------------------------------------------------------------------ one argmax per collection
    def _shared_argmax_plan(self) -> Optional[Tuple[Optional[str], List[str]]]:
        """``(confusion-matrix head, [stat-score heads])`` that can be fed by ONE pass over the logits.

        The reference updates every compute group separately (collections.py:200-226), so e.g. ``Accuracy`` +
        ``ConfusionMatrix`` (different states, hence different groups) read ``preds`` twice and run argmax twice.
        A multiclass confusion-matrix family head (ConfusionMatrix, JaccardIndex, CohenKappa, MatthewsCorrCoef) and
        every multiclass stat-scores family head (StatScores, Accuracy, FBeta / F1, Precision, Recall, Specificity,
        HammingDistance; ``top_k == 1``, global averaging) with the same ``num_classes`` and ``ignore_index`` are
        updated together instead: the kernel fills the batch confusion matrix and the three stat-score histograms
        in one launch, small finalize kernels add them into each metric's own states.
        """
        from torchmetrics_forked_b200.classification.confusion_matrix import MulticlassConfusionMatrix
        from torchmetrics_forked_b200.classification.stat_scores import MulticlassStatScores

        heads = [members[0] for members in self._groups.values()]
        for name in heads:
            cm = getattr(self, name)
            if not isinstance(cm, MulticlassConfusionMatrix) or cm.num_classes > 2048:
                continue
            subs = [
                n for n in heads
                if isinstance(getattr(self, n), MulticlassStatScores)
                and getattr(self, n).top_k == 1 and getattr(self, n).multidim_average == "global"
                and getattr(self, n).num_classes == cm.num_classes and getattr(self, n).ignore_index == cm.ignore_index
            ]  # fmt: skip
            if subs:
                return name, subs
        # no confusion-matrix head: stat-score heads that differ only in their averaging (Accuracy "micro" keeps
        # [1] states, F1 "macro" [C] states — two groups, two argmax passes in the reference) still share one pass
        stats = [
            n for n in heads
            if isinstance(getattr(self, n), MulticlassStatScores)
            and getattr(self, n).top_k == 1 and getattr(self, n).multidim_average == "global"
            and getattr(self, n).num_classes <= 2048
        ]  # fmt: skip
        for first in stats:
            a = getattr(self, first)
            same = [n for n in stats if getattr(self, n).num_classes == a.num_classes
                    and getattr(self, n).ignore_index == a.ignore_index]
            if len(same) >= 2:
                return None, same
        return None

    def _shared_argmax_update(self, args: Tuple[Any, ...], kwargs: Dict[str, Any]) -> List[str]:
        """Run the shared update if the collection and this call qualify; returns the heads it has served."""
        if self._shared_plan_cache == "unset":
            self._shared_plan_cache = self._shared_argmax_plan()
        plan = self._shared_plan_cache
        if plan is None or kwargs or len(args) != 2:
            return []
        preds, target = args
        if not (isinstance(preds, Tensor) and isinstance(target, Tensor) and preds.is_cuda and target.is_cuda):
            return []
        kernels, _multiclass_confusion_matrix_tensor_validation, _multiclass_hist_update_, \
            _multiclass_stat_scores_tensor_validation = _shared_update_functions()

        cm = getattr(self, plan[0]) if plan[0] is not None else None
        subs = [getattr(self, n) for n in plan[1]]
        members = ([cm] if cm is not None else []) + subs
        owner = members[0]  # holds the scratch and the error word of the shared launch
        if any(m.compute_on_cpu or m.dist_sync_on_step for m in members):
            return []
        dev, C = preds.device, owner.num_classes
        if (cm is not None and cm.confmat.device != dev) or any(m.tp.device != dev for m in subs):
            return []  # let the individual updates raise their usual device error
        if cm is not None and cm.validate_args:
            _multiclass_confusion_matrix_tensor_validation(preds, target, C, cm.ignore_index)
        for m in subs:
            if m.validate_args:
                _multiclass_stat_scores_tensor_validation(preds, target, C, m.multidim_average, m.ignore_index)
        for m in members:  # what Metric._wrap_update does around an update
            m.__dict__["_computed"] = None
            m.__dict__["_update_count"] += 1
        with torch.no_grad():
            scratch = owner._scratch_buffer("both", kernels.mc_scratch_numel(C, kernels.MC_BOTH), dev)
            err = owner._error_word(dev)
            if cm is not None:
                _multiclass_hist_update_(preds, target, C, owner.ignore_index, kernels.MC_BOTH, scratch, err)
                kernels.mc_shared_confmat(scratch, C, cm.confmat)
            else:  # the three histograms only, at the place the shared finalize expects them
                _multiclass_hist_update_(preds, target, C, owner.ignore_index, kernels.MC_STATS, scratch[C * C :], err)
            for i, m in enumerate(subs):
                kernels.mc_shared_stats(
                    scratch, C, m.average == "micro", i == len(subs) - 1, m.tp, m.fp, m.tn, m.fn
                )
        if any(m.validate_args for m in members):
            owner._raise_on_data_errors()
        return [m for m in plan[1]] + ([plan[0]] if plan[0] is not None else [])

    # ------------------------------------------------------------------ compute groups
    def _merge_compute_groups(self) -> None:
        """Union groups whose leading metrics hold equal states; O(n^2) comparisons, done once."""
        merged_something = True
        while merged_something:
            merged_something = False
            keys = list(self._groups)
            for i, a in enumerate(keys):
                for b in keys[i + 1 :]:
                    m_a = getattr(self, self._groups[a][0])
                    m_b = getattr(self, self._groups[b][0])
                    if self._equal_metric_states(m_a, m_b):
                        self._groups[a].extend(self._groups.pop(b))
                        merged_something = True
                        break
                if merged_something:
                    break
        self._groups = dict(enumerate(self._groups.values()))

    @staticmethod
    def _equal_metric_states(metric1: Metric, metric2: Metric) -> bool:
        """Same state names, types, shapes and values."""
        if len(metric1._defaults) == 0 or len(metric2._defaults) == 0:
            return False
        if metric1._defaults.keys() != metric2._defaults.keys():
            return False
        for key in metric1._defaults:
            s1, s2 = getattr(metric1, key), getattr(metric2, key)
            if type(s1) != type(s2):  # noqa: E721
                return False
            # EVERY state has to agree.  The reference returns from inside this loop (collections.py:281-285), i.e.
            # it only ever compares the first state: two metrics whose `tp` coincide after the first batch (top-1
            # and top-2 accuracy with few hits, say) are merged and one of them silently reports the other's
            # numbers from then on.  Deliberate deviation.
            if isinstance(s1, Tensor) and isinstance(s2, Tensor):
                if not (s1.shape == s2.shape and allclose(s1, s2)):
                    return False
            elif isinstance(s1, list) and isinstance(s2, list):
                if len(s1) != len(s2) or not all(a.shape == b.shape and allclose(a, b) for a, b in zip(s1, s2)):
                    return False
        return True

    def _compute_groups_create_state_ref(self, copy: bool = False) -> None:
        """Point every group member at the head's state objects, or (``copy``) hand every follower deep copies.

        Linking always happens on ``copy=False``; copies are made once and the flag then stays up until the next
        link.  (The reference skips the link while the flag is up and clears it, collections.py:289-307, so the link
        comes one call late; the states it ends up with are the same.)
        """
        if copy and self._state_is_copy:
            return
        for members in self._groups.values():
            head = getattr(self, members[0])
            for name in members[1:]:
                follower = getattr(self, name)
                for state in head._defaults:
                    value = getattr(head, state)
                    setattr(follower, state, deepcopy(value) if copy else value)
                follower._update_count = head._update_count
                # data errors recorded by the head's kernels must also surface through the followers
                follower._err_word = head._err_word
        self._state_is_copy = copy

    @property
    def compute_groups(self) -> Dict[int, List[str]]:
        return self._groups

    def _init_compute_groups(self) -> None:
        if isinstance(self._enable_compute_groups, list):
            self._groups = dict(enumerate(self._enable_compute_groups))
            for members in self._groups.values():
                for name in members:
                    if name not in self:
                        raise ValueError(
                            f"Input {name} in `compute_groups` argument does not match a metric in the collection."
                            f" Please make sure that {self._enable_compute_groups} matches {self.keys(keep_base=True)}"
                        )
            self._groups_checked = True
        else:
            self._groups = {i: [str(k)] for i, k in enumerate(self.keys(keep_base=True))}

    # ------------------------------------------------------------------ results
    def _compute_and_reduce(self, method_name: str, *args: Any, **kwargs: Any) -> Dict[str, Any]:
        if method_name not in ("compute", "forward"):
            raise ValueError(f"method_name should be either 'compute' or 'forward', but got {method_name}")
        raw: Dict[str, Any] = {}
        if method_name == "forward" and self._groups_checked:
            # Members of a compute group alias ONE set of state tensors, and `forward` folds the batch into the
            # global state in place (`Metric._forward_sum_states`): run on linked members the shared tensors would
            # take the batch once per member.  Give every member its own copy first; each then accumulates the batch
            # itself (what the reference's `_reduce_states` ends up with, metric.py:392-421) and the next `update`
            # links them to the head again.
            self._compute_groups_create_state_ref(copy=True)
            members = self._modules.items()
        else:
            members = self.items(keep_base=True, copy_state=False)
        for k, m in members:
            raw[k] = m.compute() if method_name == "compute" else m(*args, **m._filter_kwargs(**kwargs))
        duplicates = _find_duplicate_keys(raw)
        flat: Dict[str, Any] = {}
        for k, m in self._modules.items():
            res = raw[k]
            if not isinstance(res, dict):
                flat[k] = res
                continue
            for key, v in res.items():
                if duplicates:
                    stripped = k.replace(getattr(m, "prefix", "") or "", "")
                    stripped = stripped.replace(getattr(m, "postfix", "") or "", "")
                    key = f"{stripped}_{key}"
                if getattr(m, "_from_collection", None) and m.prefix is not None:
                    key = f"{m.prefix}{key}"
                if getattr(m, "_from_collection", None) and m.postfix is not None:
                    key = f"{key}{m.postfix}"
                flat[key] = v
        return {self._set_name(k): v for k, v in flat.items()}

    # ------------------------------------------------------------------ construction
    def add_metrics(
        self, metrics: Union[Metric, Sequence[Metric], Dict[str, Metric]], *additional_metrics: Metric
    ) -> None:
        """Add metrics (or whole collections) to this collection."""
        if isinstance(metrics, Metric):
            metrics = [metrics]
        if isinstance(metrics, Sequence):
            metrics = list(metrics)
            ignored: list = []
            for m in additional_metrics:
                (metrics if isinstance(m, Metric) else ignored).append(m)
            if ignored:
                rank_zero_warn(
                    f"You have passes extra arguments {ignored} which are not `Metric` so they will be ignored."
                )
        elif additional_metrics:
            raise ValueError(
                f"You have passes extra arguments {additional_metrics} which are not compatible"
                f" with first passed dictionary {metrics} so they will be ignored."
            )

        if isinstance(metrics, dict):
            for name in sorted(metrics.keys()):  # deterministic order
                metric = metrics[name]
                if not isinstance(metric, (Metric, MetricCollection)):
                    raise ValueError(
                        f"Value {metric} belonging to key {name} is not an instance of"
                        " `torchmetrics.Metric` or `torchmetrics.MetricCollection`"
                    )
                if isinstance(metric, Metric):
                    self[name] = metric
                else:
                    self._absorb(metric, key_prefix=f"{name}_")
        elif isinstance(metrics, Sequence):
            for metric in metrics:
                if not isinstance(metric, (Metric, MetricCollection)):
                    raise ValueError(
                        f"Input {metric} to `MetricCollection` is not a instance of"
                        " `torchmetrics.Metric` or `torchmetrics.MetricCollection`"
                    )
                if isinstance(metric, Metric):
                    name = metric.__class__.__name__
                    if name in self:
                        raise ValueError(f"Encountered two metrics both named {name}")
                    self[name] = metric
                else:
                    self._absorb(metric, key_prefix="")
        else:
            raise ValueError(
                "Unknown input to MetricCollection. Expected, `Metric`, `MetricCollection` or `dict`/`sequence` of the"
                f" previous, but got {metrics}"
            )

        self._groups_checked = False
        self._shared_plan_cache = "unset"
        if self._enable_compute_groups:
            self._init_compute_groups()
        else:
            self._groups = {}

    def _absorb(self, other: "MetricCollection", key_prefix: str) -> None:
        """Flatten a nested collection into this one, remembering its prefix / postfix per metric."""
        for k, v in other.items(keep_base=False):
            v.postfix = other.postfix
            v.prefix = other.prefix
            v._from_collection = True
            self[f"{key_prefix}{k}"] = v

    # ------------------------------------------------------------------ dict protocol
    def _set_name(self, base: str) -> str:
        name = base if self.prefix is None else self.prefix + base
        return name if self.postfix is None else name + self.postfix

    def _to_renamed_ordered_dict(self) -> OrderedDict:
        return OrderedDict((self._set_name(k), v) for k, v in self._modules.items())

    def __iter__(self) -> Iterator[Hashable]:
        return iter(self.keys())

    def keys(self, keep_base: bool = False) -> Iterable[Hashable]:
        """Metric names, with prefix/postfix unless ``keep_base``."""
        if keep_base:  # one view type either way (`Module._modules` is a plain dict in current torch, an OrderedDict before)
            return OrderedDict(self._modules).keys()
        return self._to_renamed_ordered_dict().keys()

    def items(self, keep_base: bool = False, copy_state: bool = True) -> Iterable[Tuple[str, Metric]]:
        """(name, metric) pairs; ``copy_state=True`` un-links compute-group members first."""
        self._compute_groups_create_state_ref(copy_state)
        if keep_base:
            return OrderedDict(self._modules).items()
        return self._to_renamed_ordered_dict().items()

    def values(self, copy_state: bool = True) -> Iterable[Metric]:
        self._compute_groups_create_state_ref(copy_state)
        return self._modules.values()

    def __getitem__(self, key: str, copy_state: bool = True) -> Metric:
        self._compute_groups_create_state_ref(copy_state)
        return self._modules[key]

    @staticmethod
    def _check_arg(arg: Optional[str], name: str) -> Optional[str]:
        if arg is None or isinstance(arg, str):
            return arg
        raise ValueError(f"Expected input `{name}` to be a string, but got {type(arg)}")

    def clone(self, prefix: Optional[str] = None, postfix: Optional[str] = None) -> "MetricCollection":
        """Deep copy, optionally with a new prefix / postfix."""
        mc = deepcopy(self)
        if prefix:
            mc.prefix = self._check_arg(prefix, "prefix")
        if postfix:
            mc.postfix = self._check_arg(postfix, "postfix")
        return mc

    def persistent(self, mode: bool = True) -> None:
        for m in self.values(copy_state=False):
            m.persistent(mode)

    def set_dtype(self, dst_type: Union[str, torch.dtype]) -> "MetricCollection":
        for m in self.values(copy_state=False):
            m.set_dtype(dst_type)
        return self

    def __repr__(self) -> str:
        text = super().__repr__()[:-2]
        if self.prefix:
            text += f",\n  prefix={self.prefix}{',' if self.postfix else ''}"
        if self.postfix:
            text += f"{',' if not self.prefix else ''}\n  postfix={self.postfix}"
        return text + "\n)"
