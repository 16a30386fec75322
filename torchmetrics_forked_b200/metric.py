"""`Metric` — the drop-in boundary of the classification-update hot path.

Same contract as the reference base class (metric.py:50-1073 of torchmetrics 1.3.0dev): states are
declared with ``add_state(name, default, dist_reduce_fx, persistent)``, ``update`` accumulates,
``compute`` reduces (cached, synchronised across ranks), ``forward`` returns the batch value while
accumulating globally, ``reset`` restores the defaults.  What is different underneath:

* subclasses in :mod:`torchmetrics_forked_b200.classification` accumulate **in place** into their int64
  states from sm_100a kernels (no temporaries, no host sync when ``validate_args=False``);
* data errors found by a kernel land in a per-metric device error word that is read at the next
  natural sync point (``compute``) or right away when ``validate_args=True``;
* ``_sync_dist`` (reference metric.py:423-453 + utilities/distributed.py:97-147) is replaced by ONE
  packed all-reduce for all "sum" states and one all-gather per "cat" state.  A user supplied
  ``dist_sync_fn`` still gets the reference's per-state call pattern.

Operator overloading (``CompositionalMetric``) and plotting are outside the hot path and not provided.
"""
import functools
import inspect
from abc import ABC, abstractmethod
from contextlib import contextmanager
from copy import deepcopy
from typing import Any, Callable, ClassVar, Dict, Generator, List, Optional, Sequence, Union

import torch
from torch import Tensor
from torch.nn import Module

from torchmetrics_forked_b200.utilities.data import (
    _flatten,
    _squeeze_if_scalar,
    apply_to_collection,
    dim_zero_cat,
    dim_zero_max,
    dim_zero_mean,
    dim_zero_min,
    dim_zero_sum,
)
from torchmetrics_forked_b200.utilities.distributed import all_gather_cat, gather_all_tensors, packed_all_reduce_sum
from torchmetrics_forked_b200.utilities.exceptions import TorchMetricsUserError
from torchmetrics_forked_b200.utilities.prints import rank_zero_warn

_NAMED_REDUCTIONS = {
    "sum": dim_zero_sum,
    "mean": dim_zero_mean,
    "max": dim_zero_max,
    "min": dim_zero_min,
    "cat": dim_zero_cat,
}
_CONSTANTS = (
    "higher_is_better",
    "is_differentiable",
    "full_state_update",
    "plot_lower_bound",
    "plot_upper_bound",
    "plot_legend_name",
)
_UNPICKLED = ("update", "compute", "_update_signature")
_DATA_ERROR_TEXT = {
    1: "`target` contains values outside [0, num_classes) that are not `ignore_index`",
    2: "label-valued `preds` contains values outside [0, num_classes)",
    4: "`target` contains values other than 0, 1 (and `ignore_index`)",
    8: "negative scores reached the exact-curve sort (stored scores must be probabilities / sigmoid outputs)",
}


class _entered:
    """``with`` over a context manager whose ``__enter__`` has already run."""

    def __init__(self, ctx: Any) -> None:
        self.ctx = ctx

    def __enter__(self) -> Any:
        return self.ctx

    def __exit__(self, *exc: Any) -> Any:
        return self.ctx.__exit__(*exc)


def jit_distributed_available() -> bool:
    """True when a default process group exists."""
    return torch.distributed.is_available() and torch.distributed.is_initialized()


def _bool_kwarg(kwargs: Dict[str, Any], name: str, default: bool, article: str = "an") -> bool:
    value = kwargs.pop(name, default)
    if not isinstance(value, bool):
        raise ValueError(f"Expected keyword argument `{name}` to be {article} `bool` but got {value}")
    return value


class Metric(Module, ABC):
    """Base class: state registry, wrapped ``update``/``compute``, ``forward``, sync, reset, device moves.

    Keyword arguments (all optional, as in the reference metric.py:112-147): ``compute_on_cpu``,
    ``dist_sync_on_step``, ``process_group``, ``dist_sync_fn``, ``distributed_available_fn``,
    ``sync_on_compute``, ``compute_with_cache``.
    """

    __jit_ignored_attributes__: ClassVar[List[str]] = ["device"]
    __jit_unused_properties__: ClassVar[List[str]] = [
        "is_differentiable",
        "higher_is_better",
        "plot_lower_bound",
        "plot_upper_bound",
        "plot_legend_name",
        "metric_state",
        "update_called",
        "update_count",
    ]
    is_differentiable: Optional[bool] = None
    higher_is_better: Optional[bool] = None
    full_state_update: Optional[bool] = None
    plot_lower_bound: Optional[float] = None
    plot_upper_bound: Optional[float] = None
    plot_legend_name: Optional[str] = None

    def __init__(self, **kwargs: Any) -> None:
        super().__init__()
        self._device = torch.device("cpu")

        self.compute_on_cpu = _bool_kwarg(kwargs, "compute_on_cpu", False)
        self.dist_sync_on_step = _bool_kwarg(kwargs, "dist_sync_on_step", False)
        self.process_group = kwargs.pop("process_group", None)
        self.dist_sync_fn = kwargs.pop("dist_sync_fn", None)
        if self.dist_sync_fn is not None and not callable(self.dist_sync_fn):
            raise ValueError(
                f"Expected keyword argument `dist_sync_fn` to be an callable function but got {self.dist_sync_fn}"
            )
        self.distributed_available_fn = kwargs.pop("distributed_available_fn", None) or jit_distributed_available
        self.sync_on_compute = _bool_kwarg(kwargs, "sync_on_compute", True, article="a")
        self.compute_with_cache = _bool_kwarg(kwargs, "compute_with_cache", True, article="a")
        if kwargs:
            names = ", ".join(f"`{k}`" for k in sorted(kwargs))
            raise ValueError(f"Unexpected keyword arguments: {names}")

        self._install_wrappers()
        self._computed: Any = None
        self._forward_cache: Any = None
        self._update_count = 0
        self._to_sync = self.sync_on_compute
        self._should_unsync = True
        self._enable_grad = False
        self._dtype_convert = False

        self._defaults: Dict[str, Union[List, Tensor]] = {}
        self._persistent: Dict[str, bool] = {}
        self._reductions: Dict[str, Union[str, Callable[..., Any], None]] = {}

        self._is_synced = False
        self._cache: Optional[Dict[str, Union[List[Tensor], Tensor]]] = None

        # kernel plumbing: device error word and zero-initialised scratch buffers, created on first use
        self._err_word: Optional[Tensor] = None
        self._scratch: Dict[Any, Tensor] = {}

    # ------------------------------------------------------------------ wrappers around user code
    def _install_wrappers(self) -> None:
        self._update_signature = inspect.signature(self.update)
        self.update: Callable = self._wrap_update(self.update)  # type: ignore[method-assign]
        self.compute: Callable = self._wrap_compute(self.compute)  # type: ignore[method-assign]

    def _wrap_update(self, update: Callable) -> Callable:
        @functools.wraps(update)
        def wrapped_func(*args: Any, **kwargs: Any) -> None:
            # plain (non-tensor, non-module) attributes: written through __dict__, because nn.Module.__setattr__
            # costs microseconds and a small update is only tens of microseconds of host work in total
            attrs = self.__dict__
            attrs["_computed"] = None
            attrs["_update_count"] += 1
            with torch.set_grad_enabled(self._enable_grad):
                try:
                    update(*args, **kwargs)
                except RuntimeError as err:
                    if "Expected all tensors to be on" in str(err):
                        name = self.__class__.__name__
                        raise RuntimeError(
                            "Encountered different devices in metric calculation (see stacktrace for details)."
                            " This could be due to the metric class not being on the same device as input."
                            f" Instead of `metric={name}(...)` try to do"
                            f" `metric={name}(...).to(device)` where"
                            " device corresponds to the device of the input."
                        ) from err
                    raise err
            if self.compute_on_cpu:
                self._move_list_states_to_cpu()

        return wrapped_func

    def _wrap_compute(self, compute: Callable) -> Callable:
        @functools.wraps(compute)
        def wrapped_func(*args: Any, **kwargs: Any) -> Any:
            if self._update_count == 0:
                rank_zero_warn(
                    f"The ``compute`` method of metric {self.__class__.__name__}"
                    " was called before the ``update`` method which may lead to errors,"
                    " as metric states have not yet been updated.",
                    UserWarning,
                )
            if self._computed is not None:
                return self._computed
            # natural sync point: surface data errors the kernels recorded while validate_args was off.  The batch
            # value inside `forward` is not such a point (its `.item()` would put a host sync into every step): the
            # epoch-level `compute()` reports.  When the states are about to be synchronised the error bits travel
            # with them and every rank raises AFTER the collective — a rank that raised before it would leave the
            # others waiting in the all-reduce.
            in_forward = self.__dict__.get("_in_batch", False)
            travels = not in_forward and self._errors_travel_with_sync()
            if not in_forward and not travels:
                self._raise_on_data_errors()
            self.__dict__["_in_compute_sync"] = True  # lets a subclass tell compute()'s sync from an explicit sync()
            try:
                ctx = self.sync_context(
                    dist_sync_fn=self.dist_sync_fn, should_sync=self._to_sync, should_unsync=self._should_unsync
                )
                ctx.__enter__()
            finally:
                self.__dict__["_in_compute_sync"] = False
            with _entered(ctx):
                if travels:
                    self._raise_on_data_errors(merged=self.__dict__.pop("_err_merged", None))
                value = _squeeze_if_scalar(compute(*args, **kwargs))
            if self.compute_with_cache:
                self._computed = value
            return value

        return wrapped_func

    @abstractmethod
    def update(self, *_: Any, **__: Any) -> None:
        """Accumulate one batch into the metric state."""

    @abstractmethod
    def compute(self) -> Any:
        """Reduce the (synchronised) state to the metric value."""

    # ------------------------------------------------------------------ kernel plumbing
    def _error_word(self, device: torch.device) -> Tensor:
        """int32[1] device word the kernels OR their data-error bits into."""
        if self._err_word is None or self._err_word.device != device:
            self._err_word = torch.zeros(1, dtype=torch.int32, device=device)
        return self._err_word

    def _scratch_buffer(self, key: str, numel: int, device: torch.device, dtype: torch.dtype = torch.int64) -> Tensor:
        """Zero-initialised scratch that persists between updates (the finalize kernels re-zero it)."""
        full_key = (key, numel, dtype, str(device))
        buf = self._scratch.get(full_key)
        if buf is None:
            buf = torch.zeros(numel, dtype=dtype, device=device)
            self._scratch[full_key] = buf
        return buf

    def _raise_on_data_errors(self, merged: Optional[Tensor] = None) -> None:
        """Read the device error word (one 4-byte copy) and raise ``RuntimeError`` if a kernel set it.  ``merged``:
        the error bits of all ranks as they came back from the state sync (OR-ed with this rank's)."""
        if self._err_word is None and merged is None:
            return
        bits = int(self._err_word.item()) if self._err_word is not None else 0
        if merged is not None:
            bits |= int(merged.item())
        if bits:
            if self._err_word is not None:
                self._err_word.zero_()
            if self._is_synced:
                self.unsync()
            reasons = "; ".join(text for bit, text in _DATA_ERROR_TEXT.items() if bits & bit)
            raise RuntimeError(f"Invalid values detected while updating {self.__class__.__name__}: {reasons}.")

    _ERR_BITS = 4  # number of distinct data-error bits (_DATA_ERROR_TEXT)

    def _errors_travel_with_sync(self) -> bool:
        """True when this ``compute()`` will run the packed state sync and that sync has room for the error bits:
        an int64 "sum" tensor state (the bits ride in the packed all-reduce) or a "cat" list state (they ride in its
        size exchange).  Decided from the state declarations alone, so every rank decides alike."""
        if not (self._to_sync and not self._is_synced and self.dist_sync_fn is None):
            return False
        if not (callable(self.distributed_available_fn) and self.distributed_available_fn()):
            return False
        return self._error_carrier() is not None

    def _error_carrier(self) -> Optional[str]:
        for n, fn in self._reductions.items():
            v = getattr(self, n)
            if fn == dim_zero_sum and isinstance(v, Tensor) and v.dtype == torch.int64:
                return "sum"
        for n, fn in self._reductions.items():
            if fn == dim_zero_cat and isinstance(getattr(self, n), list):
                return "cat"
        return None

    def _error_bits_vector(self, device: torch.device) -> Tensor:
        """int64[_ERR_BITS]: one 0/1 entry per error bit of this rank (summed / maxed over ranks by the sync)."""
        shifts = torch.arange(self._ERR_BITS, device=device)
        if self._err_word is None or self._err_word.device != device:
            return torch.zeros(self._ERR_BITS, dtype=torch.int64, device=device)
        return (self._err_word.to(torch.int64) >> shifts) & 1

    # ------------------------------------------------------------------ state registry
    @property
    def update_called(self) -> bool:
        """``True`` once ``update`` or ``forward`` ran since construction / the last ``reset``."""
        return self._update_count > 0

    @property
    def update_count(self) -> int:
        return self._update_count

    @property
    def metric_state(self) -> Dict[str, Union[List[Tensor], Tensor]]:
        return {name: getattr(self, name) for name in self._defaults}

    @property
    def device(self) -> "torch.device":
        return self._device

    def add_state(
        self,
        name: str,
        default: Union[list, Tensor],
        dist_reduce_fx: Optional[Union[str, Callable]] = None,
        persistent: bool = False,
    ) -> None:
        """Register a state: a tensor, or an empty list that ``update`` appends tensors to.

        ``dist_reduce_fx`` is one of ``"sum" | "mean" | "cat" | "min" | "max"``, a callable applied to the
        ``[world, ...]`` stack, or ``None`` (keep the stack).
        """
        if not isinstance(default, (Tensor, list)) or (isinstance(default, list) and default):
            raise ValueError("state variable must be a tensor or any empty list (where you can append tensors)")
        if isinstance(dist_reduce_fx, str) and dist_reduce_fx in _NAMED_REDUCTIONS:
            dist_reduce_fx = _NAMED_REDUCTIONS[dist_reduce_fx]
        elif dist_reduce_fx is not None and not callable(dist_reduce_fx):
            raise ValueError("`dist_reduce_fx` must be callable or one of ['mean', 'sum', 'cat', 'min', 'max', None]")
        if isinstance(default, Tensor):
            default = default.contiguous()
        setattr(self, name, default)
        self._defaults[name] = deepcopy(default)
        self._persistent[name] = persistent
        self._reductions[name] = dist_reduce_fx

    # ------------------------------------------------------------------ forward
    @torch.jit.unused
    def forward(self, *args: Any, **kwargs: Any) -> Any:
        """Accumulate the batch globally and return the metric value of this batch alone."""
        if self._is_synced:
            raise TorchMetricsUserError(
                "The Metric shouldn't be synced when performing ``forward``. HINT: Did you forget to call ``unsync`` ?."
            )
        if self.full_state_update or self.full_state_update is None or self.dist_sync_on_step:
            self._forward_cache = self._forward_full_state_update(*args, **kwargs)
        else:
            self._forward_cache = self._forward_reduce_state_update(*args, **kwargs)
        return self._forward_cache

    @contextmanager
    def _batch_mode(self) -> Generator:
        """Temporarily switch sync / unsync / compute_on_cpu / grad flags to their per-batch values."""
        keep_cpu = self.compute_on_cpu
        self._to_sync = self.dist_sync_on_step
        self._should_unsync = False
        self.compute_on_cpu = False
        self._enable_grad = True
        self.__dict__["_in_batch"] = True
        try:
            yield
        finally:
            self.__dict__["_in_batch"] = False
            self._is_synced = False
            self._should_unsync = True
            self._to_sync = self.sync_on_compute
            self._computed = None
            self._enable_grad = False
            self.compute_on_cpu = keep_cpu
            if keep_cpu:
                self._move_list_states_to_cpu()

    def _forward_full_state_update(self, *args: Any, **kwargs: Any) -> Any:
        """Safe strategy: update the global state, then recompute on a scratch copy for the batch value."""
        self.update(*args, **kwargs)
        seen = self._update_count
        with self._batch_mode():
            saved = {name: getattr(self, name) for name in self._defaults}
            self._reset_keeping_errors()
            self.update(*args, **kwargs)
            batch_val = self.compute()
            for name, val in saved.items():
                setattr(self, name, val)
            self._update_count = seen
        return batch_val

    def _reset_keeping_errors(self) -> None:
        self.__dict__["_in_forward_reset"] = True
        try:
            self.reset()
        finally:
            self.__dict__["_in_forward_reset"] = False

    def _forward_sum_states(self, *args: Any, **kwargs: Any) -> Any:
        """``forward`` for metrics whose states are all sum-reduced tensors of one dtype (every K1 / K2 metric): the
        other half of SURVEY §8 row f4.  The reference resets (one fill per state), updates, computes and then adds
        global and batch state into NEW tensors (one add per state, metric.py:352-390).  Here the batch states are
        views of ONE zero-filled buffer, the kernels accumulate into them, and one fused ``_foreach_add_`` folds them
        into the global tensors in place — which therefore keep their identity (compute groups stay linked)."""
        names = list(self._defaults)
        glob = [getattr(self, n) for n in names]
        flat = torch.zeros(sum(g.numel() for g in glob), dtype=glob[0].dtype, device=glob[0].device)
        batch, at = [], 0
        for g in glob:
            batch.append(flat[at : at + g.numel()].view(g.shape))
            at += g.numel()
        seen, done = self._update_count, False
        try:
            for n, b in zip(names, batch):
                setattr(self, n, b)
            with self._batch_mode():
                self._computed = None
                self.update(*args, **kwargs)
                batch_val = self.compute()
                with torch.no_grad():
                    torch._foreach_add_(glob, batch)
            done = True
        finally:  # a failing update leaves the global state and the counter as they were
            for n, g in zip(names, glob):
                setattr(self, n, g)
            self._update_count = seen + 1 if done else seen
        return batch_val

    def _all_sum_tensor_states(self) -> bool:
        states = [getattr(self, n) for n in self._defaults]
        return (
            len(states) > 0
            and all(isinstance(s, Tensor) for s in states)
            and all(self._reductions[n] == dim_zero_sum for n in self._defaults)
            and len({(s.dtype, s.device) for s in states}) == 1
        )

    def _forward_reduce_state_update(self, *args: Any, **kwargs: Any) -> Any:
        """Fast strategy (``full_state_update=False``): batch state from scratch, then merge into the global one."""
        if self._all_sum_tensor_states():
            return self._forward_sum_states(*args, **kwargs)
        global_state = {name: getattr(self, name) for name in self._defaults}
        seen = self._update_count
        self._reset_keeping_errors()
        with self._batch_mode():
            self.update(*args, **kwargs)
            batch_val = self.compute()
            self._update_count = seen + 1
            with torch.no_grad():
                self._reduce_states(global_state)
        return batch_val

    def _reduce_states(self, incoming_state: Dict[str, Any]) -> None:
        """Merge ``incoming_state`` (the global state so far) with the current (batch) state."""
        for name in self._defaults:
            local = getattr(self, name)
            glob = incoming_state[name]
            fn = self._reductions[name]
            if fn == dim_zero_sum:
                merged = glob + local
            elif fn == dim_zero_mean:
                merged = ((self._update_count - 1) * glob + local).float() / self._update_count
            elif fn == dim_zero_max:
                merged = torch.max(glob, local)
            elif fn == dim_zero_min:
                merged = torch.min(glob, local)
            elif fn == dim_zero_cat:
                merged = glob + local
            elif fn is None and isinstance(glob, Tensor):
                merged = torch.stack([glob, local])
            elif fn is None and isinstance(glob, list):
                merged = _flatten([glob, local])
            elif fn and callable(fn):
                merged = fn(torch.stack([glob, local]))
            else:
                raise TypeError(f"Unsupported reduce_fn: {fn}")
            setattr(self, name, merged)

    # ------------------------------------------------------------------ distributed sync
    def _sync_dist(self, dist_sync_fn: Optional[Callable] = None, process_group: Optional[Any] = None) -> None:
        """Replace every state by its cross-rank reduction.

        ``dist_sync_fn=None`` takes the packed path; a callable takes the reference's generic path
        (called once per state tensor with ``group=...``; metric.py:431-436).
        """
        group = process_group or self.process_group
        if dist_sync_fn is None:
            self._sync_dist_packed(group)
        else:
            self._sync_dist_generic(dist_sync_fn, group)

    def _sync_dist_generic(self, dist_sync_fn: Callable, group: Optional[Any]) -> None:
        inputs = {name: getattr(self, name) for name in self._reductions}
        for name, fn in self._reductions.items():
            # fewer collectives: one tensor per list state
            if fn == dim_zero_cat and isinstance(inputs[name], list) and len(inputs[name]) > 1:
                inputs[name] = [dim_zero_cat(inputs[name])]
        gathered = apply_to_collection(inputs, Tensor, dist_sync_fn, group=group)
        for name, fn in self._reductions.items():
            out = gathered[name]
            if isinstance(out, list) and len(out) == 0:
                setattr(self, name, [])
                continue
            if isinstance(out[0], Tensor):
                out = torch.stack(out)
            elif isinstance(out[0], list):
                out = _flatten(out)
            if not (callable(fn) or fn is None):
                raise TypeError("reduction_fn must be callable or None")
            setattr(self, name, fn(out) if fn is not None else out)

    def _sync_dist_packed(self, group: Optional[Any]) -> None:
        """One all-reduce for every "sum" tensor state, one all-gather per "cat" list state.

        Anything else (mean/min/max/None/custom reductions, tensor states with "cat") goes through the
        generic gather so that the API stays complete.
        """
        sum_names = [
            n for n, fn in self._reductions.items() if fn == dim_zero_sum and isinstance(getattr(self, n), Tensor)
        ]
        cat_names = [
            n for n, fn in self._reductions.items() if fn == dim_zero_cat and isinstance(getattr(self, n), list)
        ]
        rest = [n for n in self._reductions if n not in sum_names and n not in cat_names]

        carrier = self._error_carrier()
        if sum_names:
            tensors = [getattr(self, n) for n in sum_names]
            if carrier == "sum":  # the error bits of this rank ride in the same all-reduce
                dev = next(t.device for t in tensors if t.dtype == torch.int64)
                tensors.append(self._error_bits_vector(dev))
            reduced = packed_all_reduce_sum(tensors, group)
            if carrier == "sum":
                counts = reduced.pop()
                self.__dict__["_err_merged"] = ((counts > 0).to(torch.int64) << torch.arange(
                    self._ERR_BITS, device=counts.device)).sum()
            for n, t in zip(sum_names, reduced):
                setattr(self, n, t)
        for i, n in enumerate(cat_names):
            parts = getattr(self, n)
            local = dim_zero_cat(parts) if parts else None
            device = local.device if local is not None else self._device
            if carrier == "cat" and i == 0:  # the error word rides in the size exchange of the first list state
                word = int(self._err_word.item()) if self._err_word is not None else 0
                merged, err_or = all_gather_cat(local, device, group, extra_word=word)
                self.__dict__["_err_merged"] = torch.tensor(err_or)
            else:
                merged = all_gather_cat(local, device, group)
            setattr(self, n, [] if merged is None else merged)
        if rest:
            keep = self._reductions
            try:
                self._reductions = {n: keep[n] for n in rest}
                self._sync_dist_generic(gather_all_tensors, group)
            finally:
                self._reductions = keep

    def sync(
        self,
        dist_sync_fn: Optional[Callable] = None,
        process_group: Optional[Any] = None,
        should_sync: bool = True,
        distributed_available: Optional[Callable] = None,
    ) -> None:
        """Synchronise the states across ranks; the local values are cached for ``unsync``."""
        if self._is_synced and should_sync:
            raise TorchMetricsUserError("The Metric has already been synced.")
        if distributed_available is None and self.distributed_available_fn is not None:
            distributed_available = self.distributed_available_fn
        is_distributed = distributed_available() if callable(distributed_available) else None
        if not should_sync or not is_distributed:
            return
        self._cache = {name: getattr(self, name) for name in self._defaults}
        self._sync_dist(dist_sync_fn, process_group=process_group)
        self._is_synced = True

    def unsync(self, should_unsync: bool = True) -> None:
        """Restore the local (pre-sync) states."""
        if not should_unsync:
            return
        if not self._is_synced:
            raise TorchMetricsUserError("The Metric has already been un-synced.")
        if self._cache is None:
            raise TorchMetricsUserError("The internal cache should exist to unsync the Metric.")
        for name, val in self._cache.items():
            setattr(self, name, val)
        self._is_synced = False
        self._cache = None

    @contextmanager
    def sync_context(
        self,
        dist_sync_fn: Optional[Callable] = None,
        process_group: Optional[Any] = None,
        should_sync: bool = True,
        should_unsync: bool = True,
        distributed_available: Optional[Callable] = None,
    ) -> Generator:
        """``sync`` on entry, ``unsync`` on exit (if a sync actually happened)."""
        self.sync(
            dist_sync_fn=dist_sync_fn,
            process_group=process_group,
            should_sync=should_sync,
            distributed_available=distributed_available,
        )
        yield
        self.unsync(should_unsync=self._is_synced and should_unsync)

    # ------------------------------------------------------------------ housekeeping
    def _move_list_states_to_cpu(self) -> None:
        for name in self._defaults:
            val = getattr(self, name)
            if isinstance(val, Sequence):
                setattr(self, name, [v.to("cpu") for v in val])

    def reset(self) -> None:
        """Back to the registered defaults (tensor states keep their current device)."""
        self._update_count = 0
        self._forward_cache = None
        self._computed = None
        for name, default in self._defaults.items():
            current = getattr(self, name)
            if isinstance(default, Tensor):
                setattr(self, name, default.detach().clone().to(current.device))
            else:
                setattr(self, name, [])
        self._cache = None
        self._is_synced = False
        # data errors recorded for samples that are no longer in the state must not surface after a reset (the
        # resets `forward` does around a batch keep them: those samples are still in the global state)
        if self._err_word is not None and not self.__dict__.get("_in_forward_reset", False):
            self._err_word.zero_()

    def clone(self) -> "Metric":
        return deepcopy(self)

    def __getstate__(self) -> Dict[str, Any]:
        # the wrapped bound methods cannot be pickled; scratch buffers are re-created on demand
        state = {k: v for k, v in self.__dict__.items() if k not in _UNPICKLED}
        state["_scratch"] = {}
        return state

    def __setstate__(self, state: Dict[str, Any]) -> None:
        self.__dict__.update(state)
        self._install_wrappers()

    def __setattr__(self, name: str, value: Any) -> None:
        if name in _CONSTANTS:
            raise RuntimeError(f"Can't change const `{name}`.")
        super().__setattr__(name, value)

    # dtype casts through the nn.Module shortcuts are ignored on purpose; use set_dtype
    def type(self, dst_type: Union[str, torch.dtype]) -> "Metric":  # noqa: A003
        return self

    def float(self) -> "Metric":  # noqa: A003
        return self

    def double(self) -> "Metric":
        return self

    def half(self) -> "Metric":
        return self

    def set_dtype(self, dst_type: Union[str, torch.dtype]) -> "Metric":
        """Convert floating point states (and defaults) to ``dst_type``."""
        self._dtype_convert = True
        out = super().type(dst_type)
        out._dtype_convert = False
        return out

    def _apply(self, fn: Callable, exclude_state: Sequence[str] = "") -> Module:
        """Move states, defaults and cached results together with the module (``.to()``, ``.cuda()`` ...)."""
        this = super()._apply(fn)
        fn_name = str(fn)
        is_cast = any(
            f in fn_name for f in ("Module.type", "Module.half", "Module.float", "Module.double", "Module.bfloat16")
        )
        if is_cast and not self._dtype_convert:
            return this
        for name, default in this._defaults.items():
            if name in exclude_state:
                continue
            if isinstance(default, Tensor):
                this._defaults[name] = fn(default)
            elif isinstance(default, Sequence):
                this._defaults[name] = [fn(v) for v in default]
            current = getattr(this, name)
            if isinstance(current, Tensor):
                setattr(this, name, fn(current))
            elif isinstance(current, Sequence):
                setattr(this, name, [fn(v) for v in current])
            else:
                raise TypeError(
                    f"Expected metric state to be either a Tensor or a list of Tensor, but encountered {current}"
                )
        self._device = fn(torch.zeros(1, device=self.device)).device
        if this._computed is not None:
            this._computed = apply_to_collection(this._computed, Tensor, fn)
        if this._forward_cache is not None:
            this._forward_cache = apply_to_collection(this._forward_cache, Tensor, fn)
        # device-side plumbing is re-created lazily on the new device
        this._err_word = None
        this._scratch = {}
        return this

    def persistent(self, mode: bool = False) -> None:
        """Switch whether the states are part of ``state_dict``."""
        for name in self._persistent:
            self._persistent[name] = mode

    def state_dict(  # type: ignore[override]
        self,
        destination: Optional[Dict[str, Any]] = None,
        prefix: str = "",
        keep_vars: bool = False,
    ) -> Dict[str, Any]:
        destination = super().state_dict(destination=destination, prefix=prefix, keep_vars=keep_vars)  # type: ignore[arg-type]
        for name in self._defaults:
            if not self._persistent[name]:
                continue
            val = getattr(self, name)
            if not keep_vars:
                if isinstance(val, Tensor):
                    val = val.detach()
                elif isinstance(val, list):
                    val = [v.detach() if isinstance(v, Tensor) else v for v in val]
            destination[prefix + name] = deepcopy(val)
        return destination

    def _load_from_state_dict(
        self,
        state_dict: dict,
        prefix: str,
        local_metadata: dict,
        strict: bool,
        missing_keys: List[str],
        unexpected_keys: List[str],
        error_msgs: List[str],
    ) -> None:
        for name in self._defaults:
            key = prefix + name
            if key in state_dict:
                setattr(self, name, state_dict.pop(key))
        super()._load_from_state_dict(state_dict, prefix, local_metadata, True, missing_keys, unexpected_keys, error_msgs)

    def _filter_kwargs(self, **kwargs: Any) -> Dict[str, Any]:
        """Keep only the keyword arguments that ``update`` accepts (everything if it takes ``**kwargs``)."""
        variadic = (inspect.Parameter.VAR_POSITIONAL, inspect.Parameter.VAR_KEYWORD)
        params = self._update_signature.parameters
        if any(p.kind == inspect.Parameter.VAR_KEYWORD for p in params.values()):
            return kwargs
        return {k: v for k, v in kwargs.items() if k in params and params[k].kind not in variadic}

    def __hash__(self) -> int:
        # id() keeps two instances apart even when their states are equal (nn.Module child discovery needs that)
        parts: List[Any] = [self.__class__.__name__, id(self)]
        for name in self._defaults:
            val = getattr(self, name)
            if hasattr(val, "__iter__") and not isinstance(val, Tensor):
                parts.extend(val)
            else:
                parts.append(val)
        return hash(tuple(parts))

    __iter__ = None
