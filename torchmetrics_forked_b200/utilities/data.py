"""State helpers: dim-zero reductions, collection mapping, and the integer histogram `_bincount`.

Reference: utilities/data.py:28-55 (dim_zero_*), :58-60 (_flatten), :160-166 (_squeeze_if_scalar),
:169-199 (_bincount), :230-237 (allclose).  `apply_to_collection` restates the one helper the reference
takes from the un-vendored `lightning_utilities` package (metric.py:26, :431-436).
"""
from collections import OrderedDict, defaultdict
from typing import Any, Callable, List, Optional, Sequence, Tuple, Union

import torch
from torch import Tensor

METRIC_EPS = 1e-6


def dim_zero_cat(x: Union[Tensor, List[Tensor]]) -> Tensor:
    """Concatenate a list state along dim 0 (0-d one-element tensors are promoted to 1-d first)."""
    if isinstance(x, Tensor):
        return x
    parts = [y.unsqueeze(0) if y.numel() == 1 and y.ndim == 0 else y for y in x]
    if not parts:
        raise ValueError("No samples to concatenate")
    return torch.cat(parts, dim=0)


def dim_zero_sum(x: Tensor) -> Tensor:
    return torch.sum(x, dim=0)


def dim_zero_mean(x: Tensor) -> Tensor:
    return torch.mean(x, dim=0)


def dim_zero_max(x: Tensor) -> Tensor:
    return torch.max(x, dim=0).values


def dim_zero_min(x: Tensor) -> Tensor:
    return torch.min(x, dim=0).values


def _flatten(x: Sequence) -> list:
    """[[a, b], [c]] -> [a, b, c]."""
    return [item for sub in x for item in sub]


def apply_to_collection(data: Any, dtype: Union[type, Tuple[type, ...]], function: Callable, *args: Any, **kwargs: Any) -> Any:
    """Apply ``function`` to every leaf of type ``dtype`` inside nested dicts / lists / tuples / namedtuples."""
    if isinstance(data, dtype):
        return function(data, *args, **kwargs)
    if isinstance(data, (dict, OrderedDict, defaultdict)):
        mapped = {k: apply_to_collection(v, dtype, function, *args, **kwargs) for k, v in data.items()}
        if isinstance(data, defaultdict):
            return defaultdict(data.default_factory, mapped)
        return type(data)(mapped)
    if isinstance(data, tuple) and hasattr(data, "_fields"):  # namedtuple
        return type(data)(*(apply_to_collection(v, dtype, function, *args, **kwargs) for v in data))
    if isinstance(data, (list, tuple)):
        return type(data)(apply_to_collection(v, dtype, function, *args, **kwargs) for v in data)
    return data


def _squeeze_if_scalar(data: Any) -> Any:
    """One-element tensors anywhere in ``data`` become 0-d."""
    return apply_to_collection(data, Tensor, lambda t: t.squeeze() if t.numel() == 1 else t)


def allclose(tensor1: Tensor, tensor2: Tensor) -> bool:
    """`torch.allclose` that tolerates a dtype difference (used when discovering compute groups)."""
    if tensor1.dtype != tensor2.dtype:
        tensor2 = tensor2.to(dtype=tensor1.dtype)
    return torch.allclose(tensor1, tensor2)


def _bincount(x: Tensor, minlength: Optional[int] = None) -> Tensor:
    """Count occurrences of each value of the non-negative integer tensor ``x``.

    CUDA tensors go through the `tmb_bincount` kernel (integer atomics, hence deterministic; the
    reference needs a Python loop in deterministic mode, utilities/data.py:194-198).  Like
    ``torch.bincount`` the result has ``max(minlength, x.max() + 1)`` entries and negative values raise.
    """
    from torchmetrics_forked_b200 import kernels

    if minlength is None:
        minlength = len(torch.unique(x))
    return kernels.bincount(x, int(minlength))
