"""Small float reducers applied to the (tiny) integer states at ``compute()`` time.

These stay as torch expressions on purpose: the states they read are a few kB, and evaluating the same
fp32 expressions the reference evaluates (utilities/compute.py:46-69, 88-96, 134-157) makes the derived
metrics agree to the last bit when both run on the same device.
"""
from typing import Optional

import torch
from torch import Tensor


def _safe_divide(num: Tensor, denom: Tensor) -> Tensor:
    """``num / denom`` in floating point with zero denominators replaced by one (in place, like the reference)."""
    denom[denom == 0.0] = 1
    num = num if num.is_floating_point() else num.float()
    denom = denom if denom.is_floating_point() else denom.float()
    return num / denom


def _adjust_weights_safe_divide(
    score: Tensor, average: Optional[str], multilabel: bool, tp: Tensor, fp: Tensor, fn: Tensor
) -> Tensor:
    """Macro / weighted averaging of a per-class score; classes that never occur get weight zero (multiclass)."""
    if average is None or average == "none":
        return score
    if average == "weighted":
        weights = tp + fn
    else:
        weights = torch.ones_like(score)
        if not multilabel:
            weights[tp + fp + fn == 0] = 0.0
    return _safe_divide(weights * score, weights.sum(-1, keepdim=True)).sum(-1)


def _auc_compute_without_check(x: Tensor, y: Tensor, direction: float, axis: int = -1) -> Tensor:
    """Trapezoidal area under ``y(x)``; ``direction`` is -1 for decreasing ``x``."""
    with torch.no_grad():
        return torch.trapz(y, x, dim=axis) * direction


def interp(x: Tensor, xp: Tensor, fp: Tensor) -> Tensor:
    """Piecewise-linear interpolation of the points ``(xp, fp)`` (``xp`` increasing) evaluated at ``x``."""
    slope = _safe_divide(fp[1:] - fp[:-1], xp[1:] - xp[:-1])
    offset = fp[:-1] - (slope * xp[:-1])
    seg = torch.sum(torch.ge(x[:, None], xp[None, :]), 1) - 1
    seg = torch.clamp(seg, 0, len(slope) - 1)
    return slope[seg] * x + offset[seg]
