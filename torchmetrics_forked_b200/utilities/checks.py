"""Input checks that do not touch tensor contents (reference: utilities/checks.py:39-44)."""
from torch import Tensor


def _check_same_shape(preds: Tensor, target: Tensor) -> None:
    """Raise ``RuntimeError`` unless both tensors have the same shape."""
    if preds.shape != target.shape:
        raise RuntimeError(
            f"Predictions and targets are expected to have the same shape, but got {preds.shape} and {target.shape}."
        )


def _allclose_recursive(res1, res2, atol: float = 1e-6) -> bool:
    """Are two metric results equal within ``atol``?  Results are tensors, strings, or sequences / mappings of results
    (reference: utilities/checks.py:621-632; its own tests import it)."""
    if isinstance(res1, Tensor):
        return bool(res1.shape == res2.shape and (res1 - res2).abs().le(atol).all()) if res1.is_floating_point() else bool(
            (res1 == res2).all()
        )
    if isinstance(res1, str):
        return res1 == res2
    if isinstance(res1, dict):
        return all(_allclose_recursive(v, res2[k], atol) for k, v in res1.items())
    if isinstance(res1, (list, tuple)):
        return len(res1) == len(res2) and all(_allclose_recursive(a, b, atol) for a, b in zip(res1, res2))
    return res1 == res2
