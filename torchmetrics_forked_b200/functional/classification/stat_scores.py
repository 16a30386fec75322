"""True/false positive/negative counts for binary, multiclass and multilabel tasks.

Reference: functional/classification/stat_scores.py.  ``_format`` + ``_update`` of the reference are one
fused kernel per task that adds into the int64 ``tp/fp/tn/fn`` states in place:

* multiclass (global, ``top_k == 1``): argmax + three C-bin histograms (target, pred, match) — the
  reference builds a full C x C confusion matrix and reduces it (:405-418); the identities are
  ``tp = match``, ``fp = hist_pred - match``, ``fn = hist_target - match``, ``tn = n - fp - fn - tp``;
  ``average="micro"`` only counts matches (:394-404);
* binary / multilabel: conditional sigmoid + strict ``> threshold`` + four counters per label (:90-131,
  :647-681).

``top_k > 1`` (:363-393) goes through the top-k selection kernels (row f3), ``multidim_average="samplewise"``
through per-sample histogram blocks, and their combination through the per-sample top-k kernel.
"""
from typing import Optional, Tuple

import torch
from torch import Tensor
from typing_extensions import Literal

from torchmetrics_forked_b200 import kernels
from torchmetrics_forked_b200.functional.classification.confusion_matrix import (
    _check_ignore_index,
    _multiclass_hist_update_,
    _multiclass_shape_validation,
)
from torchmetrics_forked_b200.utilities.checks import _check_same_shape
from torchmetrics_forked_b200.utilities.enums import ClassificationTask

_AVERAGES = ("micro", "macro", "weighted", "none", None)
_MULTIDIM = ("global", "samplewise")


def _check_multidim(multidim_average: str) -> None:
    if multidim_average not in _MULTIDIM:
        raise ValueError(f"Expected argument `multidim_average` to be one of {_MULTIDIM}, but got {multidim_average}")


def _require_global(multidim_average: str, top_k: int = 1) -> None:
    """Kept for the callers' sake: every combination of ``multidim_average`` and ``top_k`` of the reference's one-hot
    branch (:363-393) is built (``samplewise`` + ``top_k > 1`` through ``tmb_mc_topk_hist_samplewise``)."""


def _samplewise_thresh_stats(
    preds: Tensor, target: Tensor, rows: int, threshold: float, ignore_index: Optional[int], err: Tensor
) -> Tuple[Tensor, Tensor, Tensor, Tensor]:
    """Binary / multilabel ``multidim_average="samplewise"`` (reference :116-129, :672-679 with ``sum_dim = 1``):
    every (sample[, label]) pair is one "label" of the single-threshold kernel — ``rows`` of them, each owning the
    trailing positions — so one launch returns the four ``[rows]`` count vectors.  The sigmoid test still looks at
    the whole batch, as the reference's ``_format`` does."""
    dev = preds.device
    outs = tuple(torch.zeros(rows, dtype=torch.long, device=dev) for _ in range(4))
    if preds.numel() == 0:
        return outs
    p, t = preds.reshape(1, rows, -1), target.reshape(1, rows, -1)
    scratch = torch.zeros(kernels.thresh_count_scratch_numel(rows), dtype=torch.long, device=dev)
    kernels.thresh_count(
        p, t, num_labels=rows, threshold=threshold, ignore_index=ignore_index, flag_counts_ignored=True,
        layout=kernels.TC_STATS, scratch=scratch, out=outs, err=err, binary=False,
    )  # fmt: skip
    return outs


def _stack_stats(
    tp: Tensor, fp: Tensor, tn: Tensor, fn: Tensor, average: Optional[str], micro_sums: bool,
    multidim_average: str = "global",
) -> Tensor:  # fmt: skip
    """``[tp, fp, tn, fn, support]`` stacked on the last dim, then averaged over classes / labels (dim 0, or dim 1 of
    the per-sample ``[N, C]`` counts)."""
    res = torch.stack([tp, fp, tn, fn, tp + fn], dim=-1)
    dim = 0 if multidim_average == "global" else 1
    if average == "micro":
        return res.sum(dim) if (micro_sums or res.ndim > 1) else res
    if average == "macro":
        return res.float().mean(dim)
    if average == "weighted":
        weight = tp + fn
        # multilabel (micro_sums) normalises by the grand total even per sample (reference :704), multiclass by the
        # sample's own total (:443-445)
        total = weight.sum() if (multidim_average == "global" or micro_sums) else weight.sum(-1, keepdim=True)
        return (res * (weight / total).reshape(*weight.shape, 1)).sum(dim)
    if average is None or average == "none":
        return res
    return None  # type: ignore[return-value]


# ------------------------------------------------------------------------------------------ binary
def _binary_stat_scores_arg_validation(
    threshold: float = 0.5, multidim_average: str = "global", ignore_index: Optional[int] = None
) -> None:
    """Reference: stat_scores.py:25-46."""
    if not (isinstance(threshold, float) and (0 <= threshold <= 1)):
        raise ValueError(f"Expected argument `threshold` to be a float in the [0,1] range, but got {threshold}.")
    _check_multidim(multidim_average)
    _check_ignore_index(ignore_index)


def _binary_stat_scores_tensor_validation(
    preds: Tensor, target: Tensor, multidim_average: str = "global", ignore_index: Optional[int] = None
) -> None:
    """Shape checks of stat_scores.py:63-64, 86-87; the {0,1} content checks run in the kernel."""
    _check_same_shape(preds, target)
    if multidim_average != "global" and preds.ndim < 2:
        raise ValueError("Expected input to be at least 2D when multidim_average is set to `samplewise`")


def _binary_stat_scores_update_(
    preds: Tensor, target: Tensor, threshold: float, ignore_index: Optional[int],
    tp: Tensor, fp: Tensor, tn: Tensor, fn: Tensor, scratch: Tensor, err: Tensor,
) -> None:  # fmt: skip
    """In-place ``tp/fp/tn/fn += counts`` (reference ``_format`` + ``_update``, :90-131).

    The sigmoid test runs over all elements, ignored ones included (:103-107, H2).
    """
    kernels.thresh_count(
        preds, target, num_labels=1, threshold=threshold, ignore_index=ignore_index, flag_counts_ignored=True,
        layout=kernels.TC_STATS, scratch=scratch, out=(tp, fp, tn, fn), err=err, binary=True,
    )  # fmt: skip


def _binary_stat_scores_compute(tp: Tensor, fp: Tensor, tn: Tensor, fn: Tensor, multidim_average: str = "global") -> Tensor:
    """Reference: stat_scores.py:134-138."""
    return torch.stack([tp, fp, tn, fn, tp + fn], dim=0 if multidim_average == "global" else 1).squeeze()


def binary_stat_scores(
    preds: Tensor,
    target: Tensor,
    threshold: float = 0.5,
    multidim_average: Literal["global", "samplewise"] = "global",
    ignore_index: Optional[int] = None,
    validate_args: bool = True,
) -> Tensor:
    """``[tp, fp, tn, fn, support]`` of a binary task (reference :141-212)."""
    if validate_args:
        _binary_stat_scores_arg_validation(threshold, multidim_average, ignore_index)
        _binary_stat_scores_tensor_validation(preds, target, multidim_average, ignore_index)
    tp, fp, tn, fn = _binary_stat_scores_functional_state(
        preds, target, threshold, ignore_index, validate_args, "binary_stat_scores", multidim_average
    )
    return _binary_stat_scores_compute(tp, fp, tn, fn, multidim_average)


# ------------------------------------------------------------------------------------------ multiclass
def _multiclass_stat_scores_arg_validation(
    num_classes: int,
    top_k: int = 1,
    average: Optional[str] = "macro",
    multidim_average: str = "global",
    ignore_index: Optional[int] = None,
) -> None:
    """Reference: stat_scores.py:215-248."""
    if not isinstance(num_classes, int) or num_classes < 2:
        raise ValueError(f"Expected argument `num_classes` to be an integer larger than 1, but got {num_classes}")
    if not isinstance(top_k, int) and top_k < 1:
        raise ValueError(f"Expected argument `top_k` to be an integer larger than or equal to 1, but got {top_k}")
    if top_k > num_classes:
        raise ValueError(
            f"Expected argument `top_k` to be smaller or equal to `num_classes` but got {top_k} and {num_classes}"
        )
    if average not in _AVERAGES:
        raise ValueError(f"Expected argument `average` to be one of {_AVERAGES}, but got {average}")
    _check_multidim(multidim_average)
    _check_ignore_index(ignore_index)


def _multiclass_stat_scores_tensor_validation(
    preds: Tensor, target: Tensor, num_classes: int, multidim_average: str = "global", ignore_index: Optional[int] = None
) -> None:
    """Shape checks of stat_scores.py:270-304; label-range checks run in the kernel."""
    _multiclass_shape_validation(preds, target, num_classes)
    if multidim_average != "global":
        if preds.ndim == target.ndim + 1 and preds.ndim < 3:
            raise ValueError(
                "If `preds` have one dimension more than `target`, the shape of `preds` should "
                " at least 3D when multidim_average is set to `samplewise`"
            )
        if preds.ndim == target.ndim and preds.ndim < 2:
            raise ValueError(
                "When `preds` and `target` have the same shape, the shape of `preds` should "
                " at least 2D when multidim_average is set to `samplewise`"
            )


def _multiclass_stats_mode(average: Optional[str], top_k: int = 1) -> int:
    """Scalar match counters for ``micro`` with ``top_k == 1`` (reference :394-403), per-class histograms otherwise
    (``top_k > 1`` keeps per-class states even for ``micro``, classification/stat_scores.py:318-320)."""
    return kernels.MC_MICRO if (average == "micro" and top_k == 1) else kernels.MC_STATS


def _multiclass_stat_scores_update_(
    preds: Tensor, target: Tensor, num_classes: int, average: Optional[str], ignore_index: Optional[int],
    tp: Tensor, fp: Tensor, tn: Tensor, fn: Tensor, scratch: Tensor, err: Tensor, top_k: int = 1,
) -> None:  # fmt: skip
    """In-place ``tp/fp/tn/fn += counts``, global averaging (reference :325-341, :394-418; ``top_k > 1``:
    :363-393 through the top-k selection kernel)."""
    mode = _multiclass_stats_mode(average, top_k)
    if top_k > 1:
        kernels.mc_topk_hist(preds, target, num_classes, top_k, ignore_index, scratch, err)
    else:
        _multiclass_hist_update_(preds, target, num_classes, ignore_index, mode, scratch, err)
    kernels.mc_stats_finalize(scratch, num_classes, mode, tp, fp, tn, fn)


def _multiclass_stat_scores_compute(
    tp: Tensor, fp: Tensor, tn: Tensor, fn: Tensor, average: Optional[str] = "macro", multidim_average: str = "global"
) -> Tensor:
    """Reference: stat_scores.py:421-448."""
    return _stack_stats(tp, fp, tn, fn, average, micro_sums=False, multidim_average=multidim_average)


def multiclass_stat_scores(
    preds: Tensor,
    target: Tensor,
    num_classes: int,
    average: Optional[Literal["micro", "macro", "weighted", "none"]] = "macro",
    top_k: int = 1,
    multidim_average: Literal["global", "samplewise"] = "global",
    ignore_index: Optional[int] = None,
    validate_args: bool = True,
) -> Tensor:
    """``[tp, fp, tn, fn, support]`` per class or averaged (reference :451-563)."""
    if validate_args:
        _multiclass_stat_scores_arg_validation(num_classes, top_k, average, multidim_average, ignore_index)
        _multiclass_stat_scores_tensor_validation(preds, target, num_classes, multidim_average, ignore_index)
    _require_global(multidim_average, top_k)
    tp, fp, tn, fn = _multiclass_stat_scores_functional_state(preds, target, num_classes, average, ignore_index,
                                                              validate_args, "multiclass_stat_scores", top_k,
                                                              multidim_average=multidim_average,
)
    return _multiclass_stat_scores_compute(tp, fp, tn, fn, average, multidim_average)


def _multiclass_stat_scores_functional_state(
    preds: Tensor, target: Tensor, num_classes: int, average: Optional[str], ignore_index: Optional[int],
    validate_args: bool, where: str, top_k: int = 1, multidim_average: str = "global",
) -> Tuple[Tensor, Tensor, Tensor, Tensor]:  # fmt: skip
    """Fresh zero states + one fused update; shared by the functional accuracy / f-beta entry points."""
    dev = preds.device
    if multidim_average == "samplewise":
        # per-sample, per-class counts [N, C]: the reference's one-hot branch (:363-393) with sum_dim = [1]
        err = torch.zeros(1, dtype=torch.int32, device=dev)
        out = kernels.mc_stats_samplewise(preds, target, num_classes, ignore_index, err, top_k)
        if validate_args:
            kernels.raise_on_error_word(err, where)
        return out
    mode = _multiclass_stats_mode(average, top_k)
    size = 1 if mode == kernels.MC_MICRO else num_classes
    tp, fp, tn, fn = (torch.zeros(size, dtype=torch.long, device=dev) for _ in range(4))
    scratch = torch.zeros(kernels.mc_scratch_numel(num_classes, mode), dtype=torch.long, device=dev)
    err = torch.zeros(1, dtype=torch.int32, device=dev)
    _multiclass_stat_scores_update_(
        preds, target, num_classes, average, ignore_index, tp, fp, tn, fn, scratch, err, top_k
    )
    if validate_args:
        kernels.raise_on_error_word(err, where)
    if mode == kernels.MC_MICRO:
        tp, fp, tn, fn = tp.squeeze(0), fp.squeeze(0), tn.squeeze(0), fn.squeeze(0)
    return tp, fp, tn, fn


# ------------------------------------------------------------------------------------------ multilabel
def _multilabel_stat_scores_arg_validation(
    num_labels: int,
    threshold: float = 0.5,
    average: Optional[str] = "macro",
    multidim_average: str = "global",
    ignore_index: Optional[int] = None,
) -> None:
    """Reference: stat_scores.py:566-592."""
    if not isinstance(num_labels, int) or num_labels < 2:
        raise ValueError(f"Expected argument `num_labels` to be an integer larger than 1, but got {num_labels}")
    if not (isinstance(threshold, float) and (0 <= threshold <= 1)):
        raise ValueError(f"Expected argument `threshold` to be a float, but got {threshold}.")
    if average not in _AVERAGES:
        raise ValueError(f"Expected argument `average` to be one of {_AVERAGES}, but got {average}")
    _check_multidim(multidim_average)
    _check_ignore_index(ignore_index)


def _multilabel_stat_scores_tensor_validation(
    preds: Tensor, target: Tensor, num_labels: int, multidim_average: str, ignore_index: Optional[int] = None
) -> None:
    """Shape checks of stat_scores.py:611-644; the {0,1} content checks run in the kernel."""
    _check_same_shape(preds, target)
    if preds.shape[1] != num_labels:
        raise ValueError(
            "Expected both `target.shape[1]` and `preds.shape[1]` to be equal to the number of labels"
            f" but got {preds.shape[1]} and expected {num_labels}"
        )
    if multidim_average != "global" and preds.ndim < 3:
        raise ValueError("Expected input to be at least 3D when multidim_average is set to `samplewise`")


def _multilabel_stat_scores_update_(
    preds: Tensor, target: Tensor, num_labels: int, threshold: float, ignore_index: Optional[int],
    tp: Tensor, fp: Tensor, tn: Tensor, fn: Tensor, scratch: Tensor, err: Tensor,
) -> None:  # fmt: skip
    """In-place per-label ``tp/fp/tn/fn += counts`` (reference ``_format`` + ``_update``, :647-681)."""
    kernels.thresh_count(
        preds, target, num_labels=num_labels, threshold=threshold, ignore_index=ignore_index,
        flag_counts_ignored=True, layout=kernels.TC_STATS, scratch=scratch, out=(tp, fp, tn, fn), err=err,
        binary=False,
    )  # fmt: skip


def _multilabel_stat_scores_compute(
    tp: Tensor, fp: Tensor, tn: Tensor, fn: Tensor, average: Optional[str] = "macro", multidim_average: str = "global"
) -> Tensor:
    """Reference: stat_scores.py:684-708."""
    return _stack_stats(tp, fp, tn, fn, average, micro_sums=True, multidim_average=multidim_average)


def multilabel_stat_scores(
    preds: Tensor,
    target: Tensor,
    num_labels: int,
    threshold: float = 0.5,
    average: Optional[Literal["micro", "macro", "weighted", "none"]] = "macro",
    multidim_average: Literal["global", "samplewise"] = "global",
    ignore_index: Optional[int] = None,
    validate_args: bool = True,
) -> Tensor:
    """``[tp, fp, tn, fn, support]`` per label or averaged (reference :711-822)."""
    if validate_args:
        _multilabel_stat_scores_arg_validation(num_labels, threshold, average, multidim_average, ignore_index)
        _multilabel_stat_scores_tensor_validation(preds, target, num_labels, multidim_average, ignore_index)
    _require_global(multidim_average)
    tp, fp, tn, fn = _multilabel_stat_scores_functional_state(preds, target, num_labels, threshold, ignore_index,
                                                              validate_args, "multilabel_stat_scores",
                                                              multidim_average=multidim_average,
)
    return _multilabel_stat_scores_compute(tp, fp, tn, fn, average, multidim_average)


def _multilabel_stat_scores_functional_state(
    preds: Tensor, target: Tensor, num_labels: int, threshold: float, ignore_index: Optional[int],
    validate_args: bool, where: str, multidim_average: str = "global",
) -> Tuple[Tensor, Tensor, Tensor, Tensor]:  # fmt: skip
    dev = preds.device
    err = torch.zeros(1, dtype=torch.int32, device=dev)
    if multidim_average == "samplewise":
        n = preds.shape[0]
        out = tuple(
            x.reshape(n, num_labels)
            for x in _samplewise_thresh_stats(preds, target, n * num_labels, threshold, ignore_index, err)
        )
    else:
        tp, fp, tn, fn = (torch.zeros(num_labels, dtype=torch.long, device=dev) for _ in range(4))
        scratch = torch.zeros(kernels.thresh_count_scratch_numel(num_labels), dtype=torch.long, device=dev)
        _multilabel_stat_scores_update_(preds, target, num_labels, threshold, ignore_index, tp, fp, tn, fn, scratch, err)
        out = (tp, fp, tn, fn)
    if validate_args:
        kernels.raise_on_error_word(err, where)
    return out


def _binary_stat_scores_functional_state(
    preds: Tensor, target: Tensor, threshold: float, ignore_index: Optional[int], validate_args: bool, where: str,
    multidim_average: str = "global",
) -> Tuple[Tensor, Tensor, Tensor, Tensor]:  # fmt: skip
    dev = preds.device
    err = torch.zeros(1, dtype=torch.int32, device=dev)
    if multidim_average == "samplewise":
        # `.squeeze()` like the reference (:127-130): a batch of one sample gives 0-d counts, and a 0-d result
        out = tuple(x.squeeze() for x in _samplewise_thresh_stats(preds, target, preds.shape[0], threshold, ignore_index, err))
    else:
        tp, fp, tn, fn = (torch.zeros(1, dtype=torch.long, device=dev) for _ in range(4))
        scratch = torch.zeros(kernels.thresh_count_scratch_numel(1), dtype=torch.long, device=dev)
        _binary_stat_scores_update_(preds, target, threshold, ignore_index, tp, fp, tn, fn, scratch, err)
        out = (tp.squeeze(), fp.squeeze(), tn.squeeze(), fn.squeeze())
    if validate_args:
        kernels.raise_on_error_word(err, where)
    return out


def stat_scores(
    preds: Tensor,
    target: Tensor,
    task: Literal["binary", "multiclass", "multilabel"],
    threshold: float = 0.5,
    num_classes: Optional[int] = None,
    num_labels: Optional[int] = None,
    average: Optional[Literal["micro", "macro", "weighted", "none"]] = "micro",
    multidim_average: Optional[Literal["global", "samplewise"]] = "global",
    top_k: Optional[int] = 1,
    ignore_index: Optional[int] = None,
    validate_args: bool = True,
) -> Tensor:
    """Task dispatcher (reference :825-1129)."""
    task = ClassificationTask.from_str(task)
    assert multidim_average is not None  # noqa: S101
    if task == ClassificationTask.BINARY:
        return binary_stat_scores(preds, target, threshold, multidim_average, ignore_index, validate_args)
    if task == ClassificationTask.MULTICLASS:
        if not isinstance(num_classes, int):
            raise ValueError(f"`num_classes` is expected to be `int` but `{type(num_classes)} was passed.`")
        if not isinstance(top_k, int):
            raise ValueError(f"`top_k` is expected to be `int` but `{type(top_k)} was passed.`")
        return multiclass_stat_scores(
            preds, target, num_classes, average, top_k, multidim_average, ignore_index, validate_args
        )
    if task == ClassificationTask.MULTILABEL:
        if not isinstance(num_labels, int):
            raise ValueError(f"`num_labels` is expected to be `int` but `{type(num_labels)} was passed.`")
        return multilabel_stat_scores(
            preds, target, num_labels, threshold, average, multidim_average, ignore_index, validate_args
        )
    raise ValueError(f"Unsupported task `{task}`")
