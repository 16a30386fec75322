"""float64 scores in front of the float32 / 16-bit kernels, exactly.

The kernels read float32, float16 and bfloat16 scores (the dtypes a model emits).  The reference also takes float64
``preds`` and then evaluates everything in float64: the range test, sigmoid / softmax, ``argmax``, the comparisons
with the thresholds, the sort.  This module is the front end for that dtype.  It does NOT round the scores and hope:
each helper hands the kernels something whose result is, by construction, the float64 result.

* labels (``preds > threshold``, ``argmax``, top-k) are decided in float64 by torch and the kernels count integer
  labels — the label entry points they already have;
* binned curves compare a float64 score ``p`` with float32 thresholds ``t`` (promoted to float64, so exactly):
  ``p >= t  <=>  rd(p) >= t`` where ``rd`` rounds toward minus infinity to float32, because ``t`` itself is a float32
  (``p > t  <=>  ru(p) > t`` likewise).  The normalisation runs in float64, the directed rounding follows, the kernels
  bin the rounded scores;
* exact curves only need the ORDER of the scores and their ties: every score is replaced by its dense rank among the
  distinct float64 values, handed to the radix sort as the bit pattern of a non-negative float32 (bit patterns of
  non-negative floats are ordered like their values), and the thresholds are looked up again from the rank.  NaN
  scores are passed on as NaN, for which the sort has its own rule.

It is a convenience path, not a measured one: torch's elementwise kernels and ``torch.unique`` do the float64 work,
some helpers read a flag back.  Nothing here runs for float32 / 16-bit scores.
"""
from typing import List, Optional, Sequence, Tuple

import torch
from torch import Tensor

_MAX_RANK = 0x7F800000  # bit patterns below this are finite non-negative float32 values


def is_f64(t: Tensor) -> bool:
    return isinstance(t, Tensor) and t.dtype == torch.float64


def _inside_everywhere(p: Tensor, keep: Optional[Tensor]) -> Tensor:
    """0-d bool: every element (that takes part) lies in [0, 1]; NaN counts as outside, like in the reference's
    ``(preds >= 0) * (preds <= 1)``."""
    inside = (p >= 0) & (p <= 1)
    if keep is not None:
        inside = inside | ~keep
    return inside.all()


def sigmoid_where_needed(preds: Tensor, target: Optional[Tensor], ignore_index: Optional[int]) -> Tensor:
    """``preds.sigmoid()`` when any element outside the ignored positions leaves [0, 1], else ``preds``: decided and
    applied on the device (both branches are evaluated, nothing is read back)."""
    keep = None if target is None or ignore_index is None else (target != ignore_index)
    return torch.where(_inside_everywhere(preds, keep), preds, preds.sigmoid())


def softmax_where_needed(rows: Tensor, target: Optional[Tensor], ignore_index: Optional[int]) -> Tensor:
    """Row softmax of ``[M, C]`` scores when any element of a row that is not ignored leaves [0, 1]."""
    keep = None if target is None or ignore_index is None else (target.reshape(-1, 1) != ignore_index)
    return torch.where(_inside_everywhere(rows, keep), rows, rows.softmax(1))


def round_directed(p: Tensor, up: bool) -> Tensor:
    """float64 -> float32 toward plus (``up``) or minus infinity."""
    q = p.to(torch.float32)
    back = q.to(torch.float64)
    if up:
        return torch.where(back < p, torch.nextafter(q, torch.full_like(q, float("inf"))), q)
    return torch.where(back > p, torch.nextafter(q, torch.full_like(q, float("-inf"))), q)


def scores_for_binning(p: Tensor, strict: bool) -> Tensor:
    """Normalised float64 scores as float32 values that compare with float32 thresholds like the originals
    (``>=`` -> round down, ``>`` -> round up).

    NaN scores are refused: a float32 NaN would make the kernel's own range test fire and normalise a second time.
    (The label metrics and the exact curves take float64 NaN scores; only this one combination does not.)"""
    if bool(torch.isnan(p).any()):
        raise ValueError(
            "float64 `preds` holding NaN cannot be binned against `thresholds`; pass float32 scores or drop the NaN"
        )
    return round_directed(p, up=strict)


def binary_labels(
    preds: Tensor, target: Tensor, threshold: float, ignore_index: Optional[int], test_every_element: bool
) -> Tensor:
    """``(sigmoid?)(preds) > threshold`` in float64 as int64 0/1 labels.  ``test_every_element``: the range test of
    the stat-scores family looks at ignored positions too, the confusion matrix's does not."""
    p = sigmoid_where_needed(preds, None if test_every_element else target, ignore_index)
    return (p > threshold).long()


def multiclass_labels(preds: Tensor) -> Tensor:
    """``argmax`` over dim 1 of ``[N, C, ...]`` float64 scores."""
    return preds.argmax(dim=1)


def topk_marks(preds: Tensor, top_k: int) -> Tensor:
    """float32 ``[N, C, ...]`` with 1.0 at the ``top_k`` largest float64 scores of every position and 0.0 elsewhere:
    the top-k kernels then select exactly those classes (1 > 0, no ties between a mark and a blank)."""
    idx = preds.topk(top_k, dim=1).indices
    return torch.zeros(preds.shape, dtype=torch.float32, device=preds.device).scatter_(1, idx, 1.0)


def rank_scores(chunks: Sequence[Tensor]) -> Tuple[Tensor, Tensor]:
    """float64 score chunks ``[n_i, L]`` -> (float32 ``[N, L]`` whose bit patterns are the dense ranks of the scores
    among all distinct values, the distinct values ascending).  Order and ties are those of the float64 scores.  NaN
    scores stay NaN: the sort kernels have their own rule for them (above all numbers, each a threshold of its own, in
    input order), the one that matches the reference's sort."""
    cat = chunks[0] if len(chunks) == 1 else torch.cat(list(chunks))
    values, rank = torch.unique(cat, return_inverse=True)
    if values.numel() >= _MAX_RANK:
        raise ValueError("too many distinct float64 scores for one exact curve computation")
    ranked = rank.to(torch.int32).view(torch.float32).reshape(cat.shape)
    return torch.where(cat.isnan(), torch.full_like(ranked, float("nan")), ranked), values


def scores_of_ranks(ranked: Tensor, values: Tensor) -> Tensor:
    """The float64 scores behind float32 rank patterns (the ``thresholds`` output of a curve)."""
    rank = ranked.contiguous().view(torch.int32).long()
    held = rank < values.numel()  # anything else is the NaN pattern
    found = values[torch.where(held, rank, torch.zeros_like(rank))]
    return torch.where(held, found, torch.full_like(found, float("nan")))


def split_like(whole: Tensor, chunks: Sequence[Tensor]) -> List[Tensor]:
    out, start = [], 0
    for c in chunks:
        out.append(whole[start : start + c.shape[0]])
        start += c.shape[0]
    return out
