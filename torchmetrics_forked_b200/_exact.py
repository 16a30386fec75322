"""Host-side launchers of the exact-curve kernels (K3): conditional sigmoid, ordered ignore-filter, batched
sort + scan, area reductions.  Re-exported through :mod:`torchmetrics_forked_b200.kernels`."""
from typing import List, Optional, Tuple

import torch
from torch import Tensor

from torchmetrics_forked_b200 import _native as nat


def _as_kernel_input(t: Tensor) -> Tensor:
    return t if t.is_contiguous() else t.contiguous()


def sigmoid_normalize(preds: Tensor, target: Optional[Tensor], ignore_index: Optional[int], err: Tensor) -> Tensor:
    """``preds.sigmoid()`` if any element (that takes part in the test) lies outside [0,1], else a copy.

    The decision is taken on the device, so ``update`` never synchronises.  ``target`` / ``ignore_index``
    exclude ignored positions from the test (pass ``None`` to test every element).
    """
    nat.require_cuda(preds, "preds")
    preds = _as_kernel_input(preds)
    out = torch.empty_like(preds)
    if preds.numel() == 0:
        return out
    dev = preds.device
    flag = torch.zeros(1, dtype=torch.int64, device=dev)  # fresh per call: concurrent streams never share it
    has_ign = int(target is not None and ignore_index is not None)
    if has_ign:
        nat.require_cuda(target, "target")
        target = _as_kernel_input(target)
    with torch.cuda.device(dev):
        rc = nat.lib().tmb_sigmoid_normalize(
            nat.ptr(preds), nat.float_code(preds), preds.numel(),
            nat.ptr(target) if has_ign else 0, nat.label_code(target) if has_ign else 0, has_ign,
            int(ignore_index) if has_ign else 0, nat.ptr(out), nat.ptr(flag), 1,
            nat.stream_ptr(dev),
        )  # fmt: skip
    nat.check(rc, "tmb_sigmoid_normalize")
    return out


def compact_not_ignored(preds: Tensor, target: Tensor, ignore_index: int) -> Tuple[Tensor, Tensor]:
    """``preds[target != ignore_index], target[target != ignore_index]`` (1-d), order preserved.

    The output length depends on the data, so this reads one count back (the reference's boolean-mask
    indexing synchronises in the same place).
    """
    nat.require_cuda(preds, "preds")
    nat.require_cuda(target, "target")
    preds = _as_kernel_input(preds.reshape(-1))
    target = _as_kernel_input(target.reshape(-1))
    n = preds.numel()
    if target.numel() != n:
        raise ValueError("`preds` and `target` must have the same number of elements")
    dev = preds.device
    out_p, out_t = torch.empty_like(preds), torch.empty_like(target)
    if n == 0:
        return out_p, out_t
    count = torch.zeros(1, dtype=torch.int64, device=dev)
    scratch = torch.empty(int(nat.lib().tmb_compact_scratch_bytes(n)), dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        rc = nat.lib().tmb_compact_not_ignored(
            nat.ptr(preds), preds.element_size(), nat.ptr(target), nat.label_code(target), n, int(ignore_index),
            nat.ptr(out_p), nat.ptr(out_t), nat.ptr(count), nat.ptr(scratch), scratch.numel(), nat.stream_ptr(dev),
        )  # fmt: skip
    nat.check(rc, "tmb_compact_not_ignored")
    kept = int(count.item())
    return out_p[:kept], out_t[:kept]


_SCORE_ERROR = (
    "exact curves need non-negative scores (probabilities / sigmoid outputs, NaN allowed): the metric's `update` "
    "always stores such values; negative scores passed directly are not supported"
)


def _run_curves(preds: Tensor, target: Tensor, ignore_index: Optional[int], pos_label: int):
    """pack -> batched radix sort -> boundary / prefix scan; everything stays on the device.

    Returns ``(fps, tps, thr)`` as ``[L, N]`` float32 buffers, ``counts`` (int64 ``[L]``, valid prefix of each
    row) and the int32 error word.
    """
    nat.require_cuda(preds, "preds")
    nat.require_cuda(target, "target")
    if preds.ndim != 2 or preds.shape != target.shape:
        raise ValueError("`preds` and `target` must both have shape (N, L)")
    preds = _as_kernel_input(preds)
    target = _as_kernel_input(target)
    n, num_labels = preds.shape
    dev = preds.device
    fps = torch.empty(num_labels, n, dtype=torch.float32, device=dev)
    tps = torch.empty(num_labels, n, dtype=torch.float32, device=dev)
    thr = torch.empty(num_labels, n, dtype=torch.float32, device=dev)
    counts = torch.zeros(num_labels, dtype=torch.int64, device=dev)
    err = torch.zeros(1, dtype=torch.int32, device=dev)
    has_ign = int(ignore_index is not None)
    if n > 0:
        scratch = torch.empty(int(nat.lib().tmb_clf_curve_scratch_bytes(n, num_labels)), dtype=torch.uint8, device=dev)
        with torch.cuda.device(dev):
            rc = nat.lib().tmb_clf_curves(
                nat.ptr(preds), nat.float_code(preds), nat.ptr(target), nat.label_code(target), n, num_labels,
                has_ign, int(ignore_index) if has_ign else 0, int(pos_label), nat.ptr(scratch), scratch.numel(),
                nat.ptr(fps), nat.ptr(tps), nat.ptr(thr), nat.ptr(counts), nat.ptr(err), nat.stream_ptr(dev),
            )  # fmt: skip
        nat.check(rc, "tmb_clf_curves")
    return fps, tps, thr, counts, err


def clf_curves(
    preds: Tensor, target: Tensor, ignore_index: Optional[int], pos_label: int = 1
) -> Tuple[List[Tensor], List[Tensor], List[Tensor]]:
    """Per-label ``(fps, tps, thresholds)`` of ``[N, L]`` scores / targets: batched radix sort + scan.

    Returns three lists of ``L`` 1-d tensors (views into three ``[L, N]`` float32 buffers); thresholds come
    back in the dtype of ``preds``.  One host read of the ``L`` curve lengths is unavoidable because the
    outputs have data-dependent sizes.
    """
    fps, tps, thr, counts, err = _run_curves(preds, target, ignore_index, pos_label)
    lengths = counts.cpu().tolist()  # the one synchronisation of the exact compute path
    if int(err.item()) & nat.ERR_SCORE_RANGE:
        raise RuntimeError(_SCORE_ERROR)
    num_labels, out_dtype = preds.shape[1], preds.dtype
    f_list = [fps[l, : lengths[l]] for l in range(num_labels)]
    t_list = [tps[l, : lengths[l]] for l in range(num_labels)]
    h_list = [thr[l, : lengths[l]] if out_dtype == torch.float32 else thr[l, : lengths[l]].to(out_dtype)
              for l in range(num_labels)]
    return f_list, t_list, h_list


def clf_curve_scores(preds: Tensor, target: Tensor, ignore_index: Optional[int], pos_label: int = 1) -> Tensor:
    """Per-label area reductions of the exact curves, float64 ``[L, 4]``, without materialising Python lists:

    column 0: trapezoid area under (fps, tps) from (0,0), un-normalised; column 1: ``sum (tps_b - tps_{b-1}) *
    precision_b``; column 2 / 3: total negatives / positives.  ``AUROC = c0 / (c2 * c3)``, ``AP = c1 / c3``.
    Nothing is read back to the host here.
    """
    fps, tps, _thr, counts, err = _run_curves(preds, target, ignore_index, pos_label)
    n, num_labels = preds.shape
    out = torch.empty(num_labels, 4, dtype=torch.float64, device=preds.device)
    with torch.cuda.device(preds.device):
        rc = nat.lib().tmb_curve_reduce(
            nat.ptr(fps), nat.ptr(tps), n, num_labels, nat.ptr(counts), nat.ptr(out), nat.stream_ptr(preds.device)
        )
    nat.check(rc, "tmb_curve_reduce")
    # a negative score poisons the result instead of costing a host sync here: NaN propagates to the metric
    return torch.where(err.bool(), torch.full_like(out, float("nan")), out)
