"""Host-side launchers: torch tensors in, C-ABI calls (include/tmb200.h) out.

Each function checks devices / dtypes / contiguity, picks up torch's current stream and calls one entry
point of libtmb200.so.  Nothing here computes anything on the host and nothing falls back to torch ops:
a CPU tensor or a missing library raises.
"""
from typing import Optional, Tuple

import torch
from torch import Tensor

from torchmetrics_forked_b200 import _f64
from torchmetrics_forked_b200 import _native as nat
from torchmetrics_forked_b200._exact import (  # noqa: F401
    as_columns,
    clf_curve_scores,
    clf_curves,
    compact_not_ignored,
    sigmoid_normalize,
    softmax_normalize,
)

MC_CONFMAT, MC_STATS, MC_MICRO, MC_BOTH = nat.MC_CONFMAT, nat.MC_STATS, nat.MC_MICRO, nat.MC_BOTH


def _ignore_args(ignore_index: Optional[int]):
    return (0, 0) if ignore_index is None else (1, int(ignore_index))


def _as_kernel_input(t: Tensor) -> Tensor:
    """Kernels index dense row-major memory; anything else is made contiguous first (one copy)."""
    return t if t.is_contiguous() else t.contiguous()


_NOT_A_LABEL = -(2.0**62)  # exactly representable in every float dtype that can hold it; float16 targets saturate to -inf -> int64 min


def _as_label_input(t: Tensor) -> Tensor:
    """Targets as the kernels read them: an integer dtype, dense.  The reference takes float-typed targets holding
    integer values too (its updates end in ``.to(torch.long)`` / ``target == 1``; its own tests pass
    ``torch.zeros(10)`` as a target): they are truncated to int64 here, which is what ``.long()`` does there."""
    if t.is_floating_point():
        # a value with a fraction (or NaN) is no label: it becomes one that every task's range check rejects
        t = torch.where(t == t.trunc(), t, torch.full_like(t, _NOT_A_LABEL)).long()
    return t if t.is_contiguous() else t.contiguous()


def mc_scratch_numel(num_classes: int, mode: int) -> int:
    """int64 elements of scratch the multiclass kernels need for ``mode`` (0 for the confusion matrix)."""
    return int(nat.lib().tmb_mc_scratch_bytes(int(num_classes), int(mode))) // 8


def mc_argmax_hist(
    preds: Tensor,
    target: Tensor,
    num_classes: int,
    ignore_index: Optional[int],
    mode: int,
    out: Tensor,
    err: Tensor,
) -> None:
    """Fused argmax over dim 1 of float ``preds [N, C, ...]`` + histogram against ``target [N, ...]``.

    ``out`` is the int64 confusion-matrix state (accumulated in place) for ``MC_CONFMAT`` or the int64
    scratch for ``MC_STATS`` / ``MC_MICRO``.
    """
    nat.require_cuda(preds, "preds")
    nat.require_cuda(target, "target")
    nat.require_cuda(out, "state")
    if preds.ndim < 2 or preds.shape[1] != num_classes:
        raise ValueError("`preds` must have shape (N, C, ...) with C equal to `num_classes`")
    if _f64.is_f64(preds):  # argmax decided in float64, the label kernel counts
        if target.numel() != preds.numel() // max(num_classes, 1):
            raise RuntimeError("`target` must have one entry per (N, ...) position of `preds`")
        return mc_label_hist(_f64.multiclass_labels(preds), target, num_classes, ignore_index, mode, out, err)
    preds = _as_kernel_input(preds)
    target = _as_label_input(target)
    n_outer = preds.shape[0]
    n_inner = 1
    for d in preds.shape[2:]:
        n_inner *= d
    if target.numel() != n_outer * n_inner:
        raise RuntimeError("`target` must have one entry per (N, ...) position of `preds`")
    if n_inner == 0:
        return
    has_ign, ign = _ignore_args(ignore_index)
    with nat.on_device(preds.device):
        rc = nat.lib().tmb_mc_argmax_hist(
            nat.ptr(preds), nat.float_code(preds), nat.ptr(target), nat.label_code(target),
            n_outer, int(num_classes), n_inner, has_ign, ign, int(mode),
            nat.ptr(out), nat.ptr(err), nat.stream_ptr(preds.device),
        )  # fmt: skip
    nat.check(rc, "tmb_mc_argmax_hist")


def mc_stats_samplewise(
    preds: Tensor, target: Tensor, num_classes: int, ignore_index: Optional[int], err: Tensor, top_k: int = 1
) -> Tuple[Tensor, Tensor, Tensor, Tensor]:
    """Per-sample ``tp/fp/tn/fn`` (int64 ``[N, C]``) of ``preds [N, C, ...]`` float scores (argmax fused, or the
    ``top_k`` largest scores of every position) or ``[N, ...]`` integer labels against ``target [N, ...]``:
    ``multidim_average="samplewise"``."""
    nat.require_cuda(preds, "preds")
    nat.require_cuda(target, "target")
    if _f64.is_f64(preds) and preds.ndim >= 2 and preds.shape[1] == num_classes:
        preds = _f64.topk_marks(preds, top_k) if top_k > 1 else _f64.multiclass_labels(preds)
    is_label = not preds.is_floating_point()
    if top_k > 1:
        if is_label or preds.ndim < 2 or preds.shape[1] != num_classes:
            raise ValueError("`top_k > 1` needs float `preds` of shape (N, C, ...) with C equal to `num_classes`")
        # the positions of a sample become consecutive rows [S, C] (torch transposes; the reference's one-hot does too)
        n = preds.shape[0]
        rows = _as_kernel_input(preds.reshape(n, num_classes, -1).movedim(1, -1))
        per_sample = rows.shape[1]
        tgt = _as_label_input(target.reshape(n, -1))
        if tgt.shape[1] != per_sample:
            raise RuntimeError("`target` must have one entry per (N, ...) position of `preds`")
        dev = preds.device
        outs = tuple(torch.zeros(n, num_classes, dtype=torch.int64, device=dev) for _ in range(4))
        if n == 0 or per_sample == 0:
            return outs
        step = max(1, (1 << 25) // (3 * num_classes))
        scratch = torch.zeros(min(n, step) * 3 * num_classes, dtype=torch.int64, device=dev)
        has_ign, ign = _ignore_args(ignore_index)
        with nat.on_device(dev):
            for s0 in range(0, n, step):
                p, t = rows[s0 : s0 + step], tgt[s0 : s0 + step]
                rc = nat.lib().tmb_mc_topk_hist_samplewise(
                    nat.ptr(p), nat.float_code(p), nat.ptr(t), nat.label_code(t), p.shape[0], per_sample,
                    int(num_classes), int(top_k), has_ign, ign, nat.ptr(scratch), nat.ptr(err), nat.stream_ptr(dev),
                )  # fmt: skip
                nat.check(rc, "tmb_mc_topk_hist_samplewise")
                rc = nat.lib().tmb_mc_stats_finalize_samplewise(
                    nat.ptr(scratch), p.shape[0], int(num_classes), *(nat.ptr(o[s0 : s0 + step]) for o in outs),
                    nat.stream_ptr(dev),
                )
                nat.check(rc, "tmb_mc_stats_finalize_samplewise")
        return outs
    if is_label:
        if preds.shape != target.shape:
            raise ValueError("label-valued `preds` must have the shape of `target`")
    elif preds.ndim < 2 or preds.shape[1] != num_classes or preds.ndim != target.ndim + 1:
        raise ValueError("`preds` must have shape (N, C, ...) with C equal to `num_classes`")
    elif preds.shape[0] != target.shape[0] or preds.shape[2:] != target.shape[1:]:
        raise RuntimeError("`target` must have one entry per (N, ...) position of `preds`")
    preds = _as_kernel_input(preds)
    target = _as_label_input(target)
    n = preds.shape[0]
    inner = target.numel() // n if n else 1
    dev = preds.device
    outs = tuple(torch.empty(n, num_classes, dtype=torch.int64, device=dev) for _ in range(4))
    if n == 0:
        return outs
    if inner == 0:
        for o in outs:
            o.zero_()
        return outs
    # one [3C] histogram block per sample: the scratch is capped at 256 MB and the samples go through in slices
    step = max(1, (1 << 25) // (3 * num_classes))
    scratch = torch.zeros(min(n, step) * 3 * num_classes, dtype=torch.int64, device=dev)
    has_ign, ign = _ignore_args(ignore_index)
    pcode = nat.label_code(preds, "preds") if is_label else nat.float_code(preds)
    with nat.on_device(dev):
        for s0 in range(0, n, step):
            p, t = preds[s0 : s0 + step], target[s0 : s0 + step]  # leading-dimension slices stay contiguous
            rc = nat.lib().tmb_mc_hist_samplewise(
                nat.ptr(p), pcode, int(is_label), nat.ptr(t), nat.label_code(target), p.shape[0], int(num_classes),
                inner, has_ign, ign, nat.ptr(scratch), nat.ptr(err), nat.stream_ptr(dev),
            )  # fmt: skip
            nat.check(rc, "tmb_mc_hist_samplewise")
            rc = nat.lib().tmb_mc_stats_finalize_samplewise(
                nat.ptr(scratch), p.shape[0], int(num_classes), *(nat.ptr(o[s0 : s0 + step]) for o in outs),
                nat.stream_ptr(dev),
            )
            nat.check(rc, "tmb_mc_stats_finalize_samplewise")
    return outs


def mc_shared_confmat(scratch: Tensor, num_classes: int, confmat: Tensor) -> None:
    """Add the batch confusion matrix held in a ``MC_BOTH`` scratch into the ``[C, C]`` state and zero it."""
    nat.require_cuda(scratch, "scratch")
    nat.require_cuda(confmat, "state")
    with nat.on_device(scratch.device):
        rc = nat.lib().tmb_mc_shared_confmat(
            nat.ptr(scratch), int(num_classes), nat.ptr(confmat), nat.stream_ptr(scratch.device)
        )
    nat.check(rc, "tmb_mc_shared_confmat")


def mc_shared_stats(
    scratch: Tensor, num_classes: int, micro: bool, clear: bool, tp: Tensor, fp: Tensor, tn: Tensor, fn: Tensor
) -> None:
    """Add the stat scores held in a ``MC_BOTH`` scratch into one metric's ``tp/fp/tn/fn`` (``micro``: summed
    over classes into 1-element states); ``clear`` zeroes the histograms (last subscriber)."""
    for t in (scratch, tp, fp, tn, fn):
        nat.require_cuda(t, "state")
    with nat.on_device(scratch.device):
        rc = nat.lib().tmb_mc_shared_stats(
            nat.ptr(scratch), int(num_classes), int(micro), int(clear), nat.ptr(tp), nat.ptr(fp), nat.ptr(tn),
            nat.ptr(fn), nat.stream_ptr(scratch.device),
        )  # fmt: skip
    nat.check(rc, "tmb_mc_shared_stats")


TOPK_MAX_CLASSES = 1024  # row width the top-k kernels keep in registers; wider rows are walked once per round


def mc_topk_hist(
    preds: Tensor, target: Tensor, num_classes: int, top_k: int, ignore_index: Optional[int], out: Tensor, err: Tensor
) -> None:
    """Top-k selection over dim 1 of float ``preds [N, C, ...]`` + the three ``MC_STATS`` histograms against
    ``target [N, ...]`` (rows per target class | rows with the class in the top k | hits), added into ``out``."""
    nat.require_cuda(preds, "preds")
    nat.require_cuda(target, "target")
    nat.require_cuda(out, "state")
    if preds.ndim < 2 or preds.shape[1] != num_classes or not preds.is_floating_point():
        raise ValueError("`top_k > 1` needs float `preds` of shape (N, C, ...) with C equal to `num_classes`")
    if _f64.is_f64(preds):
        preds = _f64.topk_marks(preds, top_k)
    if preds.ndim > 2:  # extra dimensions become rows (torch transposes; the reference's one-hot does the same)
        preds = preds.movedim(1, -1).reshape(-1, num_classes)
    preds = _as_kernel_input(preds)
    target = _as_label_input(target.reshape(-1))
    if target.numel() != preds.shape[0]:
        raise RuntimeError("`target` must have one entry per (N, ...) position of `preds`")
    if preds.shape[0] == 0:
        return
    has_ign, ign = _ignore_args(ignore_index)
    with nat.on_device(preds.device):
        rc = nat.lib().tmb_mc_topk_hist(
            nat.ptr(preds), nat.float_code(preds), nat.ptr(target), nat.label_code(target), preds.shape[0],
            int(num_classes), int(top_k), has_ign, ign, nat.ptr(out), nat.ptr(err), nat.stream_ptr(preds.device),
        )  # fmt: skip
    nat.check(rc, "tmb_mc_topk_hist")


def mc_label_hist(
    preds: Tensor,
    target: Tensor,
    num_classes: int,
    ignore_index: Optional[int],
    mode: int,
    out: Tensor,
    err: Tensor,
) -> None:
    """Histogram of integer ``preds`` against ``target`` (same shape), no argmax."""
    nat.require_cuda(preds, "preds")
    nat.require_cuda(target, "target")
    nat.require_cuda(out, "state")
    preds = _as_kernel_input(preds)
    target = _as_label_input(target)
    if preds.numel() != target.numel():
        raise ValueError("`preds` and `target` must have the same number of elements")
    has_ign, ign = _ignore_args(ignore_index)
    with nat.on_device(preds.device):
        rc = nat.lib().tmb_mc_label_hist(
            nat.ptr(preds), nat.label_code(preds, "preds"), nat.ptr(target), nat.label_code(target),
            preds.numel(), int(num_classes), has_ign, ign, int(mode),
            nat.ptr(out), nat.ptr(err), nat.stream_ptr(preds.device),
        )  # fmt: skip
    nat.check(rc, "tmb_mc_label_hist")


def mc_stats_finalize(scratch: Tensor, num_classes: int, mode: int, tp: Tensor, fp: Tensor, tn: Tensor, fn: Tensor) -> None:
    """``tp/fp/tn/fn += f(scratch)`` in place, scratch re-zeroed."""
    for name, t in (("scratch", scratch), ("tp", tp), ("fp", fp), ("tn", tn), ("fn", fn)):
        nat.require_cuda(t, name)
        if t.dtype != torch.int64 or not t.is_contiguous():
            raise ValueError(f"`{name}` must be a contiguous int64 tensor")
    with nat.on_device(scratch.device):
        rc = nat.lib().tmb_mc_stats_finalize(
            nat.ptr(scratch), int(num_classes), int(mode), nat.ptr(tp), nat.ptr(fp), nat.ptr(tn), nat.ptr(fn),
            nat.stream_ptr(scratch.device),
        )  # fmt: skip
    nat.check(rc, "tmb_mc_stats_finalize")


# ------------------------------------------------------------------------------------------ K2 / K1b
TB_CURVE, TC_CONFMAT, TC_STATS = 0, 1, 2

_scalar_thresholds = {}


def raise_on_error_word(err: Tensor, where: str) -> None:
    """Read a device error word (4-byte copy, synchronises) and raise ``RuntimeError`` when it is set."""
    bits = int(err.item())
    if bits:
        err.zero_()
        reasons = []
        if bits & nat.ERR_TARGET_RANGE:
            reasons.append("`target` contains values outside [0, num_classes) that are not `ignore_index`")
        if bits & nat.ERR_PREDS_RANGE:
            reasons.append("label-valued `preds` contains values outside the valid label range")
        if bits & nat.ERR_TARGET_BINARY:
            reasons.append("`target` contains values other than 0, 1 (and `ignore_index`)")
        raise RuntimeError(f"Invalid values detected in {where}: {'; '.join(reasons)}.")


def thresh_bin_scratch_numel(num_labels: int, num_thresholds: int) -> int:
    return int(nat.lib().tmb_thresh_bin_scratch_elems(int(num_labels), int(num_thresholds)))


def thresh_count_scratch_numel(num_labels: int) -> int:
    return thresh_bin_scratch_numel(num_labels, 1)


def _label_layout(preds: Tensor, num_labels: int, binary: bool):
    """(n_outer, L, n_inner) of a ``[N, L, ...]`` multilabel tensor, or a flattened binary one."""
    if binary:
        return preds.numel(), 1, 1
    if preds.ndim < 2 or preds.shape[1] != num_labels:
        raise ValueError("`preds` must have shape (N, L, ...) with L equal to `num_labels`")
    inner = 1
    for d in preds.shape[2:]:
        inner *= d
    return preds.shape[0], num_labels, inner


def thresh_bin_update(
    preds: Tensor,
    target: Tensor,
    num_labels: int,
    binary: bool,
    thresholds: Tensor,
    cmp_gt: bool,
    ignore_index: Optional[int],
    flag_counts_ignored: bool,
    scratch: Tensor,
    err: Tensor,
) -> None:
    """One pass of (conditional sigmoid +) threshold binning into the int64 ``scratch`` histogram.

    ``thresholds`` is a float32 CUDA tensor in ascending order.  Float ``preds`` are binned; integer
    ``preds`` are taken as 0/1 labels (only meaningful with a single threshold).
    """
    nat.require_cuda(preds, "preds")
    nat.require_cuda(target, "target")
    nat.require_cuda(thresholds, "thresholds")
    if preds.shape != target.shape:
        raise RuntimeError(
            f"Predictions and targets are expected to have the same shape, but got {preds.shape} and {target.shape}."
        )
    if thresholds.dtype != torch.float32 or not thresholds.is_contiguous():
        raise ValueError("`thresholds` must be a contiguous float32 tensor")
    if _f64.is_f64(preds) and preds.numel():
        # normalised in float64, then rounded toward the side that keeps every comparison with a float32 threshold
        normal = _f64.sigmoid_where_needed(preds, None if flag_counts_ignored else target, ignore_index)
        preds = _f64.scores_for_binning(normal, strict=bool(cmp_gt))
    preds = _as_kernel_input(preds)
    target = _as_label_input(target)
    n_outer, n_labels, n_inner = _label_layout(preds, num_labels, binary)
    if preds.numel() == 0:
        return
    is_label = not preds.is_floating_point()
    pcode = nat.label_code(preds, "preds") if is_label else nat.float_code(preds)
    has_ign, ign = _ignore_args(ignore_index)
    with nat.on_device(preds.device):
        rc = nat.lib().tmb_thresh_bin_update(
            nat.ptr(preds), pcode, int(is_label), nat.ptr(target), nat.label_code(target),
            n_outer, n_labels, n_inner, nat.ptr(thresholds), thresholds.numel(), int(cmp_gt),
            has_ign, ign, int(flag_counts_ignored), nat.ptr(scratch), scratch.numel(), nat.ptr(err),
            nat.stream_ptr(preds.device),
        )  # fmt: skip
    nat.check(rc, "tmb_thresh_bin_update")


MC_CURVE_MAX_CLASSES = 1024  # row width the fused softmax + binning kernel keeps in registers


def mc_curve_update(
    preds: Tensor,
    target: Tensor,
    thresholds: Tensor,
    ignore_index: Optional[int],
    micro: bool,
    scratch: Tensor,
    err: Tensor,
) -> None:
    """One pass of (conditional row softmax +) threshold binning of ``[M, C]`` multiclass scores with one-hot
    targets ``class == target[row]`` into the int64 ``scratch`` histogram (``C`` labels, or 1 with ``micro``).
    Follow with :func:`thresh_bin_finalize`."""
    nat.require_cuda(preds, "preds")
    nat.require_cuda(target, "target")
    nat.require_cuda(thresholds, "thresholds")
    if preds.ndim != 2 or target.ndim != 1 or target.shape[0] != preds.shape[0]:
        raise ValueError("`preds` must have shape (M, C) and `target` shape (M,)")
    if thresholds.dtype != torch.float32 or not thresholds.is_contiguous():
        raise ValueError("`thresholds` must be a contiguous float32 tensor")
    if _f64.is_f64(preds) and preds.numel():
        preds = _f64.scores_for_binning(_f64.softmax_where_needed(preds, target, ignore_index), strict=False)
    preds = _as_kernel_input(preds)
    target = _as_label_input(target)
    if preds.numel() == 0:
        return
    has_ign, ign = _ignore_args(ignore_index)
    with nat.on_device(preds.device):
        rc = nat.lib().tmb_mc_curve_update(
            nat.ptr(preds), nat.float_code(preds), nat.ptr(target), nat.label_code(target), preds.shape[0],
            preds.shape[1], nat.ptr(thresholds), thresholds.numel(), has_ign, ign, int(micro), nat.ptr(scratch),
            scratch.numel(), nat.ptr(err), nat.stream_ptr(preds.device),
        )  # fmt: skip
    nat.check(rc, "tmb_mc_curve_update")


def thresh_bin_finalize(
    scratch: Tensor, num_labels: int, num_thresholds: int, perm: Optional[Tensor], layout: int, out
) -> None:
    """Add the per-threshold counts held in ``scratch`` into the state tensor(s) ``out`` and clear ``scratch``."""
    outs = list(out) + [None] * (4 - len(out))
    for t in out:
        nat.require_cuda(t, "state")
        if t.dtype != torch.int64 or not t.is_contiguous():
            raise ValueError("metric states updated by the kernels must be contiguous int64 tensors")
    with nat.on_device(scratch.device):
        rc = nat.lib().tmb_thresh_bin_finalize(
            nat.ptr(scratch), scratch.numel(), int(num_labels), int(num_thresholds), nat.ptr(perm), int(layout),
            nat.ptr(outs[0]), nat.ptr(outs[1]), nat.ptr(outs[2]), nat.ptr(outs[3]), nat.stream_ptr(scratch.device),
        )  # fmt: skip
    nat.check(rc, "tmb_thresh_bin_finalize")


def _scalar_threshold(threshold: float, like: Tensor) -> Tensor:
    """A cached one-element float32 device tensor holding ``threshold`` rounded to the dtype of ``like``.

    The reference evaluates ``preds > threshold`` in the dtype of ``preds`` (the Python scalar is cast),
    so for half / bfloat16 inputs the threshold is rounded first.
    """
    value = float(torch.tensor(threshold, dtype=like.dtype if like.is_floating_point() else torch.float32).item())
    key = (value, str(like.device))
    t = _scalar_thresholds.get(key)
    if t is None:
        t = torch.tensor([value], dtype=torch.float32, device=like.device)
        _scalar_thresholds[key] = t
    return t


def thresh_count(
    preds: Tensor,
    target: Tensor,
    num_labels: int,
    threshold: float,
    ignore_index: Optional[int],
    flag_counts_ignored: bool,
    layout: int,
    scratch: Tensor,
    out,
    err: Tensor,
    binary: Optional[bool] = None,
) -> None:
    """K1b: ``(sigmoid?)(preds) > threshold`` against 0/1 targets, counted per label into ``out`` in place.

    ``layout`` = ``TC_CONFMAT`` (``out = (confmat[L,2,2],)``) or ``TC_STATS`` (``out = (tp, fp, tn, fn)``).
    """
    if binary is None:
        binary = num_labels == 1
    if _f64.is_f64(preds):  # the threshold is a Python float: compared in float64, the kernel counts 0/1 labels
        nat.require_cuda(preds, "preds")
        if preds.shape != target.shape:
            raise RuntimeError(
                f"Predictions and targets are expected to have the same shape, but got {preds.shape} and {target.shape}."
            )
        preds = _f64.binary_labels(preds, target, threshold, ignore_index, bool(flag_counts_ignored))
    thr = _scalar_threshold(threshold, preds)
    thresh_bin_update(preds, target, num_labels, binary, thr, True, ignore_index, flag_counts_ignored, scratch, err)
    thresh_bin_finalize(scratch, 1 if binary else num_labels, 1, None, layout, out)


def bincount(x: Tensor, minlength: int) -> Tensor:
    """``torch.bincount(x, minlength=minlength)`` for CUDA integer tensors (deterministic integer atomics)."""
    nat.require_cuda(x, "x")
    x = _as_kernel_input(x.reshape(-1))
    code = nat.label_code(x, "x")
    dev = x.device
    with nat.on_device(dev):
        top = torch.empty(1, dtype=torch.int64, device=dev)
        nat.check(nat.lib().tmb_label_max(nat.ptr(x), code, x.numel(), nat.ptr(top), nat.stream_ptr(dev)), "tmb_label_max")
        biggest = int(top.item()) if x.numel() else -1  # the output size depends on the data, like torch.bincount
        length = max(int(minlength), biggest + 1)
        out = torch.zeros(length, dtype=torch.int64, device=dev)
        err = torch.zeros(1, dtype=torch.int32, device=dev)
        nat.check(
            nat.lib().tmb_bincount(nat.ptr(x), code, x.numel(), length, nat.ptr(out), nat.ptr(err), nat.stream_ptr(dev)),
            "tmb_bincount",
        )
        if int(err.item()):
            raise RuntimeError("bincount only supports 1-d non-negative integral inputs.")
    return out
