"""Host-side launchers of the exact-curve kernels (K3): conditional sigmoid, ordered ignore-filter, batched
sort + scan, area reductions.  Re-exported through :mod:`torchmetrics_forked_b200.kernels`."""
from typing import List, Optional, Sequence, Tuple, Union

import torch
from torch import Tensor

from torchmetrics_forked_b200 import _f64
from torchmetrics_forked_b200 import _native as nat


def _as_kernel_input(t: Tensor) -> Tensor:
    return t if t.is_contiguous() else t.contiguous()


_NOT_A_LABEL = -(2.0**62)  # exactly representable in every float dtype that can hold it; float16 targets saturate to -inf -> int64 min


def _as_label_input(t: Tensor) -> Tensor:
    """Integer, dense targets (float-typed targets holding integer values are truncated like the reference's
    ``.long()``)."""
    if t.is_floating_point():
        # a value with a fraction (or NaN) is no label: it becomes one that every task's range check rejects
        t = torch.where(t == t.trunc(), t, torch.full_like(t, _NOT_A_LABEL)).long()
    return t if t.is_contiguous() else t.contiguous()


def sigmoid_normalize(preds: Tensor, target: Optional[Tensor], ignore_index: Optional[int], err: Tensor) -> Tensor:
    """``preds.sigmoid()`` if any element (that takes part in the test) lies outside [0,1], else a copy.

    The decision is taken on the device, so ``update`` never synchronises.  ``target`` / ``ignore_index``
    exclude ignored positions from the test (pass ``None`` to test every element).
    """
    nat.require_cuda(preds, "preds")
    if _f64.is_f64(preds):
        return _f64.sigmoid_where_needed(preds, target, ignore_index)
    preds = _as_kernel_input(preds)
    out = torch.empty_like(preds)
    if preds.numel() == 0:
        return out
    dev = preds.device
    flag = torch.zeros(1, dtype=torch.int64, device=dev)  # fresh per call: concurrent streams never share it
    has_ign = int(target is not None and ignore_index is not None)
    if has_ign:
        nat.require_cuda(target, "target")
        target = _as_label_input(target)
    with nat.on_device(dev):
        rc = nat.lib().tmb_sigmoid_normalize(
            nat.ptr(preds), nat.float_code(preds), preds.numel(),
            nat.ptr(target) if has_ign else 0, nat.label_code(target) if has_ign else 0, has_ign,
            int(ignore_index) if has_ign else 0, nat.ptr(out), nat.ptr(flag), 1,
            nat.stream_ptr(dev),
        )  # fmt: skip
    nat.check(rc, "tmb_sigmoid_normalize")
    return out


SOFTMAX_MAX_CLASSES = 1024  # rows up to here go through tmb_softmax_rows (bit-equal to torch); wider ones through torch


def softmax_normalize(
    preds: Tensor, target: Optional[Tensor] = None, ignore_index: Optional[int] = None
) -> Tensor:
    """``preds.softmax(1)`` of an ``[M, C]`` matrix if any element of a row that is not ignored lies outside [0,1],
    else a copy — the multiclass twin of :func:`sigmoid_normalize`, decided on the device.

    ``target`` holds one class index per row; rows with ``target == ignore_index`` take no part in the test (the
    reference removes them before it looks, precision_recall_curve.py:445-452).
    """
    nat.require_cuda(preds, "preds")
    if preds.ndim != 2:
        raise ValueError("`preds` must have shape (M, C)")
    if _f64.is_f64(preds):
        return _f64.softmax_where_needed(preds, target, ignore_index)
    preds = _as_kernel_input(preds)
    rows, num_classes = preds.shape
    if preds.numel() == 0:
        return torch.empty_like(preds)
    dev = preds.device
    flag = torch.zeros(1, dtype=torch.int64, device=dev)
    has_ign = int(target is not None and ignore_index is not None)
    if has_ign:
        nat.require_cuda(target, "target")
        target = _as_label_input(target.reshape(-1))
        if target.numel() != rows:
            raise ValueError("`target` must hold one class index per row of `preds`")
    t_ptr, t_code, ign = (nat.ptr(target), nat.label_code(target), int(ignore_index)) if has_ign else (0, 0, 0)
    with nat.on_device(dev):
        if num_classes > SOFTMAX_MAX_CLASSES:
            rc = nat.lib().tmb_rows_outside_flag(
                nat.ptr(preds), nat.float_code(preds), rows, num_classes, t_ptr, t_code, has_ign, ign, nat.ptr(flag), 1,
                nat.stream_ptr(dev),
            )
            nat.check(rc, "tmb_rows_outside_flag")
            # torch uses a different (block-reduction) kernel for wide rows; keep its bits by keeping its kernel
            return torch.where(flag.bool(), preds.softmax(1), preds)
        out = torch.empty_like(preds)
        rc = nat.lib().tmb_softmax_normalize(
            nat.ptr(preds), nat.float_code(preds), rows, num_classes, t_ptr, t_code, has_ign, ign, nat.ptr(out),
            nat.ptr(flag), 1, nat.stream_ptr(dev),
        )
        nat.check(rc, "tmb_softmax_normalize")
    return out


def compact_not_ignored(preds: Tensor, target: Tensor, ignore_index: int) -> Tuple[Tensor, Tensor]:
    """``preds[target != ignore_index], target[target != ignore_index]`` (1-d), order preserved.

    The output length depends on the data, so this reads one count back (the reference's boolean-mask
    indexing synchronises in the same place).
    """
    nat.require_cuda(preds, "preds")
    nat.require_cuda(target, "target")
    preds = _as_kernel_input(preds.reshape(-1))
    target = _as_label_input(target.reshape(-1))
    n = preds.numel()
    if target.numel() != n:
        raise ValueError("`preds` and `target` must have the same number of elements")
    dev = preds.device
    out_p, out_t = torch.empty_like(preds), torch.empty_like(target)
    if n == 0:
        return out_p, out_t
    count = torch.zeros(1, dtype=torch.int64, device=dev)
    scratch = torch.empty(int(nat.lib().tmb_compact_scratch_bytes(n)), dtype=torch.uint8, device=dev)
    with nat.on_device(dev):
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


Chunks = Union[Tensor, Sequence[Tensor]]


def as_columns(x: Chunks) -> Chunks:
    """1-d (binary) scores / targets, or a list of such chunks, as ``[n, 1]`` column(s)."""
    if isinstance(x, Tensor):
        return x.reshape(-1, 1)
    cols = [c.reshape(-1, 1) for c in x]
    info = getattr(x, "shard_info", None)
    if info is not None:  # keep the label-sharded marker (see _sharded.ShardedChunks)
        return type(x)(cols, info)
    return cols


def _chunk_lists(preds: Chunks, target: Chunks, class_targets: bool = False) -> Tuple[List[Tensor], List[Tensor]]:
    """Validated, contiguous ``[n_i, L]`` chunk lists (the per-update buffers of an exact-mode state); with
    ``class_targets`` the target chunks are ``[n_i]`` class indices."""
    ps = [preds] if isinstance(preds, Tensor) else list(preds)
    ts = [target] if isinstance(target, Tensor) else list(target)
    if len(ps) != len(ts) or not ps:
        raise ValueError("`preds` and `target` must hold the same, non-zero number of chunks")
    for p, t in zip(ps, ts):
        nat.require_cuda(p, "preds")
        nat.require_cuda(t, "target")
        if class_targets:
            if p.ndim != 2 or t.ndim != 1 or t.shape[0] != p.shape[0]:
                raise ValueError("`preds` must have shape (N, C) and `target` shape (N,)")
        elif p.ndim != 2 or p.shape != t.shape:
            raise ValueError("`preds` and `target` must both have shape (N, L)")
        if p.shape[1] != ps[0].shape[1] or p.dtype != ps[0].dtype or t.dtype != ts[0].dtype:
            raise ValueError("all chunks must agree in the number of labels and in dtype")
        if p.device != ps[0].device or t.device != ps[0].device:
            raise ValueError("all chunks must live on one device")
    return [_as_kernel_input(p) for p in ps], [_as_label_input(t) for t in ts]


def _chunk_table(ps: List[Tensor], ts: List[Tensor]) -> Tuple[Tensor, int]:
    """Device table for the chunked entry points (layout: include/tmb200.h) and the total row count."""
    starts = [0]
    for p in ps:
        starts.append(starts[-1] + p.shape[0])
    table = torch.tensor(
        starts + [p.data_ptr() for p in ps] + [t.data_ptr() for t in ts], dtype=torch.int64, device=ps[0].device
    )
    return table, starts[-1]


def _run_curves(
    preds: Chunks, target: Chunks, ignore_index: Optional[int], pos_label: int, class_targets: bool = False
):
    """pack -> batched radix sort -> boundary / prefix scan; everything stays on the device.

    ``preds`` / ``target`` are ``[N, L]`` tensors or lists of ``[n_i, L]`` chunks (read in place, never
    concatenated).  Returns ``(fps, tps, thr)`` as ``[L, N]`` float32 buffers, ``counts`` (int64 ``[L]``, valid
    prefix of each row) and the int32 error word.
    """
    ps, ts = _chunk_lists(preds, target, class_targets)
    values = None
    if _f64.is_f64(ps[0]):  # float64 scores go through the sort as their dense ranks (see _f64)
        ranked, values = _f64.rank_scores(ps)
        ps = [_as_kernel_input(c) for c in _f64.split_like(ranked, ps)]
    num_labels = ps[0].shape[1]
    n = sum(p.shape[0] for p in ps)
    dev = ps[0].device
    fps = torch.empty(num_labels, n, dtype=torch.float32, device=dev)
    tps = torch.empty(num_labels, n, dtype=torch.float32, device=dev)
    thr = torch.empty(num_labels, n, dtype=torch.float32, device=dev)
    counts = torch.zeros(num_labels, dtype=torch.int64, device=dev)
    err = torch.zeros(1, dtype=torch.int32, device=dev)
    has_ign = int(ignore_index is not None)
    if n > 0:
        scratch = torch.empty(int(nat.lib().tmb_clf_curve_scratch_bytes(n, num_labels)), dtype=torch.uint8, device=dev)
        head = (n, num_labels, has_ign, int(ignore_index) if has_ign else 0, int(pos_label))
        tail = (
            nat.ptr(scratch), scratch.numel(), nat.ptr(fps), nat.ptr(tps), nat.ptr(thr), nat.ptr(counts),
            nat.ptr(err), nat.stream_ptr(dev),
        )  # fmt: skip
        with nat.on_device(dev):
            if len(ps) == 1 and not class_targets:
                rc = nat.lib().tmb_clf_curves(
                    nat.ptr(ps[0]), nat.float_code(ps[0]), nat.ptr(ts[0]), nat.label_code(ts[0]), *head, *tail
                )
                nat.check(rc, "tmb_clf_curves")
            else:
                table, _ = _chunk_table(ps, ts)
                rc = nat.lib().tmb_clf_curves_chunked(
                    nat.ptr(table), len(ps), nat.float_code(ps[0]), nat.label_code(ts[0]), *head,
                    int(class_targets), *tail
                )
                nat.check(rc, "tmb_clf_curves_chunked")
    return fps, tps, thr, counts, err, (ps[0].dtype if values is None else values)


def clf_curves(
    preds: Chunks, target: Chunks, ignore_index: Optional[int], pos_label: int = 1, class_targets: bool = False
) -> Tuple[List[Tensor], List[Tensor], List[Tensor]]:
    """Per-label ``(fps, tps, thresholds)`` of ``[N, L]`` scores / targets (or lists of ``[n_i, L]`` chunks):
    batched radix sort + scan.  ``class_targets=True``: multiclass form, ``target`` chunks are ``[n_i]`` class
    indices and label ``l`` is positive where ``target == l``.

    Returns three lists of ``L`` 1-d tensors (views into three ``[L, N]`` float32 buffers); thresholds come
    back in the dtype of ``preds``.  One host read of the ``L`` curve lengths is unavoidable because the
    outputs have data-dependent sizes.
    """
    fps, tps, thr, counts, err, out_dtype = _run_curves(preds, target, ignore_index, pos_label, class_targets)
    lengths = counts.cpu().tolist()  # the one synchronisation of the exact compute path
    if int(err.item()) & nat.ERR_SCORE_RANGE:
        raise RuntimeError(_SCORE_ERROR)
    num_labels = fps.shape[0]
    f_list = [fps[l, : lengths[l]] for l in range(num_labels)]
    t_list = [tps[l, : lengths[l]] for l in range(num_labels)]
    if isinstance(out_dtype, Tensor):  # float64 scores: `thr` holds rank patterns, `out_dtype` the distinct values
        h_list = [_f64.scores_of_ranks(thr[l, : lengths[l]], out_dtype) for l in range(num_labels)]
    else:
        h_list = [thr[l, : lengths[l]] if out_dtype == torch.float32 else thr[l, : lengths[l]].to(out_dtype)
                  for l in range(num_labels)]
    return f_list, t_list, h_list


def clf_curve_scores(
    preds: Chunks, target: Chunks, ignore_index: Optional[int], pos_label: int = 1, want_ap: bool = True,
    class_targets: bool = False,
) -> Tensor:  # fmt: skip
    """Per-label area reductions of the exact curves, float64 ``[L, 4]``, straight from the sorted keys — the
    curves themselves are never written:

    column 0: trapezoid area under (fps, tps) from (0,0), un-normalised; column 1: ``sum (tps_b - tps_{b-1}) *
    precision_b``; column 2 / 3: total negatives / positives.  ``AUROC = c0 / (c2 * c3)``, ``AP = c1 / c3``.
    ``preds`` / ``target`` as in :func:`clf_curves`; ``want_ap=False`` leaves column 1 at zero (AUROC does not need
    its float64 division per curve point).  Nothing is read back to the host here.

    When ``preds`` carries a ``shard_info`` (the metric's ``compute()`` under a process group, see
    :mod:`torchmetrics_forked_b200._sharded`) the chunks are this rank's LOCAL rows and the result covers the rows of
    all ranks: the labels are dealt out to the ranks, each sorts its share.
    """
    info = getattr(preds, "shard_info", None)
    if info is not None:
        from torchmetrics_forked_b200 import _sharded

        if len(preds) > 0:
            ps, ts = _chunk_lists(list(preds), list(target), class_targets)
        else:
            ps, ts = [], []
        return _sharded.sharded_curve_scores(ps, ts, info, ignore_index, pos_label, want_ap, class_targets)
    ps, ts = _chunk_lists(preds, target, class_targets)
    if _f64.is_f64(ps[0]):
        ps = [_as_kernel_input(c) for c in _f64.split_like(_f64.rank_scores(ps)[0], ps)]
    num_labels = ps[0].shape[1]
    dev = ps[0].device
    table, n = _chunk_table(ps, ts)
    out = torch.empty(num_labels, 4, dtype=torch.float64, device=dev)
    err = torch.zeros(1, dtype=torch.int32, device=dev)
    has_ign = int(ignore_index is not None)
    nbytes = int(nat.lib().tmb_clf_curve_scratch_bytes(n, num_labels))
    scratch = torch.empty(max(nbytes, 1), dtype=torch.uint8, device=dev)
    with nat.on_device(dev):
        rc = nat.lib().tmb_clf_curve_scores_chunked(
            nat.ptr(table), len(ps), nat.float_code(ps[0]), nat.label_code(ts[0]), n, num_labels, has_ign,
            int(ignore_index) if has_ign else 0, int(pos_label), int(class_targets), int(want_ap), nat.ptr(scratch),
            scratch.numel(),
            nat.ptr(out), nat.ptr(err), nat.stream_ptr(dev),
        )  # fmt: skip
    nat.check(rc, "tmb_clf_curve_scores_chunked")
    # a negative score poisons the result instead of costing a host sync here: NaN propagates to the metric
    return torch.where(err.bool(), torch.full_like(out, float("nan")), out)
