"""ORACLE (timing port) — the reference's update path restated with the SAME torch CPU (ATen) calls it makes.

TEST / MEASUREMENT INFRASTRUCTURE, NOT PRODUCT CODE.  `bench.py` times these functions on the GPU box's host
cores as the CPU baseline and as the `--impl reference` arm (the reference itself cannot travel to the GPU
box and is pure Python over ATen, so an op-for-op port with all intra-op threads is the fair stand-in);
`tests/test_oracle_vs_reference.py` checks them against the live reference in the build container and
`tests/test_oracle_golden.py` against the committed golden vectors.  Each function lists the reference lines
whose op sequence it repeats (paths relative to /root/reference/src/torchmetrics/).
"""
from typing import Optional

import torch
from torch import Tensor


def multiclass_confusion_matrix_update(
    preds: Tensor, target: Tensor, num_classes: int, ignore_index: Optional[int] = None, validate_args: bool = True
) -> Tensor:
    """One `MulticlassConfusionMatrix.update` (classification/confusion_matrix.py:275-281):
    tensor validation (functional/classification/confusion_matrix.py:288-303: `unique`), format (:306-330:
    `argmax`, flatten, boolean mask), update (:333-337: `target*C+pred`, `bincount`, reshape)."""
    if validate_args:
        n_unique = len(torch.unique(target))
        limit = num_classes if ignore_index is None else num_classes + 1
        if n_unique > limit:
            raise RuntimeError("Detected more unique values in `target` than `num_classes`.")
    if preds.ndim == target.ndim + 1:
        preds = preds.argmax(dim=1)
    preds = preds.flatten()
    target = target.flatten()
    if ignore_index is not None:
        keep = target != ignore_index
        preds, target = preds[keep], target[keep]
    mapping = target.to(torch.long) * num_classes + preds.to(torch.long)
    return torch.bincount(mapping, minlength=num_classes**2).reshape(num_classes, num_classes)


def multiclass_stat_scores_update(
    preds: Tensor, target: Tensor, num_classes: int, average: Optional[str] = "macro",
    ignore_index: Optional[int] = None, validate_args: bool = True,
):  # fmt: skip
    """One `MulticlassStatScores.update` (functional/classification/stat_scores.py:307-322, :325-341,
    :394-418), global, top_k=1."""
    if validate_args:
        len(torch.unique(target))
    if preds.ndim == target.ndim + 1:
        preds = preds.argmax(dim=1)
    preds, target = preds.flatten(), target.flatten()
    if ignore_index is not None:
        keep = target != ignore_index
        preds, target = preds[keep], target[keep]
    if average == "micro":
        tp = (preds == target).sum()
        fp = (preds != target).sum()
        fn = (preds != target).sum()
        tn = num_classes * preds.numel() - (fp + fn + tp)
        return tp, fp, tn, fn
    mapping = target.to(torch.long) * num_classes + preds.to(torch.long)
    confmat = torch.bincount(mapping, minlength=num_classes**2).reshape(num_classes, num_classes)
    tp = confmat.diag()
    fp = confmat.sum(0) - tp
    fn = confmat.sum(1) - tp
    tn = confmat.sum() - (fp + fn + tp)
    return tp, fp, tn, fn


def binary_curve_update(
    preds: Tensor, target: Tensor, thresholds: Tensor, ignore_index: Optional[int] = None, validate_args: bool = True
) -> Tensor:
    """One binned `BinaryPrecisionRecallCurve.update` (functional/classification/precision_recall_curve.py
    :149-159 validation, :162-187 format, :190-250 update with the 50 000-sample switch)."""
    if validate_args:
        torch.unique(target)
    preds, target = preds.flatten(), target.flatten()
    if ignore_index is not None:
        keep = target != ignore_index
        preds, target = preds[keep], target[keep]
    if not torch.all((preds >= 0) * (preds <= 1)):
        preds = preds.sigmoid()
    len_t = len(thresholds)
    if preds.numel() <= 50_000:
        preds_t = (preds.unsqueeze(-1) >= thresholds.unsqueeze(0)).long()
        mapping = preds_t + 2 * target.long().unsqueeze(-1) + 4 * torch.arange(len_t)
        return torch.bincount(mapping.flatten(), minlength=4 * len_t).reshape(len_t, 2, 2)
    pos = target == 1
    confmat = thresholds.new_empty((len_t, 2, 2), dtype=torch.int64)
    for i in range(len_t):
        preds_t = preds >= thresholds[i]
        confmat[i, 1, 1] = (pos & preds_t).sum()
        confmat[i, 0, 1] = ((~pos) & preds_t).sum()
        confmat[i, 1, 0] = (pos & (~preds_t)).sum()
    confmat[:, 0, 0] = len(preds_t) - confmat[:, 0, 1] - confmat[:, 1, 0] - confmat[:, 1, 1]
    return confmat


def binary_clf_curve(preds: Tensor, target: Tensor, pos_label: int = 1):
    """`_binary_clf_curve` (functional/classification/precision_recall_curve.py:28-80), unweighted."""
    order = torch.argsort(preds, descending=True)
    preds, target = preds[order], target[order]
    distinct = torch.where(preds[1:] - preds[:-1])[0]
    idx = torch.nn.functional.pad(distinct, [0, 1], value=target.size(0) - 1)
    pos = (target == pos_label).to(torch.long)
    tps = torch.cumsum(pos * 1.0, dim=0)[idx]
    fps = 1 + idx - tps
    return fps, tps, preds[idx]


def multilabel_exact_update(preds: Tensor, target: Tensor, num_labels: int):
    """Exact-mode `MultilabelPrecisionRecallCurve.update` (precision_recall_curve.py:752-757): what is stored."""
    preds = preds.transpose(0, 1).reshape(num_labels, -1).T
    target = target.transpose(0, 1).reshape(num_labels, -1).T
    if not torch.all((preds >= 0) * (preds <= 1)):
        preds = preds.sigmoid()
    return preds, target


def multilabel_auroc_exact_compute(preds: Tensor, target: Tensor, num_labels: int) -> Tensor:
    """`_multilabel_auroc_compute`, macro, exact (auroc.py:307-332 -> roc.py:345-355 -> :53-78 -> trapz)."""
    res = []
    for i in range(num_labels):
        fps, tps, _ = binary_clf_curve(preds[:, i], target[:, i])
        tps = torch.cat([torch.zeros(1), tps])
        fps = torch.cat([torch.zeros(1), fps])
        fpr = fps / fps[-1] if fps[-1] > 0 else torch.zeros_like(fps)
        tpr = tps / tps[-1] if tps[-1] > 0 else torch.zeros_like(tps)
        res.append(torch.trapz(tpr, fpr))
    res = torch.stack(res)
    return res[~torch.isnan(res)].mean()
