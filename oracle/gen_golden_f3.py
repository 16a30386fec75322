"""Golden vectors for `top_k > 1` stat scores from the UNMODIFIED reference, CPU, build container.

TEST INFRASTRUCTURE.  Usage (/root/reference must be mounted):   python oracle/gen_golden_f3.py
Scores are continuous random numbers (no ties at the k-th place: torch.topk's tie order differs between devices).
"""
import sys
import warnings
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent))
from ref_loader import load_reference  # noqa: E402

OUT = Path(__file__).resolve().parent.parent / "tests" / "golden"


def main() -> None:
    if load_reference() is None:
        raise SystemExit("reference not mounted at /root/reference")
    warnings.filterwarnings("ignore")
    import torchmetrics.functional.classification as rf

    g = torch.Generator().manual_seed(20261109)

    def case(name, preds, target, C, k, ignore_index=None):
        d = {"preds": preds.numpy(), "target": target.numpy(), "num_classes": C, "top_k": k,
             "ignore_index": -999 if ignore_index is None else ignore_index}
        kw = dict(num_classes=C, top_k=k, ignore_index=ignore_index)
        for avg in ("micro", "macro", "weighted", "none"):
            d[f"stat_scores_{avg}"] = rf.multiclass_stat_scores(preds, target, average=avg, **kw).numpy()
            for fn in ("accuracy", "precision", "recall", "f1_score", "specificity", "hamming_distance"):
                d[f"{fn}_{avg}"] = getattr(rf, f"multiclass_{fn}")(preds, target, average=avg, **kw).numpy()
        np.savez_compressed(OUT / f"f3_topk_{name}.npz", **d)

    case("c10_k3", torch.randn(2000, 10, generator=g), torch.randint(10, (2000,), generator=g), 10, 3)
    t = torch.randint(37, (1500,), generator=g)
    t[torch.rand(1500, generator=g) < 0.2] = -1
    case("c37_k5_ignore_outside", torch.randn(1500, 37, generator=g), t, 37, 5, ignore_index=-1)
    case("c37_k5_ignore_inside", torch.randn(1500, 37, generator=g), torch.randint(37, (1500,), generator=g), 37, 5,
         ignore_index=3)
    case("multidim_k2", torch.randn(40, 6, 5, 4, generator=g), torch.randint(6, (40, 5, 4), generator=g), 6, 2)
    case("c1000_k5_probs", torch.rand(300, 1000, generator=g), torch.randint(1000, (300,), generator=g), 1000, 5)
    case("k_equals_c", torch.randn(200, 4, generator=g), torch.randint(4, (200,), generator=g), 4, 4)
    # ---- multidim_average="samplewise" ------------------------------------------------------------------
    def sw_case(name, task, preds, target, **kw):
        d = {"preds": preds.numpy(), "target": target.numpy(), "ignore_index": -999 if kw.get("ignore_index") is None else kw["ignore_index"]}
        for k in ("num_classes", "num_labels"):
            if k in kw:
                d[k] = kw[k]
        avgs = (None,) if task == "binary" else ("micro", "macro", "weighted", "none")
        for avg in avgs:
            extra = {} if task == "binary" else {"average": avg}
            tag = "" if task == "binary" else f"_{avg}"
            d[f"stat_scores{tag}"] = getattr(rf, f"{task}_stat_scores")(preds, target, multidim_average="samplewise", **kw, **extra).numpy()
            for fn in ("accuracy", "precision", "recall", "f1_score", "specificity", "hamming_distance"):
                d[f"{fn}{tag}"] = getattr(rf, f"{task}_{fn}")(preds, target, multidim_average="samplewise", **kw, **extra).numpy()
        np.savez_compressed(OUT / f"f3_samplewise_{name}.npz", **d)

    bt = torch.randint(2, (9, 7, 11), generator=g)
    bt[torch.rand(9, 7, 11, generator=g) < 0.15] = -1
    sw_case("binary_logits_ignore", "binary", torch.randn(9, 7, 11, generator=g), bt, ignore_index=-1)
    sw_case("binary_probs", "binary", torch.rand(12, 50, generator=g), torch.randint(2, (12, 50), generator=g))
    mt = torch.randint(6, (8, 5, 13), generator=g)
    mt[torch.rand(8, 5, 13, generator=g) < 0.1] = -1
    sw_case("multiclass_logits_ignore", "multiclass", torch.randn(8, 6, 5, 13, generator=g), mt, num_classes=6, ignore_index=-1)
    sw_case("multiclass_labels", "multiclass", torch.randint(5, (10, 40), generator=g), torch.randint(5, (10, 40), generator=g), num_classes=5)
    sw_case("multiclass_ignore_inside", "multiclass", torch.randn(6, 4, 30, generator=g), torch.randint(4, (6, 30), generator=g), num_classes=4, ignore_index=1)
    lt = torch.randint(2, (7, 4, 20), generator=g)
    lt[torch.rand(7, 4, 20, generator=g) < 0.1] = -1
    sw_case("multilabel_logits_ignore", "multilabel", torch.randn(7, 4, 20, generator=g), lt, num_labels=4, ignore_index=-1)
    # ---- round 2: top_k > 1 beyond 1024 classes, and top_k > 1 together with samplewise ---------------------
    case("c1500_k5_wide", torch.randn(120, 1500, generator=g), torch.randint(1500, (120,), generator=g), 1500, 5)
    case("c2049_k3_wide_ignore", torch.randn(64, 2049, generator=g),
         torch.where(torch.rand(64, generator=g) < 0.2, torch.tensor(-1), torch.randint(2049, (64,), generator=g)),
         2049, 3, ignore_index=-1)

    def sw_topk_case(name, preds, target, C, k, ignore_index=None):
        d = {"preds": preds.numpy(), "target": target.numpy(), "num_classes": C, "top_k": k,
             "ignore_index": -999 if ignore_index is None else ignore_index}
        kw = dict(num_classes=C, top_k=k, ignore_index=ignore_index, multidim_average="samplewise")
        for avg in ("micro", "macro", "weighted", "none"):
            d[f"stat_scores_{avg}"] = rf.multiclass_stat_scores(preds, target, average=avg, **kw).numpy()
            for fn in ("accuracy", "precision", "recall", "f1_score", "specificity", "hamming_distance"):
                d[f"{fn}_{avg}"] = getattr(rf, f"multiclass_{fn}")(preds, target, average=avg, **kw).numpy()
        np.savez_compressed(OUT / f"f3_swtopk_{name}.npz", **d)

    sw_topk_case("c6_k2", torch.randn(9, 6, 5, 7, generator=g), torch.randint(6, (9, 5, 7), generator=g), 6, 2)
    st = torch.randint(12, (7, 40), generator=g)
    st[torch.rand(7, 40, generator=g) < 0.15] = -1
    sw_topk_case("c12_k4_ignore", torch.randn(7, 12, 40, generator=g), st, 12, 4, ignore_index=-1)
    sw_topk_case("c5_k3_ignore_inside", torch.randn(6, 5, 30, generator=g), torch.randint(5, (6, 30), generator=g), 5, 3,
                 ignore_index=2)
    print("wrote", sorted(p.name for p in OUT.glob("f3_*.npz")))


if __name__ == "__main__":
    main()
