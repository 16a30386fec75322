"""Golden vectors for the multiclass curves from the UNMODIFIED reference, CPU, build container.

TEST INFRASTRUCTURE.  Usage (/root/reference must be mounted):   python oracle/gen_golden_f2.py

tests/golden/f2_mccurve_*.npz: seeded inputs, the reference's binned states ([T, C, 2, 2] and the micro [T, 2, 2]),
the exact-mode stored state, per-class exact curves and the scalar metrics built on them.
"""
import sys
import warnings
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent))
from ref_loader import load_reference  # noqa: E402

OUT = Path(__file__).resolve().parent.parent / "tests" / "golden"


def _np(x):
    return x.detach().cpu().numpy() if isinstance(x, torch.Tensor) else np.asarray(x)


def main() -> None:
    if load_reference() is None:
        raise SystemExit("reference not mounted at /root/reference")
    warnings.filterwarnings("ignore")
    import torchmetrics.functional.classification as rf
    from torchmetrics.functional.classification.precision_recall_curve import (
        _binary_clf_curve,
        _multiclass_precision_recall_curve_format,
        _multiclass_precision_recall_curve_update,
    )

    g = torch.Generator().manual_seed(20261108)
    OUT.mkdir(parents=True, exist_ok=True)

    def case(name, preds, target, C, ignore_index=None, T=11):
        thr = torch.linspace(0, 1, T)
        d = {"preds": _np(preds), "target": _np(target), "num_classes": C, "thresholds": _np(thr),
             "ignore_index": -999 if ignore_index is None else ignore_index}
        for avg in (None, "micro"):
            p, t, th = _multiclass_precision_recall_curve_format(preds, target, C, thr, ignore_index, avg)
            d["state_micro" if avg else "state"] = _np(_multiclass_precision_recall_curve_update(p, t, C, th, avg))
        p, t, _ = _multiclass_precision_recall_curve_format(preds, target, C, None, ignore_index, None)
        d["exact_preds"], d["exact_target"] = _np(p), _np(t)
        for c in range(C):
            fps, tps, th = _binary_clf_curve(p[:, c], t, pos_label=c)
            d[f"fps_{c}"], d[f"tps_{c}"], d[f"thr_{c}"] = _np(fps), _np(tps), _np(th)
        kw = dict(num_classes=C, ignore_index=ignore_index)
        for avg in ("macro", "weighted", "none"):
            for T_, tag in ((None, "exact"), (T, "binned")):
                d[f"auroc_{avg}_{tag}"] = _np(rf.multiclass_auroc(preds, target, average=avg, thresholds=T_, **kw))
                d[f"ap_{avg}_{tag}"] = _np(rf.multiclass_average_precision(preds, target, average=avg, thresholds=T_, **kw))
        for stem, kwn in (("recall_at_fixed_precision", "min_precision"), ("precision_at_fixed_recall", "min_recall"),
                          ("specificity_at_sensitivity", "min_sensitivity")):
            for T_, tag in ((None, "exact"), (T, "binned")):
                a, b = getattr(rf, f"multiclass_{stem}")(preds, target, **{kwn: 0.4}, thresholds=T_, **kw)
                d[f"{stem}_{tag}"] = np.stack([_np(a), _np(b)]).astype(np.float64)
        for avg in ("macro", "micro"):
            pr = rf.multiclass_precision_recall_curve(preds, target, thresholds=T, average=avg, **kw)
            roc = rf.multiclass_roc(preds, target, thresholds=T, average=avg, **kw)
            for k, v in zip(("p", "r", "t"), pr):
                d[f"prc_{avg}_{k}"] = _np(v)
            for k, v in zip(("f", "t", "h"), roc):
                d[f"roc_{avg}_{k}"] = _np(v)
        np.savez_compressed(OUT / f"f2_mccurve_{name}.npz", **d)

    n, C = 1500, 5
    case("logits", torch.randn(n, C, generator=g) * 2, torch.randint(C, (n,), generator=g), C)
    case("probs", torch.rand(n, C, generator=g).softmax(1), torch.randint(C, (n,), generator=g), C)
    t = torch.randint(C, (n,), generator=g)
    t[torch.rand(n, generator=g) < 0.2] = -1
    case("logits_ignore", torch.randn(n, C, generator=g) * 3, t, C, ignore_index=-1)
    p = torch.rand(400, 3, generator=g)
    p[7, 1] = 1.5  # only this row is outside [0,1] ...
    t = torch.randint(3, (400,), generator=g)
    t[7] = -1      # ... and it is ignored: no softmax (the reference filters rows before it looks)
    case("probs_outlier_ignored", p, t, 3, ignore_index=-1)
    case("multidim", torch.randn(40, 4, 6, 5, generator=g), torch.randint(4, (40, 6, 5), generator=g), 4)
    case("wide", torch.randn(300, 37, generator=g), torch.randint(37, (300,), generator=g), 37, T=5)
    print("wrote", sorted(p.name for p in OUT.glob("f2_*.npz")))


if __name__ == "__main__":
    main()
