"""Golden vectors for the metrics that only add a reducer to the hot-path updates, produced by
running the UNMODIFIED reference on the CPU of the build container.

TEST INFRASTRUCTURE.  Usage (/root/reference must be mounted):   python oracle/gen_golden_f1.py

Writes tests/golden/f1_*.npz: reducer inputs/outputs (states -> scores), operating-point selections on curves, and
end-to-end cases (seeded inputs -> the reference's functional results).
"""
import sys
import warnings
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent))
from ref_loader import load_reference  # noqa: E402

OUT = Path(__file__).resolve().parent.parent / "tests" / "golden"
AVERAGES = ("micro", "macro", "weighted", "none")


def _np(x):
    return x.detach().cpu().numpy() if isinstance(x, torch.Tensor) else np.asarray(x)


def main() -> None:
    if load_reference() is None:
        raise SystemExit("reference not mounted at /root/reference")
    warnings.filterwarnings("ignore")
    import torchmetrics.functional.classification as rf
    from torchmetrics.functional.classification.cohen_kappa import _cohen_kappa_reduce
    from torchmetrics.functional.classification.hamming import _hamming_distance_reduce
    from torchmetrics.functional.classification.matthews_corrcoef import _matthews_corrcoef_reduce
    from torchmetrics.functional.classification.precision_fixed_recall import _precision_at_recall
    from torchmetrics.functional.classification.precision_recall import _precision_recall_reduce
    from torchmetrics.functional.classification.recall_fixed_precision import _recall_at_precision
    from torchmetrics.functional.classification.specificity import _specificity_reduce
    from torchmetrics.functional.classification.specificity_sensitivity import _specificity_at_sensitivity

    g = torch.Generator().manual_seed(20261107)
    OUT.mkdir(parents=True, exist_ok=True)

    # ---- stat-score reducers ------------------------------------------------------------------------------
    red = {}
    states = []
    for C in (1, 4, 7, 50):
        for _ in range(3):
            st = torch.randint(0, 2000, (4, C), generator=g)
            st[:, torch.rand(C, generator=g) < 0.25] = 0  # classes without any sample: the 0/0 branches
            states.append(st)
    states.append(torch.zeros(4, 5, dtype=torch.long))
    red["n_states"] = len(states)
    reducers = {
        "precision": lambda *a, **k: _precision_recall_reduce("precision", *a, **k),
        "recall": lambda *a, **k: _precision_recall_reduce("recall", *a, **k),
        "specificity": _specificity_reduce,
        "hamming": _hamming_distance_reduce,
    }
    for i, st in enumerate(states):
        red[f"state_{i}"] = _np(st)
        tp, fp, tn, fn = st
        for name, fn_ in reducers.items():
            if st.shape[1] == 1:
                red[f"{name}_{i}_binary"] = _np(fn_(tp, fp, tn, fn, average="binary"))
            for avg in AVERAGES:
                for ml in (False, True):
                    red[f"{name}_{i}_{avg}_{int(ml)}"] = _np(fn_(tp, fp, tn, fn, average=avg, multilabel=ml))
    np.savez_compressed(OUT / "f1_stat_reducers.npz", **red)

    # ---- confusion-matrix reducers ------------------------------------------------------------------------
    cm = {}
    mats = []
    for C in (2, 2, 3, 10, 150):
        mats.append(torch.randint(0, 5000, (C, C), generator=g))
    mats += [
        torch.tensor([[5, 0], [0, 7]]), torch.tensor([[0, 4], [9, 0]]), torch.tensor([[0, 0], [0, 0]]),
        torch.tensor([[6, 3], [0, 0]]), torch.tensor([[0, 3], [0, 8]]), torch.tensor([[6, 0], [2, 0]]),
        torch.tensor([[0, 0], [4, 3]]), torch.diag(torch.tensor([3, 0, 9, 1])),
        torch.tensor([[0, 5, 0], [0, 7, 0], [0, 2, 0]]), torch.randint(0, 3_000_000, (6, 6), generator=g),
        torch.randint(0, 900, (5, 2, 2), generator=g),  # multilabel state (MCC only)
    ]
    cm["n_mats"] = len(mats)
    for i, m in enumerate(mats):
        cm[f"mat_{i}"] = _np(m)
        cm[f"mcc_{i}"] = _np(_matthews_corrcoef_reduce(m).float())
        cm[f"mcc_dtype_{i}"] = np.array(str(_matthews_corrcoef_reduce(m).dtype))
        if m.ndim == 2:
            for w in ("none", "linear", "quadratic"):
                cm[f"kappa_{w}_{i}"] = _np(_cohen_kappa_reduce(m, w))
    np.savez_compressed(OUT / "f1_confmat_reducers.npz", **cm)

    # ---- operating points on curves -----------------------------------------------------------------------
    fp_ = {}
    curves = []
    for n in (1, 5, 40, 400):
        for quant in (None, 4):
            p, r = torch.rand(n + 1, generator=g), torch.rand(n + 1, generator=g)
            t = torch.sort(torch.rand(n, generator=g)).values
            if quant:
                p, r = (p * quant).round() / quant, (r * quant).round() / quant
            curves.append((p, r, t))
    curves.append((torch.zeros(6), torch.zeros(6), torch.linspace(0, 1, 5)))
    fp_["n_curves"] = len(curves)
    levels = (0.0, 0.25, 0.5, 0.9, 1.0)
    fp_["levels"] = np.array(levels, dtype=np.float64)
    for i, (p, r, t) in enumerate(curves):
        fp_[f"p_{i}"], fp_[f"r_{i}"], fp_[f"t_{i}"] = _np(p), _np(r), _np(t)
        for j, lv in enumerate(levels):
            a, b = _recall_at_precision(p, r, t, lv)
            fp_[f"rap_{i}_{j}"] = np.array([float(a), float(b)], dtype=np.float64)
            a, b = _precision_at_recall(p, r, t, lv)
            fp_[f"par_{i}_{j}"] = np.array([float(a), float(b)], dtype=np.float64)
            # ROC style: specificity / sensitivity / thresholds of equal length
            a, b = _specificity_at_sensitivity(p[: len(t)], r[: len(t)], t, lv)
            fp_[f"sas_{i}_{j}"] = np.array([float(a), float(b)], dtype=np.float64)
    np.savez_compressed(OUT / "f1_fixed_points.npz", **fp_)

    # ---- end to end ---------------------------------------------------------------------------------------
    e = {}
    n = 3000
    bp, bt = torch.rand(n, generator=g), torch.randint(2, (n,), generator=g)
    bl = torch.randn(n, generator=g) * 2
    bti = bt.clone()
    bti[torch.rand(n, generator=g) < 0.15] = -1
    C = 7
    mp, mt = torch.randn(n, C, generator=g), torch.randint(C, (n,), generator=g)
    mti = mt.clone()
    mti[torch.rand(n, generator=g) < 0.15] = 3
    L = 5
    lp, lt = torch.rand(n, L, generator=g), torch.randint(2, (n, L), generator=g)
    ll = torch.randn(n, L, generator=g) * 2
    lti = lt.clone()
    lti[torch.rand(n, L, generator=g) < 0.15] = -1
    for k, v in dict(bp=bp, bt=bt, bl=bl, bti=bti, mp=mp, mt=mt, mti=mti, lp=lp, lt=lt, ll=ll, lti=lti).items():
        e[k] = _np(v)
    e["C"], e["L"] = C, L
    for stem in ("precision", "recall", "specificity", "hamming_distance"):
        e[f"binary_{stem}"] = _np(getattr(rf, f"binary_{stem}")(bp, bt))
        e[f"binary_{stem}_logits_ign"] = _np(getattr(rf, f"binary_{stem}")(bl, bti, threshold=0.3, ignore_index=-1))
        for avg in AVERAGES:
            e[f"multiclass_{stem}_{avg}"] = _np(getattr(rf, f"multiclass_{stem}")(mp, mt, C, average=avg))
            e[f"multiclass_{stem}_{avg}_ign"] = _np(
                getattr(rf, f"multiclass_{stem}")(mp, mti, C, average=avg, ignore_index=3))
            e[f"multilabel_{stem}_{avg}"] = _np(getattr(rf, f"multilabel_{stem}")(lp, lt, L, average=avg))
            e[f"multilabel_{stem}_{avg}_logits_ign"] = _np(
                getattr(rf, f"multilabel_{stem}")(ll, lti, L, threshold=0.6, average=avg, ignore_index=-1))
    for w in (None, "linear", "quadratic"):
        e[f"binary_cohen_kappa_{w}"] = _np(rf.binary_cohen_kappa(bp, bt, weights=w))
        e[f"multiclass_cohen_kappa_{w}"] = _np(rf.multiclass_cohen_kappa(mp, mt, C, weights=w))
        e[f"multiclass_cohen_kappa_{w}_ign"] = _np(rf.multiclass_cohen_kappa(mp, mti, C, weights=w, ignore_index=3))
    e["binary_mcc"] = _np(rf.binary_matthews_corrcoef(bp, bt))
    e["binary_mcc_logits_ign"] = _np(rf.binary_matthews_corrcoef(bl, bti, ignore_index=-1))
    e["multiclass_mcc"] = _np(rf.multiclass_matthews_corrcoef(mp, mt, C))
    e["multiclass_mcc_ign"] = _np(rf.multiclass_matthews_corrcoef(mp, mti, C, ignore_index=3))
    e["multilabel_mcc"] = _np(rf.multilabel_matthews_corrcoef(lp, lt, L))
    e["multilabel_mcc_logits_ign"] = _np(rf.multilabel_matthews_corrcoef(ll, lti, L, ignore_index=-1))
    for thr in (None, 20):
        for lv in (0.3, 0.6):
            tag = f"{'exact' if thr is None else 'binned'}_{lv}"
            for stem, kw in (("recall_at_fixed_precision", "min_precision"), ("precision_at_fixed_recall", "min_recall"),
                             ("specificity_at_sensitivity", "min_sensitivity")):
                a, b = getattr(rf, f"binary_{stem}")(bl, bt, **{kw: lv}, thresholds=thr)
                e[f"binary_{stem}_{tag}"] = np.stack([_np(a), _np(b)]).astype(np.float64)
                a, b = getattr(rf, f"multilabel_{stem}")(ll, lti, L, **{kw: lv}, thresholds=thr, ignore_index=-1)
                e[f"multilabel_{stem}_{tag}"] = np.stack([_np(a), _np(b)]).astype(np.float64)
    np.savez_compressed(OUT / "f1_end_to_end.npz", **e)
    print("wrote", sorted(p.name for p in OUT.glob("f1_*.npz")))


if __name__ == "__main__":
    main()
