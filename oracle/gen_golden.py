"""Generate the golden vectors under tests/golden/ by running the UNMODIFIED reference (CPU, this container).

TEST INFRASTRUCTURE.  Usage (build container only, /root/reference must be mounted):

    python oracle/gen_golden.py

Every fixture is a small .npz holding seeded inputs and the reference's outputs for one stage of the hot path
(SURVEY §8a rows a1-a14).  The fixtures are committed; the GPU box never needs the reference.
"""
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent))
from ref_loader import load_reference  # noqa: E402

OUT = Path(__file__).resolve().parent.parent / "tests" / "golden"


def _np(x):
    if isinstance(x, torch.Tensor):
        return x.detach().cpu().numpy()
    return np.asarray(x)


def main() -> None:
    tm = load_reference()
    if tm is None:
        raise SystemExit("reference not mounted at /root/reference")
    from torchmetrics.functional.classification import (
        binary_auroc,
        binary_average_precision,
        binary_confusion_matrix,
        binary_stat_scores,
        multiclass_accuracy,
        multiclass_confusion_matrix,
        multiclass_f1_score,
        multiclass_jaccard_index,
        multiclass_stat_scores,
        multilabel_auroc,
        multilabel_average_precision,
        multilabel_confusion_matrix,
        multilabel_stat_scores,
    )
    from torchmetrics.functional.classification.precision_recall_curve import (
        _binary_clf_curve,
        _binary_precision_recall_curve_format,
        _binary_precision_recall_curve_update,
        _multilabel_precision_recall_curve_format,
        _multilabel_precision_recall_curve_update,
    )

    OUT.mkdir(parents=True, exist_ok=True)
    g = torch.Generator().manual_seed(20261106)
    cases = {}

    # ---- K1: multiclass ------------------------------------------------------------------------------
    def mc_case(name, preds, target, C, ignore_index=None):
        d = {"preds": _np(preds.float() if preds.is_floating_point() else preds), "target": _np(target),
             "num_classes": C, "ignore_index": -999 if ignore_index is None else ignore_index}
        d["confmat"] = _np(multiclass_confusion_matrix(preds, target, C, ignore_index=ignore_index, validate_args=False))
        for avg in ("micro", "macro", "none"):
            d[f"stat_{avg}"] = _np(multiclass_stat_scores(preds, target, C, average=avg, ignore_index=ignore_index,
                                                          validate_args=False))
            d[f"acc_{avg}"] = _np(multiclass_accuracy(preds, target, C, average=avg, ignore_index=ignore_index,
                                                      validate_args=False))
            d[f"f1_{avg}"] = _np(multiclass_f1_score(preds, target, C, average=avg, ignore_index=ignore_index,
                                                     validate_args=False))
        for avg in ("micro", "macro", "weighted", "none"):
            d[f"jaccard_{avg}"] = _np(multiclass_jaccard_index(preds, target, C, average=avg,
                                                               ignore_index=ignore_index, validate_args=False))
        cases[name] = d

    p = torch.randn(257, 5, generator=g).softmax(-1)
    t = torch.randint(5, (257,), generator=g)
    mc_case("mc_probs_c5", p, t, 5)
    mc_case("mc_probs_c5_ignore", p, t, 5, ignore_index=2)
    mc_case("mc_probs_c5_ignore_outside", p, torch.where(torch.rand(257, generator=g) < 0.2, torch.tensor(-1), t), 5,
            ignore_index=-1)
    p = torch.randn(64, 10, 7, generator=g)
    t = torch.randint(10, (64, 7), generator=g)
    mc_case("mc_logits_c10_multidim", p, t, 10)
    p = torch.randn(96, 37, generator=g).bfloat16()
    t = torch.randint(37, (96,), generator=g)
    mc_case("mc_bf16_c37_ties", p, t, 37)  # bf16 rounding produces genuine ties
    p = torch.randn(40, 6, generator=g)
    p[3, 2] = float("nan")
    p[5, :] = float("-inf")
    p[7, 1] = p[7, 4] = 9.0
    p[9, 0] = float("nan")
    p[9, 5] = float("nan")
    p[11, :] = 0.25
    t = torch.randint(6, (40,), generator=g)
    mc_case("mc_nan_inf_ties", p, t, 6)
    mc_case("mc_int_preds", torch.randint(7, (300,), generator=g), torch.randint(7, (300,), generator=g), 7)
    mc_case("mc_int_preds_multidim_ignore", torch.randint(4, (30, 5), generator=g),
            torch.randint(4, (30, 5), generator=g), 4, ignore_index=0)
    p = torch.randn(33, 150, 4, 5, generator=g)
    t = torch.randint(150, (33, 4, 5), generator=g)
    mc_case("mc_c150_spatial", p, t, 150)
    p = torch.randn(50, 1000, generator=g).bfloat16()
    t = torch.randint(1000, (50,), generator=g)
    p[torch.arange(0, 50, 2), t[::2]] += 4
    mc_case("mc_bf16_c1000", p, t, 1000)

    # ---- K1b: binary / multilabel counts -----------------------------------------------------------------
    def bin_case(name, preds, target, ignore_index=None, threshold=0.5):
        cases[name] = {
            "preds": _np(preds), "target": _np(target), "threshold": threshold,
            "ignore_index": -999 if ignore_index is None else ignore_index,
            "stat": _np(binary_stat_scores(preds, target, threshold, ignore_index=ignore_index, validate_args=False)),
            "confmat": _np(binary_confusion_matrix(preds, target, threshold, ignore_index=ignore_index,
                                                   validate_args=False)),
        }

    t = torch.randint(2, (301,), generator=g)
    bin_case("bin_probs", torch.rand(301, generator=g), t)
    bin_case("bin_probs_thr03", torch.rand(301, generator=g), t, threshold=0.3)
    bin_case("bin_logits", torch.randn(301, generator=g) * 3, t)
    bin_case("bin_int", torch.randint(2, (301,), generator=g), t)
    tt = torch.where(torch.rand(301, generator=g) < 0.25, torch.tensor(-1), t)
    bin_case("bin_probs_ignore", torch.rand(301, generator=g), tt, ignore_index=-1)
    # H2: the only out-of-range pred sits at an ignored position -> stat scores take sigmoid, confmat does not
    p = torch.rand(64, generator=g)
    t2 = torch.randint(2, (64,), generator=g)
    t2[10] = -1
    p[10] = 3.5
    bin_case("bin_h2_flag_at_ignored", p, t2, ignore_index=-1)
    bin_case("bin_multidim", torch.randn(16, 3, 5, generator=g), torch.randint(2, (16, 3, 5), generator=g))

    def ml_case(name, preds, target, L, ignore_index=None, threshold=0.5):
        d = {"preds": _np(preds), "target": _np(target), "num_labels": L, "threshold": threshold,
             "ignore_index": -999 if ignore_index is None else ignore_index}
        for avg in ("micro", "macro", "none"):
            d[f"stat_{avg}"] = _np(multilabel_stat_scores(preds, target, L, threshold, average=avg,
                                                          ignore_index=ignore_index, validate_args=False))
        d["confmat"] = _np(multilabel_confusion_matrix(preds, target, L, threshold, ignore_index=ignore_index,
                                                       validate_args=False))
        cases[name] = d

    ml_case("ml_probs", torch.rand(130, 6, generator=g), torch.randint(2, (130, 6), generator=g), 6)
    ml_case("ml_logits", torch.randn(130, 6, generator=g) * 2, torch.randint(2, (130, 6), generator=g), 6)
    ml_case("ml_multidim", torch.rand(20, 4, 9, generator=g), torch.randint(2, (20, 4, 9), generator=g), 4)
    tm_ = torch.randint(2, (130, 6), generator=g)
    tm_[torch.rand(130, 6, generator=g) < 0.2] = -1
    ml_case("ml_probs_ignore", torch.rand(130, 6, generator=g), tm_, 6, ignore_index=-1)
    ml_case("ml_int", torch.randint(2, (90, 5), generator=g), torch.randint(2, (90, 5), generator=g), 5)

    # ---- K2: binned curves ---------------------------------------------------------------------------------
    def bcurve_case(name, preds, target, thresholds, ignore_index=None):
        p_, t_, thr = _binary_precision_recall_curve_format(preds, target, thresholds, ignore_index)
        state = _binary_precision_recall_curve_update(p_, t_, thr)
        cases[name] = {
            "preds": _np(preds), "target": _np(target), "thresholds": _np(thr), "state": _np(state),
            "ignore_index": -999 if ignore_index is None else ignore_index,
            "auroc": _np(binary_auroc(preds, target, thresholds=thresholds, ignore_index=ignore_index,
                                      validate_args=False)),
            "ap": _np(binary_average_precision(preds, target, thresholds=thresholds, ignore_index=ignore_index,
                                               validate_args=False)),
        }

    t = torch.randint(2, (4001,), generator=g)
    bcurve_case("bcurve_probs_t200", torch.rand(4001, generator=g), t, 200)
    bcurve_case("bcurve_probs_t5_exact_hits", torch.randint(0, 5, (4001,), generator=g).float() / 4, t, 5)
    bcurve_case("bcurve_logits_t200", torch.randn(4001, generator=g) * 3, t, 200)
    bcurve_case("bcurve_list_unsorted", torch.rand(4001, generator=g), t, [0.9, 0.1, 0.5, 0.5, 0.33])
    tt = torch.where(torch.rand(4001, generator=g) < 0.3, torch.tensor(-1), t)
    bcurve_case("bcurve_probs_ignore", torch.rand(4001, generator=g), tt, 50, ignore_index=-1)
    pn = torch.rand(500, generator=g)
    pn[::50] = float("nan")
    bcurve_case("bcurve_nan", pn, t[:500], 11)
    # above the reference's 50 000-sample switch to the per-threshold loop
    t = torch.randint(2, (60000,), generator=g)
    bcurve_case("bcurve_probs_loop_variant", torch.rand(60000, generator=g), t, 20)

    def mlcurve_case(name, preds, target, L, thresholds, ignore_index=None):
        p_, t_, thr = _multilabel_precision_recall_curve_format(preds, target, L, thresholds, ignore_index)
        state = _multilabel_precision_recall_curve_update(p_, t_, L, thr)
        d = {"preds": _np(preds), "target": _np(target), "num_labels": L, "thresholds": _np(thr),
             "state": _np(state), "ignore_index": -999 if ignore_index is None else ignore_index}
        for avg in ("micro", "macro", "none"):
            d[f"auroc_{avg}"] = _np(multilabel_auroc(preds, target, L, average=avg, thresholds=thresholds,
                                                     ignore_index=ignore_index, validate_args=False))
            d[f"ap_{avg}"] = _np(multilabel_average_precision(preds, target, L, average=avg, thresholds=thresholds,
                                                              ignore_index=ignore_index, validate_args=False))
        cases[name] = d

    mlcurve_case("mlcurve_probs", torch.rand(700, 5, generator=g), torch.randint(2, (700, 5), generator=g), 5, 33)
    mlcurve_case("mlcurve_logits", torch.randn(700, 5, generator=g) * 2, torch.randint(2, (700, 5), generator=g), 5, 33)
    tml = torch.randint(2, (700, 5), generator=g)
    tml[torch.rand(700, 5, generator=g) < 0.2] = -1
    mlcurve_case("mlcurve_probs_ignore", torch.rand(700, 5, generator=g), tml, 5, 17, ignore_index=-1)
    mlcurve_case("mlcurve_multidim", torch.rand(40, 3, 6, generator=g), torch.randint(2, (40, 3, 6), generator=g), 3, 9)

    # ---- K3: exact curves ------------------------------------------------------------------------------------
    def clf_case(name, preds, target):
        fps, tps, thr = _binary_clf_curve(preds, target)
        cases[name] = {
            "preds": _np(preds), "target": _np(target), "fps": _np(fps), "tps": _np(tps), "thr": _np(thr),
            "auroc": _np(binary_auroc(preds, target, validate_args=False)),
            "ap": _np(binary_average_precision(preds, target, validate_args=False)),
        }

    t = torch.randint(2, (3000,), generator=g)
    clf_case("clf_probs", torch.rand(3000, generator=g), t)
    clf_case("clf_ties", torch.randint(0, 40, (3000,), generator=g).float() / 39, t)
    clf_case("clf_all_equal", torch.full((100,), 0.5), t[:100])
    clf_case("clf_single", torch.tensor([0.3]), torch.tensor([1]))
    pz = torch.rand(200, generator=g)
    pz[::7] = 0.0
    pz[3::11] = -0.0
    pz[5::13] = 1.0
    clf_case("clf_zeros_ones", pz, t[:200])

    def mlexact_case(name, preds, target, L, ignore_index=None):
        d = {"preds": _np(preds), "target": _np(target), "num_labels": L,
             "ignore_index": -999 if ignore_index is None else ignore_index}
        p_, t_, _ = _multilabel_precision_recall_curve_format(preds, target, L, None, ignore_index)
        d["state_preds"] = _np(p_)
        for avg in ("micro", "macro", "none"):
            d[f"auroc_{avg}"] = _np(multilabel_auroc(preds, target, L, average=avg, thresholds=None,
                                                     ignore_index=ignore_index, validate_args=False))
            d[f"ap_{avg}"] = _np(multilabel_average_precision(preds, target, L, average=avg, thresholds=None,
                                                              ignore_index=ignore_index, validate_args=False))
        for i in range(L):
            pp, tt_ = p_[:, i], t_[:, i]
            if ignore_index is not None:
                keep = tt_ != ignore_index
                pp, tt_ = pp[keep], tt_[keep]
            fps, tps, thr = _binary_clf_curve(pp, tt_)
            d[f"fps_{i}"], d[f"tps_{i}"], d[f"thr_{i}"] = _np(fps), _np(tps), _np(thr)
        cases[name] = d

    mlexact_case("mlexact_probs", torch.rand(900, 4, generator=g), torch.randint(2, (900, 4), generator=g), 4)
    mlexact_case("mlexact_logits", torch.randn(900, 4, generator=g), torch.randint(2, (900, 4), generator=g), 4)
    tml = torch.randint(2, (900, 4), generator=g)
    tml[torch.rand(900, 4, generator=g) < 0.15] = -1
    mlexact_case("mlexact_probs_ignore", torch.rand(900, 4, generator=g), tml, 4, ignore_index=-1)

    for name, d in cases.items():
        np.savez_compressed(OUT / f"{name}.npz", **d)
    total = sum((OUT / f"{n}.npz").stat().st_size for n in cases)
    print(f"wrote {len(cases)} fixtures, {total / 1024:.0f} KiB, to {OUT}")


if __name__ == "__main__":
    main()
