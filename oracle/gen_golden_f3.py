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
    print("wrote", sorted(p.name for p in OUT.glob("f3_*.npz")))


if __name__ == "__main__":
    main()
