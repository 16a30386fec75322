"""Golden vectors added in round 2, from the UNMODIFIED reference (CPU, build container).

TEST INFRASTRUCTURE.  Usage (/root/reference must be mounted):   python oracle/gen_golden_r2.py

`switch16_*`: 16-bit scores around the reference's loop / vectorised switch of the binned curves
(functional/classification/precision_recall_curve.py:203 `preds.numel() <= 50_000`, :473 `preds.numel() * num_classes
<= 1_000_000`, both evaluated AFTER the ignore_index filter).  The loop compares against thresholds rounded to the score
dtype (0-d compare), the vectorised form against float32 thresholds, so the two differ for bf16 / fp16 scores that sit
on a rounded threshold; the scores here are drawn from the bf16 / fp16 grid near the thresholds to make that happen.
Scores are stored as float32 (exact) with the dtype name beside them.
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

    g = torch.Generator().manual_seed(20261107)

    def near_thresholds(n, T, dtype):
        """Probabilities that often coincide with a threshold rounded to `dtype` (and sometimes with the exact one)."""
        thr = torch.linspace(0, 1, T)
        pick = thr[torch.randint(T, (n,), generator=g)]
        jitter = (torch.rand(n, generator=g) - 0.5) * 0.004
        return (pick + torch.where(torch.rand(n, generator=g) < 0.5, torch.zeros(n), jitter)).clamp(0, 1).to(dtype)

    def binary(name, n, dtype, ignore_frac):
        T = 200
        preds = near_thresholds(n, T, dtype)
        target = torch.randint(2, (n,), generator=g)
        ign = None
        if ignore_frac:
            ign = -1
            target[torch.rand(n, generator=g) < ignore_frac] = -1
        p, r, t = rf.binary_precision_recall_curve(preds, target, thresholds=T, ignore_index=ign, validate_args=False)
        from torchmetrics.functional.classification.precision_recall_curve import (
            _binary_precision_recall_curve_format,
            _binary_precision_recall_curve_update,
        )

        pf, tf, thr = _binary_precision_recall_curve_format(preds, target, T, ign)
        state = _binary_precision_recall_curve_update(pf, tf, thr)
        np.savez_compressed(
            OUT / f"switch16_binary_{name}.npz", preds=preds.float().numpy(), dtype=str(dtype).split(".")[1],
            target=target.numpy(), thresholds=T, ignore_index=-999 if ign is None else ign, state=state.numpy(),
            kept=int((target != -1).sum()),
        )

    binary("bf16_loop_60000", 60_000, torch.bfloat16, 0.0)
    binary("bf16_ignore_kept_below_switch", 60_000, torch.bfloat16, 0.25)   # 60 000 rows, ~45 000 kept: vectorised
    binary("bf16_ignore_kept_above_switch", 80_000, torch.bfloat16, 0.25)   # ~60 000 kept: loop
    binary("f16_ignore_kept_below_switch", 60_000, torch.float16, 0.25)
    binary("bf16_vectorised_40000", 40_000, torch.bfloat16, 0.0)

    def multiclass(name, m, C, dtype, ignore_frac):
        T = 100
        x = near_thresholds(m * C, T, torch.float32).reshape(m, C)
        x = (x / x.sum(1, keepdim=True)).to(dtype)  # rows that look like probabilities (inside [0,1]: no softmax)
        target = torch.randint(C, (m,), generator=g)
        ign = None
        if ignore_frac:
            ign = -1
            target[torch.rand(m, generator=g) < ignore_frac] = -1
        from torchmetrics.functional.classification.precision_recall_curve import (
            _multiclass_precision_recall_curve_format,
            _multiclass_precision_recall_curve_update,
        )

        pf, tf, thr = _multiclass_precision_recall_curve_format(x, target, C, T, ign)
        state = _multiclass_precision_recall_curve_update(pf, tf, C, thr)
        np.savez_compressed(
            OUT / f"switch16_multiclass_{name}.npz", preds=x.float().numpy(), dtype=str(dtype).split(".")[1],
            target=target.numpy(), num_classes=C, thresholds=T, ignore_index=-999 if ign is None else ign,
            state=state.numpy(), kept=int((target != -1).sum()),
        )

    multiclass("bf16_ignore_kept_below_switch", 3000, 20, torch.bfloat16, 0.25)  # 3000*400 = 1.2 M, ~0.9 M kept
    multiclass("bf16_loop", 3000, 20, torch.bfloat16, 0.0)
    print("wrote", sorted(p.name for p in OUT.glob("switch16_*.npz")))


if __name__ == "__main__":
    main()
