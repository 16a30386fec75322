"""Per-step latency of `update` on training-sized batches (256 samples), this package beside the unmodified reference on
the same GPU: wall clock per call over 300 calls after 30 warm-ups, one synchronize at the end (the calls are
asynchronous; what is measured is what a training loop pays per step).  Not a bench line.

    python profiles/small_batch_latency.py
"""
import sys
import time
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "oracle")]
import ref_loader  # noqa: E402

import torchmetrics_forked_b200 as ours  # noqa: E402

ref = ref_loader.load_staged_reference()
g = torch.Generator(device="cuda").manual_seed(0)
n = 256
cases = [
    ("MulticlassAccuracy(1000)", "MulticlassAccuracy", dict(num_classes=1000), torch.randn(n, 1000, generator=g, device="cuda"), torch.randint(1000, (n,), generator=g, device="cuda")),
    ("MulticlassConfusionMatrix(100)", "MulticlassConfusionMatrix", dict(num_classes=100), torch.randn(n, 100, generator=g, device="cuda"), torch.randint(100, (n,), generator=g, device="cuda")),
    ("BinaryF1Score()", "BinaryF1Score", {}, torch.randn(n, generator=g, device="cuda"), torch.randint(2, (n,), generator=g, device="cuda")),
    ("BinaryAUROC(thresholds=200)", "BinaryAUROC", dict(thresholds=200), torch.randn(n, generator=g, device="cuda"), torch.randint(2, (n,), generator=g, device="cuda")),
    ("BinaryAUROC(thresholds=None)", "BinaryAUROC", dict(thresholds=None), torch.randn(n, generator=g, device="cuda"), torch.randint(2, (n,), generator=g, device="cuda")),
    ("MultilabelF1Score(20)", "MultilabelF1Score", dict(num_labels=20), torch.randn(n, 20, generator=g, device="cuda"), torch.randint(2, (n, 20), generator=g, device="cuda")),
    ("MulticlassAUROC(10, thresholds=100)", "MulticlassAUROC", dict(num_classes=10, thresholds=100), torch.randn(n, 10, generator=g, device="cuda"), torch.randint(10, (n,), generator=g, device="cuda")),
]


def per_call_us(metric, p, t, calls=300):
    for _ in range(30):
        metric.update(p, t)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(calls):
        metric.update(p, t)
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / calls * 1e6


print(f"{'update of a batch of 256':40s} {'ours us':>9s} {'reference on the GPU us':>24s} {'validate_args':>14s}")
for label, cls, kw, p, t in cases:
    for validate in (False, True):
        a = per_call_us(getattr(ours.classification, cls)(**kw, validate_args=validate).cuda(), p, t)
        b = per_call_us(getattr(ref.classification, cls)(**kw, validate_args=validate).cuda(), p, t)
        print(f"{label:40s} {a:9.1f} {b:24.1f} {str(validate):>14s}")
