"""Tiny drivers for ncu: run one kernel family a few times on the BASELINE shape.

    python profiles/drive.py k1   # config 2: MulticlassConfusionMatrix(1000), bf16 [2^22, 1000]
    python profiles/drive.py k2   # config 3: BinaryPrecisionRecallCurve(thresholds=200), 2^28 fp32 logits
    python profiles/drive.py k5   # config 5: MulticlassJaccardIndex(150), [16,150,512,512]
    python profiles/drive.py k3   # config 4 (scaled): MultilabelAUROC exact, compute()
"""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))

which = sys.argv[1] if len(sys.argv) > 1 else "k1"
reps = int(sys.argv[2]) if len(sys.argv) > 2 and which in ("k1", "k2", "k5") else 4
dev = torch.device("cuda")
g = torch.Generator(device=dev).manual_seed(0)

if which == "k1":
    from torchmetrics_forked_b200.classification import MulticlassConfusionMatrix

    rows = 1 << 22
    t = torch.randint(1000, (rows,), generator=g, device=dev)
    p = torch.empty(rows, 1000, dtype=torch.bfloat16, device=dev)
    for s in range(0, rows, 1 << 18):
        p[s : s + (1 << 18)] = torch.randn(1 << 18, 1000, generator=g, device=dev).bfloat16()
    m = MulticlassConfusionMatrix(1000, validate_args=False).to(dev)
    for _ in range(reps):
        m.update(p, t)
elif which == "k2":
    from torchmetrics_forked_b200.classification import BinaryPrecisionRecallCurve

    n = 1 << int(__import__("os").environ.get("LOG2N", "28"))
    p = torch.randn(n, generator=g, device=dev) * 3
    t = torch.randint(2, (n,), generator=g, device=dev)
    m = BinaryPrecisionRecallCurve(thresholds=200, validate_args=False).to(dev)
    for _ in range(reps):
        m.update(p, t)
    if "TIME" in __import__("os").environ:
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(20):
            m.update(p, t)
        b.record()
        torch.cuda.synchronize()
        print("k2 log2n", __import__("os").environ.get("LOG2N", "28"), "ms/update", a.elapsed_time(b) / 20)
elif which == "k5":
    from torchmetrics_forked_b200.classification import MulticlassJaccardIndex

    p = torch.randn(16, 150, 512, 512, generator=g, device=dev)
    t = torch.randint(150, (16, 512, 512), generator=g, device=dev)
    m = MulticlassJaccardIndex(150, validate_args=False).to(dev)
    for _ in range(reps):
        m.update(p, t)
elif which == "k3":
    from torchmetrics_forked_b200.classification import MultilabelAUROC

    n, L = 1 << 20, 100
    m = MultilabelAUROC(L, thresholds=None, validate_args=False).to(dev)
    for _ in range(2):
        m.update(torch.randn(n, L, generator=g, device=dev), torch.randint(2, (n, L), generator=g, device=dev))
    print(float(m.compute()))
elif which == "f2":
    from torchmetrics_forked_b200.classification import MulticlassAUROC

    n, C = (1 << 21, 100) if len(sys.argv) < 4 else (int(sys.argv[2]), int(sys.argv[3]))
    p = torch.randn(n, C, generator=g, device=dev)
    if __import__("os").environ.get("DT") == "bf16":
        p = p.bfloat16()
    t = torch.randint(C, (n,), generator=g, device=dev)
    m = MulticlassAUROC(C, thresholds=200, validate_args=False).to(dev)
    for _ in range(3):
        m.update(p, t)
elif which == "f2x":
    from torchmetrics_forked_b200.classification import MulticlassAUROC

    n, C = (1 << 21, 100) if len(sys.argv) < 4 else (int(sys.argv[2]), int(sys.argv[3]))
    p = torch.randn(n, C, generator=g, device=dev)
    if __import__("os").environ.get("DT") == "bf16":
        p = p.bfloat16()
    t = torch.randint(C, (n,), generator=g, device=dev)
    m = MulticlassAUROC(C, thresholds=None, validate_args=False).to(dev)
    for _ in range(3):
        m.reset()
        m.update(p, t)
    if n * C <= 1 << 28:
        print(float(m.compute()))
elif which == "k1c":  # argmax + histogram at other class counts: python profiles/drive.py k1c bf16|fp32 C
    from torchmetrics_forked_b200.classification import MulticlassAccuracy

    C = int(sys.argv[3])
    rows = (1 << 30) // C // 1024 * 1024
    t = torch.randint(C, (rows,), generator=g, device=dev)
    p = torch.randn(rows, C, generator=g, device=dev).bfloat16()
    if sys.argv[2] == "fp32":
        p = p.float()[: rows // 2]
        t = t[: rows // 2]
    m = MulticlassAccuracy(C, average="macro", validate_args=False).to(dev)
    for _ in range(3):
        m.update(p, t)
elif which == "f3":
    from torchmetrics_forked_b200.classification import MulticlassAccuracy

    C = int(sys.argv[3]) if len(sys.argv) > 3 else 1000
    rows = (1 << 30) // C // 1024 * 1024
    t = torch.randint(C, (rows,), generator=g, device=dev)
    p = torch.randn(rows, C, generator=g, device=dev).bfloat16()
    if len(sys.argv) > 2 and sys.argv[2] == "fp32":
        p = p.float()[: rows // 2]
        t = t[: rows // 2]
    m = MulticlassAccuracy(C, top_k=5, average="micro", validate_args=False).to(dev)
    for _ in range(3):
        m.update(p, t)
elif which == "evid":  # round 2 evidence set: the kernels of rows a1-a12 that had no --set full capture yet
    from torchmetrics_forked_b200.classification import (
        BinaryAccuracy,
        MulticlassJaccardIndex,
        MultilabelAUROC,
        MultilabelPrecisionRecallCurve,
        MultilabelStatScores,
    )

    t5 = torch.randint(150, (16, 512, 512), generator=g, device=dev)
    for dt in (torch.float32, torch.bfloat16):  # k1_strided<f32>, k1_strided<bf16> (config 5)
        p5 = torch.randn(16, 150, 512, 512, generator=g, device=dev).to(dt)
        m = MulticlassJaccardIndex(150, validate_args=False).to(dev)
        for _ in range(2):
            m.update(p5, t5)
        del p5, m
    n = 1 << 28  # k2_single, binary (K1b): BinaryAccuracy on config 3's logits
    p = torch.randn(n, generator=g, device=dev) * 3
    t = torch.randint(2, (n,), generator=g, device=dev)
    m = BinaryAccuracy(validate_args=False).to(dev)
    for _ in range(2):
        m.update(p, t)
    del p, t, m
    nl, L = 1 << 21, 100  # k2_single multilabel, k2_multi, k3_normalize_vec on [2^21, 100]
    pl = torch.randn(nl, L, generator=g, device=dev)
    tl = torch.randint(2, (nl, L), generator=g, device=dev)
    for m in (MultilabelStatScores(L, validate_args=False), MultilabelPrecisionRecallCurve(L, thresholds=200, validate_args=False),
              MultilabelAUROC(L, thresholds=None, validate_args=False)):
        m = m.to(dev)
        for _ in range(2):
            m.update(pl, tl)
elif which == "evid2":  # exact-mode update kernels: k3_normalize_vec ([2^21,100] sigmoid), k5_softmax_rows / _small
    from torchmetrics_forked_b200.classification import MulticlassAUROC, MultilabelAUROC

    nl, L = 1 << 21, 100
    pl = torch.randn(nl, L, generator=g, device=dev)
    tl = torch.randint(2, (nl, L), generator=g, device=dev)
    m = MultilabelAUROC(L, thresholds=None, validate_args=False).to(dev)
    for _ in range(2):
        m.reset()
        m.update(pl, tl)
    tc = torch.randint(L, (nl,), generator=g, device=dev)
    m = MulticlassAUROC(L, thresholds=None, validate_args=False).to(dev)
    for _ in range(2):
        m.reset()
        m.update(pl, tc)
    del pl, tl, m
    ps = torch.randn(1 << 24, 10, generator=g, device=dev)
    ts = torch.randint(10, (1 << 24,), generator=g, device=dev)
    m = MulticlassAUROC(10, thresholds=None, validate_args=False).to(dev)
    for _ in range(2):
        m.reset()
        m.update(ps, ts)
torch.cuda.synchronize()
if which in ("f2", "f2x", "f3", "k1c") and "TIME" in __import__("os").environ:  # quick timing while iterating (not a bench value)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(20):
        if which == "f2x":
            m.reset()
        m.update(p, t)
    b.record()
    torch.cuda.synchronize()
    print(which, sys.argv[2:], "ms/update", a.elapsed_time(b) / 20)
print("done", which)
