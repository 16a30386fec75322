"""GPU parity: K5 (multiclass curves) — the conditional row softmax must return torch's bits."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16])
def test_softmax_rows_bits_equal_torch(dtype):
    """Every lane layout of torch's persistent warp softmax: C from 1 to 1024 (all powers of two, their
    neighbours, and a spread in between), ragged row counts, large-magnitude logits, -inf entries."""
    from torchmetrics_forked_b200 import kernels

    g = torch.Generator(device="cuda").manual_seed(5)
    widths = sorted({1, 2, 3, 4, 5, 7, 8, 9, 15, 16, 17, 31, 32, 33, 63, 64, 65, 100, 127, 128, 129, 150, 255, 256,
                     257, 500, 511, 512, 513, 777, 1000, 1023, 1024})
    for C in widths:
        rows = 1237 if C <= 128 else 301
        x = (torch.randn(rows, C, generator=g, device="cuda") * 6).to(dtype)
        x[0, 0] = 40.0
        if C > 2:
            x[1, C // 2] = float("-inf")
        got = kernels.softmax_normalize(x)
        assert torch.equal(got, x.softmax(1)), (C, dtype)


def test_softmax_normalize_decision_and_special_rows():
    from torchmetrics_forked_b200 import kernels

    g = torch.Generator(device="cuda").manual_seed(6)
    probs = torch.rand(999, 10, generator=g, device="cuda")
    out = kernels.softmax_normalize(probs)
    assert torch.equal(out, probs) and out.data_ptr() != probs.data_ptr()  # all inside [0,1]: a copy
    spiked = probs.clone()
    spiked[500, 3] = 1.0001
    assert torch.equal(kernels.softmax_normalize(spiked), spiked.softmax(1))  # one value flips every row (H1)
    # ... unless its row is ignored
    target = torch.zeros(999, dtype=torch.long, device="cuda")
    target[500] = -1
    assert torch.equal(kernels.softmax_normalize(spiked, target, -1), spiked)
    assert torch.equal(kernels.softmax_normalize(spiked, target, -2), spiked.softmax(1))
    # NaN and +inf rows come out all-NaN exactly like torch's
    weird = torch.randn(64, 33, generator=g, device="cuda")
    weird[3, 5] = float("nan")
    weird[7, 0] = float("inf")
    got, want = kernels.softmax_normalize(weird), weird.softmax(1)
    assert torch.equal(got.isnan(), want.isnan()) and torch.equal(got.nan_to_num(), want.nan_to_num())
    # wide rows keep torch's kernel; empty input
    wide = torch.randn(17, 3000, generator=g, device="cuda")
    assert torch.equal(kernels.softmax_normalize(wide), wide.softmax(1))
    assert kernels.softmax_normalize(probs[:0]).shape == (0, 10)
