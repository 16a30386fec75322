"""The float64 front end (torchmetrics_forked_b200/_f64.py) is plain torch: its arithmetic claims are checked here on
the CPU against numpy — the directed rounding keeps every comparison with a float32 threshold, the rank transform
keeps order and ties and finds the scores again."""
import numpy as np
import torch

from torchmetrics_forked_b200 import _f64


def _samples(seed, n=20000):
    g = torch.Generator().manual_seed(seed)
    grid = torch.linspace(0, 1, 200).double()
    near = grid[torch.randint(200, (n,), generator=g)] + (torch.randint(5, (n,), generator=g) - 2).double() * 1e-12
    wide = torch.randn(n, generator=g, dtype=torch.float64) * 1e3
    tiny = torch.rand(n, generator=g, dtype=torch.float64) * 1e-44  # below the float32 denormals
    exact = torch.rand(n, generator=g).double()  # float32 values: nothing to round
    return torch.cat([near.clamp(0, 1), wide, tiny, exact, torch.tensor([0.0, 1.0, -0.0, 3.5e38, -3.5e38, 1e39])])


def test_directed_rounding_is_the_neighbouring_float32():
    p = _samples(1)
    down, up = _f64.round_directed(p, up=False), _f64.round_directed(p, up=True)
    assert down.dtype == up.dtype == torch.float32
    assert bool((down.double() <= p).all()) and bool((up.double() >= p).all())
    representable = p.float().double() == p
    assert torch.equal(down[representable], up[representable])
    # otherwise the two are neighbours: nothing fits between them
    between = torch.nextafter(down, torch.full_like(down, float("inf")))
    assert torch.equal(between[~representable], up[~representable])


def test_directed_rounding_keeps_every_threshold_comparison():
    p = _samples(2)
    thresholds = torch.cat([torch.linspace(0, 1, 200), torch.tensor([0.3, 0.31, 1e-45, 0.5])])  # float32
    want_ge = p.unsqueeze(1) >= thresholds.double().unsqueeze(0)
    want_gt = p.unsqueeze(1) > thresholds.double().unsqueeze(0)
    assert torch.equal(_f64.round_directed(p, up=False).unsqueeze(1) >= thresholds.unsqueeze(0), want_ge)
    assert torch.equal(_f64.round_directed(p, up=True).unsqueeze(1) > thresholds.unsqueeze(0), want_gt)
    # plain rounding to nearest does NOT: that is why the front end exists
    assert not torch.equal(p.float().unsqueeze(1) >= thresholds.unsqueeze(0), want_ge)


def test_rank_transform_keeps_order_ties_and_values():
    g = torch.Generator().manual_seed(3)
    base = torch.rand(50, generator=g, dtype=torch.float64)
    cols = base[torch.randint(50, (4000, 3), generator=g)] + torch.randint(3, (4000, 3), generator=g).double() * 1e-13
    cols[7, 1] = float("nan")
    chunks = [cols[:1000], cols[1000:1001], cols[1001:]]
    ranked, values = _f64.rank_scores(chunks)
    assert ranked.dtype == torch.float32 and ranked.shape == cols.shape
    bits = ranked.view(torch.int32)
    ok = ~cols.isnan()
    assert bool(ranked[~ok].isnan().all())
    a, b = cols[ok], bits[ok].long()
    order = np.argsort(a.numpy(), kind="stable")
    assert bool((np.diff(b.numpy()[order]) >= 0).all())  # same order ...
    same_score = np.diff(a.numpy()[order]) == 0
    assert bool(((np.diff(b.numpy()[order]) == 0) == same_score).all())  # ... and exactly the same ties
    assert bool((ranked[ok] >= 0).all()) and bool(torch.isfinite(ranked[ok]).all())
    back = _f64.scores_of_ranks(ranked, values)
    assert torch.equal(back[ok], cols[ok]) and bool(back[~ok].isnan().all())
    parts = _f64.split_like(ranked, chunks)
    assert [p.shape[0] for p in parts] == [1000, 1, 2999]


def test_normalisation_decisions():
    probs = torch.tensor([[0.2, 0.8], [1.0, 0.0]], dtype=torch.float64)
    logits = torch.tensor([[0.2, 1.8], [1.0, 0.0]], dtype=torch.float64)
    target = torch.tensor([[1, 0], [0, -1]])
    assert torch.equal(_f64.sigmoid_where_needed(probs, None, None), probs)
    assert torch.equal(_f64.sigmoid_where_needed(logits, None, None), logits.sigmoid())
    # the only value outside [0, 1] sits at an ignored position: it takes no part in the decision
    masked = torch.tensor([[0.2, 0.8], [1.0, 7.0]], dtype=torch.float64)
    assert torch.equal(_f64.sigmoid_where_needed(masked, target, -1), masked)
    assert torch.equal(_f64.sigmoid_where_needed(masked, None, None), masked.sigmoid())
    with_nan = torch.tensor([0.5, float("nan")], dtype=torch.float64)
    assert torch.equal(_f64.sigmoid_where_needed(with_nan, None, None).isnan(), torch.tensor([False, True]))
    assert float(_f64.sigmoid_where_needed(with_nan, None, None)[0]) == float(torch.tensor(0.5, dtype=torch.float64).sigmoid())
    rows = torch.tensor([[0.1, 0.9], [3.0, -1.0]], dtype=torch.float64)
    assert torch.equal(_f64.softmax_where_needed(rows, torch.tensor([0, 5]), 5), rows)  # the logit row is ignored
    assert torch.equal(_f64.softmax_where_needed(rows, torch.tensor([0, 1]), 5), rows.softmax(1))
    assert torch.equal(_f64.binary_labels(logits, target, 0.7, None, True), (logits.sigmoid() > 0.7).long())
    marks = _f64.topk_marks(torch.tensor([[0.1, 0.7, 0.3, 0.5]], dtype=torch.float64), 2)
    assert marks.tolist() == [[0.0, 1.0, 0.0, 1.0]]
