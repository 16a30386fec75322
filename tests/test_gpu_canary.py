"""Stand-in for compute-sanitizer (closed on this GPU pool, see profiles/r02_sanitizer_summary.txt): guard regions
around every buffer the kernels write, and run-to-run repeatability of every kernel family on ragged shapes.

* Guards: states, scratch and outputs are views into the middle of larger allocations filled with a canary pattern;
  after the launches the canaries must be untouched.
* Repeatability: three runs from a zeroed state on the same input must agree bit for bit (integer states, float
  outputs alike) — the kernels use lane-private counters, staged shared-memory tiles, warp votes and device flags, and
  a race in any of them shows as a difference between runs.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu

GUARD = 1024  # elements on either side


def _guarded(shape, dtype, fill):
    n = 1
    for s in shape:
        n *= s
    big = torch.full((n + 2 * GUARD,), fill, dtype=dtype, device="cuda")
    view = big[GUARD : GUARD + n].view(*shape)
    view.zero_()
    return big, view


def _guards_intact(big, n, fill):
    return bool((big[:GUARD] == fill).all()) and bool((big[GUARD + n :] == fill).all())


def _with_guarded_states(metric):
    """Re-home every tensor state of `metric` (and its defaults) in a canary-padded allocation."""
    keep = []
    for name, default in metric._defaults.items():
        cur = getattr(metric, name)
        if isinstance(cur, torch.Tensor):
            fill = 0x5A5A5A5A5A5A if cur.dtype == torch.int64 else 77
            big, view = _guarded(tuple(cur.shape), cur.dtype, fill)
            setattr(metric, name, view)
            keep.append((big, cur.numel(), fill))
    return keep


def _cases():
    from torchmetrics_forked_b200 import classification as M

    g = torch.Generator(device="cuda").manual_seed(123)
    rn = lambda *s: torch.randn(*s, generator=g, device="cuda")  # noqa: E731
    ri = lambda hi, *s: torch.randint(hi, s, generator=g, device="cuda")  # noqa: E731
    yield "k1_rows_bf16_C1000", lambda: M.MulticlassConfusionMatrix(1000, validate_args=False), (rn(4099, 1000).bfloat16(), ri(1000, 4099))
    yield "k1_rows_f32_C100", lambda: M.MulticlassStatScores(100, average=None, validate_args=False), (rn(30_001, 100), ri(100, 30_001))
    yield "k1_thread_C5", lambda: M.MulticlassConfusionMatrix(5, validate_args=False), (rn(100_003, 5), ri(5, 100_003))
    yield "k1_strided_C150", lambda: M.MulticlassJaccardIndex(150, validate_args=False), (rn(2, 150, 33, 47), ri(150, 2, 33, 47))
    yield "k1_topk", lambda: M.MulticlassAccuracy(1000, top_k=5, average="micro", validate_args=False), (rn(3001, 1000).bfloat16(), ri(1000, 3001))
    yield "k2_binary_lanes", lambda: M.BinaryPrecisionRecallCurve(thresholds=200, validate_args=False), (rn(1_000_003) * 3, ri(2, 1_000_003))
    yield "k2_single", lambda: M.BinaryStatScores(validate_args=False), (rn(1_000_003), ri(2, 1_000_003))
    yield "k2_multi", lambda: M.MultilabelPrecisionRecallCurve(100, thresholds=200, validate_args=False), (rn(10_007, 100), ri(2, 10_007, 100))
    yield "k2_generic_inner", lambda: M.MultilabelPrecisionRecallCurve(3, thresholds=17, validate_args=False), (rn(501, 3, 7), ri(2, 501, 3, 7))
    yield "k2_mc_rows", lambda: M.MulticlassPrecisionRecallCurve(100, thresholds=200, validate_args=False), (rn(20_011, 100), ri(100, 20_011))
    yield "k2_mc_small", lambda: M.MulticlassPrecisionRecallCurve(10, thresholds=50, validate_args=False), (rn(50_021, 10), ri(10, 50_021))


CASE_NAMES = ["k1_rows_bf16_C1000", "k1_rows_f32_C100", "k1_thread_C5", "k1_strided_C150", "k1_topk", "k2_binary_lanes",
              "k2_single", "k2_multi", "k2_generic_inner", "k2_mc_rows", "k2_mc_small"]


@pytest.mark.parametrize("name", CASE_NAMES)
def test_state_kernels_keep_inside_their_buffers_and_repeat(name):
    make, args = next((m, a) for n, m, a in _cases() if n == name)
    runs = []
    for _ in range(3):
        metric = make().cuda()
        guards = _with_guarded_states(metric)
        metric.update(*args)
        metric.update(*args)
        torch.cuda.synchronize()
        for big, n, fill in guards:
            assert _guards_intact(big, n, fill), f"{name}: a kernel wrote outside a state tensor"
        runs.append({k: getattr(metric, k).clone() for k in metric._defaults})
    for k in runs[0]:
        assert torch.equal(runs[0][k], runs[1][k]) and torch.equal(runs[0][k], runs[2][k]), (name, k)


def test_exact_curve_pipeline_keeps_inside_its_buffers_and_repeats():
    """K3 (sigmoid normalise, pack, four sort passes, scan, emit / scores) and K5 (row softmax) through the raw entry
    points, with canaries around scratch and outputs."""
    from torchmetrics_forked_b200 import _native as nat
    from torchmetrics_forked_b200 import kernels

    lib = nat.lib()
    g = torch.Generator(device="cuda").manual_seed(7)
    n, L = 70_003, 5
    preds = torch.randn(n, L, generator=g, device="cuda")
    preds[:, 1] = (preds[:, 1] * 4).round() / 4
    target = torch.randint(2, (n, L), generator=g, device="cuda")
    err = torch.zeros(1, dtype=torch.int32, device="cuda")
    outs = []
    for _ in range(3):
        stored = kernels.sigmoid_normalize(preds, None, None, err)
        nbytes = int(lib.tmb_clf_curve_scratch_bytes(n, L))
        big_s, scratch = _guarded((nbytes,), torch.uint8, 0xA5)
        bigs = [_guarded((L, n), torch.float32, -7.0) for _ in range(3)]
        big_c, counts = _guarded((L,), torch.int64, 0x5A5A)
        p = lambda t: t.data_ptr()  # noqa: E731
        rc = lib.tmb_clf_curves(p(stored), 0, p(target), 0, n, L, 0, 0, 1, p(scratch), nbytes, p(bigs[0][1]), p(bigs[1][1]),
                                p(bigs[2][1]), p(counts), p(err), torch.cuda.current_stream().cuda_stream)
        assert rc == 0
        torch.cuda.synchronize()
        assert _guards_intact(big_s, nbytes, 0xA5), "K3 wrote outside its scratch"
        assert _guards_intact(big_c, L, 0x5A5A)
        for big, view in bigs:
            assert _guards_intact(big, L * n, -7.0), "K3 wrote outside a curve output"
        lens = counts.tolist()
        scores = kernels.clf_curve_scores(stored, target, None, 1)
        sm = kernels.softmax_normalize(torch.randn(9973, 100, generator=torch.Generator(device="cuda").manual_seed(1), device="cuda"))
        outs.append((stored.clone(), [bigs[i][1][l, : lens[l]].clone() for i in range(3) for l in range(L)], scores.clone(), sm))
    for a, b in ((outs[0], outs[1]), (outs[0], outs[2])):
        assert torch.equal(a[0], b[0]), "sigmoid normalise differs between runs"
        assert torch.equal(a[3], b[3]), "row softmax differs between runs"
        assert all(torch.equal(x, y) for x, y in zip(a[1], b[1])), "curve arrays differ between runs"
        # columns 0, 2, 3 (trapezoid area in integer arithmetic, the two totals) are exact; column 1 is a float64 sum
        # of per-tile partial sums added with atomics in whatever order the tiles finish: equal to ~1 ulp of the sum
        assert torch.equal(a[2][:, [0, 2, 3]], b[2][:, [0, 2, 3]]), "integer reductions differ between runs"
        torch.testing.assert_close(a[2][:, 1], b[2][:, 1], rtol=1e-13, atol=0)
    assert int(err.item()) == 0
