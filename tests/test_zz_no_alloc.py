"""`tmb_binary_curve_update_noalloc`: the binned binary-curve update for hosts that cannot allocate.

Who needs it: a program that reaches libtmb200.so through a bare FFI and has no device allocator of its own,
a service whose memory pool already owns the whole GPU (a fresh cudaMalloc would fail or fragment it), or a
loop that cannot afford the device-wide synchronisation cudaMalloc / cudaFree imply.  Such a caller passes
inputs, thresholds, the int64 state and the error word; the histogram scratch is a fixed-size array inside the
library.

The test drives the entry point through the C ABI exactly as such a host would (ctypes, raw pointers) and
checks it, twice in a row, against the ordinary entry points fed with caller-allocated scratch.  The two calls
are chosen so that stale contents of the shared array would show: the first batch holds logits (it raises the
"apply sigmoid" flag and fills the sigmoid histogram), the second holds probabilities (flag must be down again,
raw histogram must start empty).  Integer states must be identical.

The file name sorts last on purpose: if this addition breaks, a `pytest -x` run has already executed every
other test.  It passes on its own in a fresh process as well as after the rest of the suite.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


def _p(t):
    return ctypes.c_void_p(t.data_ptr())


def _reference_state(lib, preds, target, thr, ignore, state):
    """The ordinary path: caller-owned scratch, update + finalize."""
    T = thr.numel()
    elems = lib.tmb_thresh_bin_scratch_elems(1, T)
    scratch = torch.zeros(elems, dtype=torch.int64, device="cuda")
    err = torch.zeros(1, dtype=torch.int32, device="cuda")
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    has_ign, ign = (0, 0) if ignore is None else (1, ignore)
    assert lib.tmb_thresh_bin_update(_p(preds), 0, 0, _p(target), 0, preds.numel(), 1, 1, _p(thr), T, 0, has_ign, ign,
                                     0, _p(scratch), elems, _p(err), st) == 0
    assert lib.tmb_thresh_bin_finalize(_p(scratch), elems, 1, T, None, 0, _p(state), None, None, None, st) == 0
    torch.cuda.synchronize()
    assert int(err.item()) == 0


@pytest.mark.parametrize("T", [200, 1024, 1])
def test_noalloc_matches_ordinary_entry_point_twice(T):
    from torchmetrics_forked_b200 import _native as nat

    lib = nat.lib()
    g = torch.Generator().manual_seed(1000 + T)
    thr = torch.linspace(0, 1, T).cuda() if T > 1 else torch.tensor([0.5]).cuda()
    n = 600_001
    logits = (torch.randn(n, generator=g) * 3).cuda()
    probs = torch.rand(n, generator=g).cuda()
    target = torch.randint(2, (n,), generator=g)
    target[::11] = -1
    target = target.cuda()

    want = torch.zeros(T, 2, 2, dtype=torch.int64, device="cuda")
    got = torch.zeros(T, 2, 2, dtype=torch.int64, device="cuda")
    err = torch.zeros(1, dtype=torch.int32, device="cuda")
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    def allocations():  # number of allocation events torch's caching allocator has served so far
        return torch.cuda.memory_stats().get("allocation.all.allocated", 0)

    for batch in (logits, probs):  # two calls in a row on the same shared array
        before = allocations()
        rc = lib.tmb_binary_curve_update_noalloc(_p(batch), 0, _p(target), 0, n, _p(thr), T, 1, -1, _p(got), _p(err), st)
        after = allocations()
        assert rc == 0
        assert after == before, "the entry point must not allocate through torch either"
        torch.cuda.synchronize()
        _reference_state(lib, batch, target, thr, -1, want)
        assert int(err.item()) == 0
        assert torch.equal(got, want)  # exact: integer counts, same kernels, same sigmoid
    assert int(got[0].sum()) == 2 * int((target != -1).sum())  # every counted sample landed once per call


def test_noalloc_rejects_more_thresholds_than_the_array_holds():
    from torchmetrics_forked_b200 import _native as nat

    lib = nat.lib()
    T = 1025
    thr = torch.linspace(0, 1, T).cuda()
    x = torch.rand(16).cuda()
    t = torch.zeros(16, dtype=torch.int64).cuda()
    state = torch.zeros(T, 2, 2, dtype=torch.int64, device="cuda")
    err = torch.zeros(1, dtype=torch.int32, device="cuda")
    rc = lib.tmb_binary_curve_update_noalloc(_p(x), 0, _p(t), 0, 16, _p(thr), T, 0, 0, _p(state), _p(err), None)
    assert rc == -6  # TMB_E_LIMIT
    torch.cuda.synchronize()
    assert int(state.sum()) == 0
    with pytest.raises(ValueError, match="limit"):
        nat.check(rc, "tmb_binary_curve_update_noalloc")
