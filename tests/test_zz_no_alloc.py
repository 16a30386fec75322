"""`tmb_binary_curve_update_noalloc`: the binned binary-curve update for hosts that cannot allocate.

Who needs it: a program that reaches libtmb200.so through a bare FFI and has no device allocator of its own,
a service whose memory pool already owns the whole GPU (a fresh cudaMalloc would fail or fragment it), or a
loop that cannot afford the device-wide synchronisation cudaMalloc / cudaFree imply.  Such a caller passes
inputs, thresholds, the int64 state and the error word; the histogram scratch is a fixed-size array inside the
library.

The test drives the entry point through the C ABI exactly as such a host would (ctypes, raw pointers) and
checks it, twice in a row, against the ordinary entry points fed with caller-allocated scratch and against the
CPU oracle (`oracle/hotpath.binary_curve_binned`, fed the device's own sigmoid as every logit parity test is).  The two calls
are chosen so that stale contents of the shared array would show: the first batch holds logits (it raises the
"apply sigmoid" flag and fills the sigmoid histogram), the second holds probabilities (flag must be down again,
raw histogram must start empty).  Integer states must be identical.

The file name sorts last on purpose: if this addition breaks, a `pytest -x` run has already executed every
other test.  It passes on its own in a fresh process as well as after the rest of the suite.
"""
import ctypes

import numpy as np
import pytest
import torch

import hotpath as O

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

    def device_free_bytes():  # cudaMemGetInfo: sees a cudaMalloc made by ANY runtime in the process, also the
        return torch.cuda.mem_get_info()[0]  # library's statically linked one, which torch's counters cannot

    sig = lambda x: torch.from_numpy(x).cuda().sigmoid().cpu().numpy()  # noqa: E731
    oracle = np.zeros((T, 2, 2), dtype=np.int64)
    lib.tmb_binary_curve_update_noalloc(_p(probs), 0, _p(target), 0, n, _p(thr), T, 1, -1, _p(got), _p(err), st)
    torch.cuda.synchronize()  # set-up call (address look-up, module load) before the memory readings
    got.zero_()
    for batch in (logits, probs):  # two calls in a row on the same shared array
        torch.cuda.synchronize()
        before, free_before = allocations(), device_free_bytes()
        rc = lib.tmb_binary_curve_update_noalloc(_p(batch), 0, _p(target), 0, n, _p(thr), T, 1, -1, _p(got), _p(err), st)
        torch.cuda.synchronize()
        after, free_after = allocations(), device_free_bytes()
        assert rc == 0
        assert after == before, "the entry point must not allocate through torch either"
        assert free_after >= free_before, "device-wide free memory shrank: something called cudaMalloc"
        _reference_state(lib, batch, target, thr, -1, want)
        assert int(err.item()) == 0
        assert torch.equal(got, want)  # exact: integer counts, same kernels, same sigmoid
        oracle += O.binary_curve_binned(batch.cpu().numpy(), target.cpu().numpy(), thr.cpu().numpy(), ignore_index=-1,
                                        sigmoid_fn=sig)
        assert np.array_equal(got.cpu().numpy(), oracle), "differs from the CPU oracle"
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


def test_noalloc_argument_errors_enqueue_nothing():
    """Bad dtype codes, a negative size or NULL inputs come back as negative codes before any kernel is launched."""
    from torchmetrics_forked_b200 import _native as nat

    lib = nat.lib()
    thr = torch.linspace(0, 1, 8).cuda()
    x, t = torch.rand(16).cuda(), torch.zeros(16, dtype=torch.int64).cuda()
    state = torch.zeros(8, 2, 2, dtype=torch.int64, device="cuda")
    err = torch.zeros(1, dtype=torch.int32, device="cuda")
    call = lib.tmb_binary_curve_update_noalloc
    assert call(_p(x), 9, _p(t), 0, 16, _p(thr), 8, 0, 0, _p(state), _p(err), None) == -1   # preds dtype
    assert call(_p(x), 0, _p(t), 9, 16, _p(thr), 8, 0, 0, _p(state), _p(err), None) == -1   # target dtype
    assert call(_p(x), 0, _p(t), 0, -1, _p(thr), 8, 0, 0, _p(state), _p(err), None) == -2   # n < 0
    assert call(None, 0, _p(t), 0, 16, _p(thr), 8, 0, 0, _p(state), _p(err), None) == -3    # NULL preds, n > 0
    assert call(_p(x), 0, _p(t), 0, 16, _p(thr), 8, 0, 0, None, _p(err), None) == -3        # NULL state
    torch.cuda.synchronize()
    assert int(state.sum()) == 0 and int(err.item()) == 0
