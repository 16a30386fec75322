"""K4 helpers on one GPU: `tmb_pack_states` / `tmb_unpack_states` (the pack of the one-all-reduce state sync) against
`torch.cat`, through the C ABI and through `utilities.distributed._pack`.  The collectives themselves are covered by
tests/test_gpu_nccl.py (two GPUs), tests/test_ddp_gloo.py (CPU) and bench.py's N > 1 parity prologue."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dtype", [torch.int64, torch.float32, torch.int32])
def test_pack_and_unpack_round_trip(dtype):
    from torchmetrics_forked_b200 import _native as nat

    lib = nat.lib()
    g = torch.Generator().manual_seed(5)
    shapes = [(1000, 1000), (7,), (), (3, 5, 2), (1,)]
    states = [(torch.randint(-1000, 1000, s, generator=g).to(dtype)).cuda() for s in shapes]
    es = states[0].element_size()
    offsets = [0]
    for s in states:
        offsets.append(offsets[-1] + s.numel() * es)
    table = torch.tensor(offsets + [s.data_ptr() for s in states], dtype=torch.int64, device="cuda")
    flat = torch.full((offsets[-1] // es + 16,), 77, dtype=dtype, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    assert lib.tmb_pack_states(table.data_ptr(), len(states), offsets[-1], flat.data_ptr(), st) == 0
    want = torch.cat([s.reshape(-1) for s in states])
    assert torch.equal(flat[: want.numel()], want)
    assert bool((flat[want.numel() :] == 77).all())  # nothing written past the packed bytes
    doubled = flat[: want.numel()] * 2
    assert lib.tmb_unpack_states(table.data_ptr(), len(states), offsets[-1], doubled.data_ptr(), st) == 0
    torch.cuda.synchronize()
    at = 0
    for s, shape in zip(states, shapes):
        assert torch.equal(s.reshape(-1), want[at : at + s.numel()] * 2)
        at += s.numel()


def test_python_pack_matches_cat_and_caches_its_table():
    from torchmetrics_forked_b200.utilities import distributed as D

    states = [torch.arange(12, device="cuda").reshape(3, 4), torch.arange(5, device="cuda") + 100, torch.tensor(9, device="cuda")]
    a = D._pack(states)
    states[1] += 1  # same storage, new values: the cached address table must still be right
    b = D._pack(states)
    assert torch.equal(a, torch.cat([torch.arange(12), torch.arange(5) + 100, torch.tensor([9])]).cuda())
    assert torch.equal(b, torch.cat([torch.arange(12), torch.arange(5) + 101, torch.tensor([9])]).cuda())
    single = D._pack([states[0]])
    assert single.data_ptr() != states[0].data_ptr() and torch.equal(single, states[0].reshape(-1))
