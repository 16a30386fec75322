"""K4 host logic on CPU: world_size-2 gloo process group, packed all-reduce for sum states, all-gather for
cat states, the sync / unsync protocol (reference tests/unittests/bases/test_ddp.py:33-274)."""
import os
import socket
import sys
from pathlib import Path

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = Path(__file__).resolve().parent.parent
WORLD = 2


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank: int, port: int, fn_name: str, result_dir: str) -> None:
    for p in (str(ROOT), str(ROOT / "oracle"), str(ROOT / "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=WORLD)
    try:
        globals()[fn_name](rank)
        Path(result_dir, f"ok{rank}").write_text("ok")
    finally:
        dist.destroy_process_group()


def _run(fn_name: str, tmp_path) -> None:
    port = _free_port()
    mp.spawn(_worker, args=(port, fn_name, str(tmp_path)), nprocs=WORLD, join=True)
    assert all((tmp_path / f"ok{r}").exists() for r in range(WORLD))


def _metric_classes():
    from torchmetrics_forked_b200 import Metric

    class SumAndCat(Metric):
        full_state_update = False

        def __init__(self, **kw):
            super().__init__(**kw)
            self.add_state("confmat", torch.zeros(3, 3, dtype=torch.long), dist_reduce_fx="sum")
            self.add_state("tp", torch.zeros(4, dtype=torch.long), dist_reduce_fx="sum")
            self.add_state("total", torch.tensor(0.0), dist_reduce_fx="sum")
            self.add_state("preds", [], dist_reduce_fx="cat")
            self.add_state("best", torch.tensor(-1e9), dist_reduce_fx="max")
            self.add_state("raw", torch.tensor(0.0), dist_reduce_fx=None)

        def update(self, confmat, tp, total, preds, best):
            self.confmat += confmat
            self.tp += tp
            self.total += total
            if preds is not None:
                self.preds.append(preds)
            self.best = torch.maximum(self.best, best)
            self.raw = self.raw + total

        def compute(self):
            return self.confmat.sum() + self.tp.sum()

    return SumAndCat


def _states_for(rank: int):
    g = torch.Generator().manual_seed(100 + rank)
    return dict(
        confmat=torch.randint(10, (3, 3), generator=g),
        tp=torch.randint(10, (4,), generator=g),
        total=torch.tensor(float(rank + 1)),
        preds=torch.rand(3 + 2 * rank, 5, generator=g),  # uneven row counts across ranks
        best=torch.tensor(float(rank * 7)),
    )


def check_packed_sync_matches_oracle(rank: int) -> None:
    import hotpath as O

    m = _metric_classes()()
    mine = _states_for(rank)
    m.update(**mine)
    every = [_states_for(r) for r in range(WORLD)]
    m.sync()
    assert m._is_synced
    assert np.array_equal(m.confmat.numpy(), O.sync_sum([s["confmat"].numpy() for s in every]))
    assert np.array_equal(m.tp.numpy(), O.sync_sum([s["tp"].numpy() for s in every]))
    assert float(m.total) == sum(float(s["total"]) for s in every)
    assert isinstance(m.preds, torch.Tensor)
    assert np.array_equal(m.preds.numpy(), O.sync_cat([s["preds"].numpy() for s in every]))
    assert float(m.best) == 7.0
    assert m.raw.shape == (WORLD,)
    m.unsync()
    assert torch.equal(m.confmat, mine["confmat"]) and len(m.preds) == 1 and torch.equal(m.preds[0], mine["preds"])
    # compute() syncs, reduces and restores the local state
    total = sum(int(s["confmat"].sum() + s["tp"].sum()) for s in every)
    assert int(m.compute()) == total
    assert torch.equal(m.confmat, mine["confmat"])


def check_matches_generic_gather(rank: int) -> None:
    """The packed path and the reference-style per-state gather give identical synced states."""
    from torchmetrics_forked_b200.utilities.distributed import gather_all_tensors

    cls = _metric_classes()
    a, b = cls(), cls(dist_sync_fn=gather_all_tensors)
    for m in (a, b):
        m.update(**_states_for(rank))
    a.sync()
    b.sync(dist_sync_fn=gather_all_tensors)
    for name in a._defaults:
        assert torch.equal(torch.as_tensor(getattr(a, name)), torch.as_tensor(getattr(b, name))), name


def check_empty_and_one_sided_lists(rank: int) -> None:
    cls = _metric_classes()
    m = cls()
    s = _states_for(rank)
    s["preds"] = None  # nobody appended: the list state stays an empty list (reference test_ddp.py:265-274)
    m.update(**s)
    m.sync()
    assert m.preds == []
    m.unsync()
    m2 = cls()
    s2 = _states_for(rank)
    if rank == 1:
        s2["preds"] = None  # only rank 0 has rows
    m2.update(**s2)
    m2.sync()
    assert torch.equal(m2.preds, _states_for(0)["preds"])


def check_sync_protocol(rank: int) -> None:
    from torchmetrics_forked_b200.utilities.exceptions import TorchMetricsUserError

    m = _metric_classes()()
    m.update(**_states_for(rank))
    m.sync()
    with pytest.raises(TorchMetricsUserError, match="The Metric has already been synced."):
        m.sync()
    with pytest.raises(TorchMetricsUserError, match="The Metric shouldn't be synced when performing"):
        m(**_states_for(rank))
    m.unsync()
    with pytest.raises(TorchMetricsUserError, match="The Metric has already been un-synced."):
        m.unsync()
    with m.sync_context():
        assert m._is_synced
        synced_total = float(m.total)
    assert not m._is_synced and synced_total == 3.0
    m.sync(should_sync=False)
    assert not m._is_synced
    off = _metric_classes()(sync_on_compute=False)
    off.update(**_states_for(rank))
    mine = _states_for(rank)
    assert int(off.compute()) == int(mine["confmat"].sum() + mine["tp"].sum())


def check_noncontiguous_and_multidim_gather(rank: int) -> None:
    from torchmetrics_forked_b200.utilities.distributed import gather_all_tensors

    t = torch.arange(12.0).reshape(3, 4).T  # non-contiguous (reference test_ddp.py:105-127)
    out = gather_all_tensors(t)
    assert len(out) == WORLD and all(torch.equal(o, t) for o in out)
    uneven = torch.ones(rank + 1, 2 - rank)
    out = gather_all_tensors(uneven)
    assert [tuple(o.shape) for o in out] == [(1, 2), (2, 1)]
    assert torch.equal(gather_all_tensors(torch.tensor(float(rank)))[1], torch.tensor(1.0))


def check_data_errors_raise_on_every_rank_after_the_collective(rank: int) -> None:
    """A kernel-recorded data error on ONE rank: its bits ride in the packed sync, `compute()` raises on BOTH ranks
    (nobody is left waiting in the all-reduce), the local state is restored, and `reset()` clears the word."""
    from torchmetrics_forked_b200 import Metric

    class SumOnly(Metric):
        def __init__(self):
            super().__init__()
            self.add_state("tp", torch.zeros(4, dtype=torch.long), dist_reduce_fx="sum")

        def update(self, tp):
            self.tp += tp

        def compute(self):
            return self.tp.sum()

    class CatOnly(Metric):
        def __init__(self):
            super().__init__()
            self.add_state("preds", [], dist_reduce_fx="cat")

        def update(self, preds):
            self.preds.append(preds)

        def compute(self):
            return torch.cat(self.preds).sum() if isinstance(self.preds, list) else self.preds.sum()

    for m, arg in ((SumOnly(), torch.arange(4)), (CatOnly(), torch.ones(2 + rank, 3)), (_metric_classes()(), None)):
        if arg is None:
            m.update(**_states_for(rank))
        else:
            m.update(arg)
        if rank == 1:  # what a kernel does when it meets a label outside [0, num_classes)
            m._err_word = torch.tensor([1], dtype=torch.int32)
        with pytest.raises(RuntimeError, match="Invalid values detected.*outside"):
            m.compute()
        assert not m._is_synced
        m.reset()
        if arg is None:
            m.update(**_states_for(rank))
        else:
            m.update(arg)
        m.compute()  # clean epoch after the reset: nothing left to report


@pytest.mark.parametrize(
    "fn",
    [
        "check_data_errors_raise_on_every_rank_after_the_collective",
        "check_packed_sync_matches_oracle",
        "check_matches_generic_gather",
        "check_empty_and_one_sided_lists",
        "check_sync_protocol",
        "check_noncontiguous_and_multidim_gather",
    ],
)
def test_gloo_world2(fn, tmp_path):
    _run(fn, tmp_path)
