"""Host logic of the drop-in boundary: Metric / MetricCollection behaviours pinned by the reference's own
tests (tests/unittests/bases/test_metric.py, test_collections.py), exercised with
user-defined metrics that only use torch ops (so they run on CPU)."""
import pickle
from copy import deepcopy

import pytest
import torch

from torchmetrics_forked_b200 import Metric, MetricCollection
from torchmetrics_forked_b200.utilities.exceptions import TorchMetricsUserError


class DummySum(Metric):
    full_state_update = False

    def __init__(self, **kw):
        super().__init__(**kw)
        self.add_state("x", torch.tensor(0.0), dist_reduce_fx="sum")

    def update(self, v):
        self.x += v

    def compute(self):
        return self.x


class DummyList(Metric):
    full_state_update = True

    def __init__(self, **kw):
        super().__init__(**kw)
        self.add_state("x", [], dist_reduce_fx="cat")

    def update(self, v):
        self.x.append(v)

    def compute(self):
        return torch.cat([t.reshape(-1) for t in self.x]).sum() if self.x else torch.tensor(0.0)


class MyAccuracy(Metric):  # the user-defined metric of the reference README (README.md:225-254)
    def __init__(self, **kw):
        super().__init__(**kw)
        self.add_state("correct", default=torch.tensor(0), dist_reduce_fx="sum")
        self.add_state("total", default=torch.tensor(0), dist_reduce_fx="sum")

    def update(self, preds, target):
        self.correct += torch.sum(preds.argmax(-1) == target)
        self.total += target.numel()

    def compute(self):
        return self.correct.float() / self.total


def test_constructor_kwarg_errors():
    for name, msg in [
        ("dist_sync_on_step", "Expected keyword argument `dist_sync_on_step` to be an `bool` but"),
        ("compute_on_cpu", "Expected keyword argument `compute_on_cpu` to be an `bool` but"),
        ("sync_on_compute", "Expected keyword argument `sync_on_compute` to be a `bool` but"),
        ("compute_with_cache", "Expected keyword argument `compute_with_cache` to be a `bool` but"),
        ("dist_sync_fn", "Expected keyword argument `dist_sync_fn` to be an callable function"),
    ]:
        with pytest.raises(ValueError, match=msg):
            DummySum(**{name: 42 if name != "dist_sync_fn" else [2, 3]})
    with pytest.raises(ValueError, match="Unexpected keyword arguments: `bar`, `foo`"):
        DummySum(foo=True, bar=42)


def test_add_state_rules():
    m = DummySum()
    m.add_state("a", torch.tensor(0), "sum")
    assert m._reductions["a"](torch.tensor([1, 1])) == 2
    m.add_state("b", torch.tensor(0), "mean")
    m.add_state("c", torch.tensor(0), "cat")
    m.add_state("d", [], None)
    with pytest.raises(ValueError, match="`dist_reduce_fx` must be callable or one of .*"):
        m.add_state("e", torch.tensor(0), "xyz")
    with pytest.raises(ValueError, match="state variable must be a tensor or any empty list .*"):
        m.add_state("f", [torch.tensor(0)], "sum")
    with pytest.raises(ValueError, match="state variable must be a tensor or any empty list .*"):
        m.add_state("g", 42, "sum")
    assert m._persistent["a"] is False
    assert set(m.metric_state) >= {"x", "a", "b", "c", "d"}


def test_update_compute_cache_reset():
    m = DummySum()
    with pytest.warns(UserWarning, match="The ``compute`` method of metric DummySum"):
        m.compute()
    m.update(1.0)
    assert m.update_called and m.update_count == 1
    assert m.compute() == 1 and m._computed == 1
    m.update(2.0)
    assert m._computed is None
    assert m.compute() == 3
    m.reset()
    assert m.x == 0 and m._computed is None and not m.update_called
    nocache = DummySum(compute_with_cache=False)
    nocache.update(2.0)
    nocache.compute()
    assert nocache._computed is None
    lst = DummyList()
    lst.update(torch.tensor([1.0]))
    lst.reset()
    assert lst.x == [] and lst._forward_cache is None


def test_forward_both_strategies():
    for cls in (DummySum, DummyList):
        m = cls()
        assert m(torch.tensor(5.0)) == 5 and m._forward_cache == 5
        assert m(torch.tensor(8.0)) == 8 and m._forward_cache == 8
        assert m.compute() == 13 and m.update_count == 2
    m = DummySum()
    m._is_synced = True
    with pytest.raises(TorchMetricsUserError, match="The Metric shouldn't be synced when performing"):
        m(torch.tensor(1.0))
    # all-sum tensor states (full_state_update=False) take the fused path: the global state tensor keeps its
    # identity (accumulated in place), a failing update leaves it untouched and restored
    m = DummySum()
    assert m._all_sum_tensor_states()
    before = m.x
    assert m(torch.tensor(2.0)) == 2 and m(torch.tensor(3.0)) == 3
    assert m.x is before and m.x == 5 and m.update_count == 2
    with pytest.raises(TypeError):
        m("not a number")
    assert m.x is before and m.x == 5 and m.update_count == 2


def test_constants_and_iteration_and_hash():
    m = DummySum()
    for const in ("is_differentiable", "higher_is_better", "full_state_update", "plot_lower_bound"):
        with pytest.raises(RuntimeError, match="Can't change const"):
            setattr(m, const, True)
    with pytest.raises(TypeError):
        for _ in m:
            pass
    assert hash(DummySum()) != hash(DummySum())
    a, b = DummyList(), DummyList()
    a.update(torch.tensor(1.0))
    b.update(torch.tensor(1.0))
    assert hash(a) != hash(b)


def test_pickle_clone_state_dict():
    m = MyAccuracy()
    m.update(torch.eye(3), torch.tensor([0, 1, 0]))
    m2 = pickle.loads(pickle.dumps(m))
    assert m2.compute() == m.compute()
    m2.update(torch.eye(3), torch.tensor([0, 1, 2]))
    assert m2.total == 6 and m.total == 3
    assert m.clone().compute() == m.compute()
    assert m.state_dict() == {}
    m.persistent(True)
    assert set(m.state_dict()) == {"correct", "total"}
    fresh = MyAccuracy()
    fresh.persistent(True)
    fresh.load_state_dict(m.state_dict())
    assert fresh.total == 3
    # as a child module: only persistent states are saved, and the child moves with the parent
    class Net(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.metric = DummySum()

    assert Net().state_dict() == {}


def test_dtype_casts_are_guarded():
    m = DummySum()
    for fn in ("half", "double", "float"):
        getattr(m, fn)()
        assert m.x.dtype == torch.float32
    m.type(torch.float16)
    assert m.x.dtype == torch.float32
    m.set_dtype(torch.float64)
    assert m.x.dtype == torch.float64
    m.reset()
    assert m.x.dtype == torch.float64


def test_sync_protocol_errors_without_a_process_group():
    m = DummySum()
    m.update(1.0)
    m.sync()  # not distributed: a no-op
    assert not m._is_synced
    with pytest.raises(TorchMetricsUserError, match="The Metric has already been un-synced."):
        m.unsync()
    m._is_synced = True
    with pytest.raises(TorchMetricsUserError, match="The Metric has already been synced."):
        m.sync()
    with pytest.raises(TorchMetricsUserError, match="The internal cache should exist to unsync the Metric."):
        m.unsync()


def test_custom_sync_fn_called_once_per_state():
    calls = []

    def fake_gather(t, group=None):
        calls.append(t)
        return [t, t]

    class Four(Metric):
        def __init__(self, **kw):
            super().__init__(**kw)
            for n in "abcd":
                self.add_state(n, torch.tensor(1.0), dist_reduce_fx="sum")

        def update(self):
            pass

        def compute(self):
            return self.a + self.b + self.c + self.d

    m = Four(dist_sync_fn=fake_gather, distributed_available_fn=lambda: True)
    m.update()
    assert m.compute() == 8 and len(calls) == 4  # reference test_metric.py:517-529
    assert m.a == 1  # unsynced again


def test_filter_kwargs_and_collection_basics():
    class TakesXY(DummySum):
        def update(self, x, y):
            self.x += x + y

    class TakesX(DummySum):
        def update(self, x):
            self.x += x

    coll = MetricCollection({"a": TakesXY(), "b": TakesX()}, prefix="p_", postfix="_s")
    coll.update(x=torch.tensor(1.0), y=torch.tensor(2.0))
    out = coll.compute()
    assert out == {"p_a_s": torch.tensor(3.0), "p_b_s": torch.tensor(1.0)}
    assert list(coll.keys()) == ["p_a_s", "p_b_s"] and list(coll.keys(keep_base=True)) == ["a", "b"]
    with pytest.raises(ValueError, match="Encountered two metrics both named DummySum"):
        MetricCollection([DummySum(), DummySum()])
    with pytest.raises(ValueError, match="Unknown input to MetricCollection."):
        MetricCollection(5)
    with pytest.raises(ValueError, match="Expected input `prefix` to be a string"):
        MetricCollection([DummySum()], prefix=3)
    nested = MetricCollection([MetricCollection([TakesX()], prefix="in_")], prefix="out_")
    nested.update(x=torch.tensor(1.0))
    assert list(nested.compute()) == ["out_in_TakesX"]


def test_compute_groups_discovery_and_copy_on_access():
    class A(DummySum):
        pass

    class B(DummySum):
        def compute(self):
            return self.x * 2

    class C(Metric):
        def __init__(self):
            super().__init__()
            self.add_state("y", torch.tensor(0.0), dist_reduce_fx="sum")

        def update(self, v):
            self.y += v

        def compute(self):
            return self.y

    coll = MetricCollection([A(), B(), C()])
    assert coll.compute_groups == {0: ["A"], 1: ["B"], 2: ["C"]}
    coll.update(torch.tensor(1.0))
    assert coll.compute_groups == {0: ["A", "B"], 1: ["C"]}
    coll.update(torch.tensor(2.0))  # only A is updated; B aliases its state
    out = coll.compute()
    assert out["A"] == 3 and out["B"] == 6 and out["C"] == 3
    assert coll["B"].x is not coll["A"].x  # item access hands out copies ...
    coll.update(torch.tensor(1.0))  # ... and the next update links them again
    assert coll.compute()["B"] == 8
    coll.reset()
    assert coll.compute_groups == {0: ["A", "B"], 1: ["C"]}
    user = MetricCollection([A(), B()], compute_groups=[["A", "B"]])
    user.update(torch.tensor(2.0))
    assert user.compute()["B"] == 4
    with pytest.raises(ValueError, match="Input Z in `compute_groups`"):
        MetricCollection([A()], compute_groups=[["Z"]])
    off = MetricCollection([A(), B()], compute_groups=False)
    off.update(torch.tensor(1.0))
    assert off.compute_groups == {}
    assert deepcopy(coll).compute_groups == coll.compute_groups


def test_forward_on_a_grouped_collection_counts_each_batch_once():
    """Members of a compute group alias one state tensor and `forward` accumulates in place: the batch must land
    once per metric, not once per member (update -> forward, reset -> forward, forward -> update)."""

    class A(DummySum):
        pass

    class B(DummySum):
        def compute(self):
            return self.x * 2

    x = torch.tensor(6.0)
    coll = MetricCollection([A(), B()])
    coll.update(x)
    assert coll.compute_groups == {0: ["A", "B"]}
    batch = coll(x)
    assert batch["A"] == 6 and batch["B"] == 12
    out = coll.compute()
    assert out["A"] == 12 and out["B"] == 24
    coll.update(x)  # links the members again; the head alone is updated
    out = coll.compute()
    assert out["A"] == 18 and out["B"] == 36
    assert coll["A"].update_count == coll["B"].update_count == 3
    coll.reset()
    coll(x)
    coll(x)
    out = coll.compute()
    assert out["A"] == 12 and out["B"] == 24
    alone_a, alone_b = A(), B()
    for m in (alone_a, alone_b):
        m(x), m(x)
    assert out["A"] == alone_a.compute() and out["B"] == alone_b.compute()


def test_task_wrappers_and_arg_validation():
    from torchmetrics_forked_b200.classification import (
        AUROC, Accuracy, BinaryAccuracy, BinaryAUROC, ConfusionMatrix, F1Score, JaccardIndex,
        MulticlassAccuracy, MulticlassConfusionMatrix, MulticlassF1Score, MulticlassJaccardIndex,
        MulticlassStatScores, MultilabelAUROC, MultilabelConfusionMatrix, StatScores,
    )  # fmt: skip

    assert isinstance(Accuracy(task="binary"), BinaryAccuracy)
    acc = Accuracy(task="multiclass", num_classes=5)
    assert isinstance(acc, MulticlassAccuracy) and acc.average == "micro" and acc.tp.shape == (1,)
    assert MulticlassAccuracy(5).average == "macro" and MulticlassAccuracy(5).tp.shape == (5,)
    assert isinstance(ConfusionMatrix(task="multiclass", num_classes=3), MulticlassConfusionMatrix)
    assert isinstance(ConfusionMatrix(task="multilabel", num_labels=3), MultilabelConfusionMatrix)
    assert isinstance(F1Score(task="multiclass", num_classes=3), MulticlassF1Score)
    assert isinstance(JaccardIndex(task="multiclass", num_classes=3), MulticlassJaccardIndex)
    assert isinstance(StatScores(task="multiclass", num_classes=3), MulticlassStatScores)
    assert isinstance(AUROC(task="binary", thresholds=5), BinaryAUROC)
    assert isinstance(AUROC(task="multilabel", num_labels=4), MultilabelAUROC)
    with pytest.raises(ValueError, match=r"Invalid *"):
        ConfusionMatrix(task="NotValidTask")
    with pytest.raises(ValueError, match="`num_classes` is expected to be `int`"):
        ConfusionMatrix(task="multiclass")
    with pytest.raises(ValueError, match="Expected argument `num_classes` to be an integer larger than 1"):
        MulticlassConfusionMatrix(1)
    with pytest.raises(ValueError, match="Expected argument `normalize`"):
        MulticlassConfusionMatrix(3, normalize="bad")
    from torchmetrics_forked_b200.classification import BinaryPrecisionRecallCurve

    with pytest.raises(ValueError, match="Expected argument `thresholds`"):
        BinaryPrecisionRecallCurve(thresholds="x")
    with pytest.raises(ValueError, match="expected it to be larger than 1"):
        BinaryAUROC(thresholds=1)
    m = BinaryAUROC(thresholds=200)
    assert torch.equal(m.thresholds, torch.linspace(0, 1, 200)) and m.confmat.shape == (200, 2, 2)
    assert m.state_dict() == {}  # thresholds is a non-persistent buffer (reference test :441-451)
    assert BinaryAUROC().preds == [] and MultilabelAUROC(3, thresholds=7).confmat.shape == (7, 3, 2, 2)
    assert MulticlassConfusionMatrix(4).confmat.dtype == torch.int64
    assert MulticlassConfusionMatrix.full_state_update is False and BinaryAUROC.full_state_update is False


def test_jit_script_does_not_fail():
    torch.jit.script(DummySum())
