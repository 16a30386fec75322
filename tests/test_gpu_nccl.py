"""K4 on real hardware: two ranks, one GPU each, NCCL.  Samples are sharded by rank; `compute()` must return
the metric of ALL samples (packed all-reduce for sum states, all-gather for cat states) and leave the local
state untouched.  Skipped on a single-GPU box; the host logic is covered on CPU by tests/test_ddp_gloo.py."""
import os
import socket
import sys
from pathlib import Path

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = Path(__file__).resolve().parent.parent
WORLD = 2


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _shard(rank: int):
    g = torch.Generator().manual_seed(500 + rank)
    n = 40_000 + 1000 * rank  # uneven shards
    return dict(
        mc_preds=torch.randn(n, 50, generator=g),
        mc_target=torch.randint(50, (n,), generator=g),
        bin_preds=torch.rand(n, generator=g),
        bin_target=torch.randint(2, (n,), generator=g),
        ml_preds=torch.rand(n // 10, 4, generator=g),
        ml_target=torch.randint(2, (n // 10, 4), generator=g),
    )


def _worker(rank: int, port: int, out_dir: str) -> None:
    for p in (str(ROOT), str(ROOT / "oracle"), str(ROOT / "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import torch.distributed as dist

    import hotpath as O
    from torchmetrics_forked_b200 import MetricCollection
    from torchmetrics_forked_b200.classification import (
        BinaryPrecisionRecallCurve,
        MulticlassAccuracy,
        MulticlassAUROC,
        MulticlassConfusionMatrix,
        MultilabelAUROC,
    )

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=WORLD, device_id=dev)
    try:
        mine = {k: v.to(dev) for k, v in _shard(rank).items()}
        every = [_shard(r) for r in range(WORLD)]
        cat = lambda key: torch.cat([s[key] for s in every]).numpy()  # noqa: E731

        coll = MetricCollection([MulticlassAccuracy(50, average="micro"), MulticlassConfusionMatrix(50)]).to(dev)
        coll.update(mine["mc_preds"], mine["mc_target"])
        local_confmat = coll["MulticlassConfusionMatrix"].confmat.clone()
        out = coll.compute()
        want = O.multiclass_confusion_matrix(cat("mc_preds"), cat("mc_target"), 50)
        assert np.array_equal(out["MulticlassConfusionMatrix"].cpu().numpy(), want)
        tp, fp, tn, fn = O.multiclass_stat_scores(cat("mc_preds"), cat("mc_target"), 50, "micro")
        assert abs(float(out["MulticlassAccuracy"]) - float(O.accuracy_reduce(tp, fp, tn, fn, "micro"))) < 1e-6
        # compute() un-syncs: the local shard's state is back
        assert torch.equal(coll["MulticlassConfusionMatrix"].confmat, local_confmat)

        curve = BinaryPrecisionRecallCurve(thresholds=64).to(dev)
        curve.update(mine["bin_preds"], mine["bin_target"])
        curve.sync()
        want = O.binary_curve_binned(cat("bin_preds"), cat("bin_target"), torch.linspace(0, 1, 64).numpy())
        assert np.array_equal(curve.confmat.cpu().numpy(), want)
        curve.unsync()

        auroc = MultilabelAUROC(4, thresholds=None, average="none").to(dev)
        auroc.update(mine["ml_preds"], mine["ml_target"])
        got = auroc.compute().cpu().numpy()
        curves = O.multilabel_clf_curves(cat("ml_preds"), cat("ml_target"))
        np.testing.assert_allclose(got, [O.binary_auroc_exact(f, t) for f, t, _ in curves], rtol=1e-5)
        with auroc.sync_context():  # rank-ordered concatenation of the uneven shards
            assert np.array_equal(auroc.preds.cpu().numpy(), cat("ml_preds"))
        # rows f2 / f3 through the same sync: exact multiclass AUROC over two updates per rank (a list state of
        # [n, C] chunks and one of [n] class indices), and top-k accuracy (per-class sum states)
        probs = lambda s: s["mc_preds"][:4000].softmax(1)  # noqa: E731  (in [0,1]: stored as they are)
        mca = MulticlassAUROC(50, thresholds=None, average="none").to(dev)
        mine_probs = probs(_shard(rank)).to(dev)
        mca.update(mine_probs[:1500], mine["mc_target"][:1500])
        mca.update(mine_probs[1500:], mine["mc_target"][1500:4000])
        got = mca.compute().cpu().numpy()
        all_probs = torch.cat([probs(s) for s in every]).numpy()
        all_t = torch.cat([s["mc_target"][:4000] for s in every]).numpy()
        want = [O.binary_auroc_exact(f, t) for f, t, _ in O.multiclass_clf_curves(all_probs, all_t)]
        np.testing.assert_allclose(got, want, rtol=1e-5)
        top3 = MulticlassAccuracy(50, top_k=3, average="micro").to(dev)
        top3.update(mine["mc_preds"], mine["mc_target"])
        tp, fp, tn, fn = O.multiclass_stat_scores_topk(cat("mc_preds"), cat("mc_target"), 50, 3)
        assert abs(float(top3.compute()) - tp.sum() / (tp.sum() + fn.sum())) < 1e-6
        # ---- label-sharded exact compute (_sharded.py): every variant against the same metric computed on ONE rank
        # from all rows (no process group involved: sync_on_compute=False), and across the exchange flavours
        import os as _os

        from torchmetrics_forked_b200.classification import (
            BinaryAUROC,
            BinaryAveragePrecision,
            MultilabelAveragePrecision,
        )

        def ml(r):
            g = torch.Generator().manual_seed(900 + r)
            n = 5000 + 777 * r
            p = torch.randn(n, 5, generator=g)
            p[:, 2] = (p[:, 2] * 4).round() / 4  # ties
            t = torch.randint(2, (n, 5), generator=g)
            t[torch.rand(n, 5, generator=g) < 0.1] = -1
            return p, t

        def variants():
            yield "ml_auroc", lambda **k: MultilabelAUROC(5, thresholds=None, average="none", ignore_index=-1, **k), ml
            yield "ml_ap", lambda **k: MultilabelAveragePrecision(5, thresholds=None, average="macro", ignore_index=-1, **k), ml
            yield "bin_auroc", lambda **k: BinaryAUROC(thresholds=None, **k), lambda r: (ml(r)[0][:, 0], (ml(r)[1][:, 0] > 0).long())
            yield "bin_ap", lambda **k: BinaryAveragePrecision(thresholds=None, **k), lambda r: (ml(r)[0][:, 1], (ml(r)[1][:, 1] > 0).long())
            yield "mc_auroc", lambda **k: MulticlassAUROC(50, thresholds=None, average="macro", **k), lambda r: (
                probs(_shard(r)), _shard(r)["mc_target"][:4000])

        for flavour in ("symm", "nccl", "gather"):
            _os.environ["TMB_EXACT_EXCHANGE"] = "nccl" if flavour == "nccl" else "symm"
            _os.environ["TMB_EXACT_SHARDED"] = "0" if flavour == "gather" else "1"
            for name, make, data in variants():
                m, alone = make().to(dev), make(sync_on_compute=False).to(dev)
                p, t = data(rank)
                half = p.shape[0] // 2
                m.update(p[:half].to(dev), t[:half].to(dev))      # two chunks per rank
                m.update(p[half:].to(dev), t[half:].to(dev))
                for r in range(WORLD):
                    pr, tr = data(r)
                    alone.update(pr.to(dev), tr.to(dev))
                got, want = m.compute(), alone.compute()
                torch.testing.assert_close(got, want, rtol=1e-6, atol=1e-7), (flavour, name)
                assert isinstance(m.preds, list) and len(m.preds) == 2, "compute() must leave the local state"
            # a rank without any rows takes part in the exchange
            m = MultilabelAUROC(5, thresholds=None, average="none", ignore_index=-1).to(dev)
            alone = MultilabelAUROC(5, thresholds=None, average="none", ignore_index=-1, sync_on_compute=False).to(dev)
            if rank == 0:
                m.update(*[x.to(dev) for x in ml(0)])
            alone.update(*[x.to(dev) for x in ml(0)])
            import warnings

            with warnings.catch_warnings():
                warnings.simplefilter("ignore")
                torch.testing.assert_close(m.compute(), alone.compute(), rtol=1e-6, atol=1e-7)
        _os.environ.pop("TMB_EXACT_EXCHANGE", None)
        _os.environ.pop("TMB_EXACT_SHARDED", None)
        Path(out_dir, f"ok{rank}").write_text("ok")
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < WORLD, reason="needs two GPUs")
def test_nccl_sync_two_ranks(tmp_path):
    import torch.multiprocessing as mp

    mp.spawn(_worker, args=(_free_port(), str(tmp_path)), nprocs=WORLD, join=True)
    assert all((tmp_path / f"ok{r}").exists() for r in range(WORLD))
