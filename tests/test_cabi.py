"""The C-ABI library loads on a CPU-only machine and exports exactly what include/tmb200.h declares.
No kernel is launched here (pure host checks only)."""
import ctypes
import re
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent


@pytest.fixture(scope="module")
def lib():
    from torchmetrics_forked_b200 import _build, _native

    _build.build_library()
    return _native.lib()


def _declared_symbols():
    text = (ROOT / "include" / "tmb200.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(tmb_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_are_exported(lib):
    from torchmetrics_forked_b200 import _native

    declared = _declared_symbols()
    assert declared, "no declarations found in include/tmb200.h"
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in tmb200.h but not exported by libtmb200.so"
    assert sorted(_native.SIGNATURES) == declared, "ctypes signatures and header disagree"


def test_built_for_sm100a(lib):
    import subprocess

    from torchmetrics_forked_b200 import _native

    assert lib.tmb_build_arch() == 100
    out = subprocess.run(["cuobjdump", "-lelf", _native.library_path()], capture_output=True, text=True)
    if out.returncode == 0:
        assert "sm_100a" in out.stdout and "sm_90" not in out.stdout


def test_host_only_queries(lib):
    assert lib.tmb_mc_scratch_bytes(1000, 0) == 0
    assert lib.tmb_mc_scratch_bytes(1000, 1) == 3 * 1000 * 8
    assert lib.tmb_mc_scratch_bytes(1000, 2) == 16
    # header + two histogram variants + a padded float copy of the thresholds
    assert lib.tmb_thresh_bin_scratch_elems(1, 200) == 8 + 2 * 2 * 201 + 101
    assert lib.tmb_thresh_bin_scratch_elems(100, 200) == 8 + 2 * 100 * 2 * 201 + 101
    assert lib.tmb_thresh_bin_scratch_elems(0, 5) == 0


def test_argument_errors_need_no_gpu(lib):
    # dtype / mode / shape errors are rejected before anything touches the device
    assert lib.tmb_mc_argmax_hist(None, 9, None, 0, 1, 4, 1, 0, 0, 0, None, None, None) == -1
    assert lib.tmb_mc_argmax_hist(None, 0, None, 0, -1, 4, 1, 0, 0, 0, None, None, None) == -2
    assert lib.tmb_mc_label_hist(None, 9, None, 0, 1, 4, 0, 0, 0, None, None, None) == -1
    assert lib.tmb_thresh_bin_update(None, 0, 0, None, 9, 1, 1, 1, None, 1, 0, 0, 0, 0, None, 0, None, None) == -1
    assert lib.tmb_thresh_bin_finalize(None, 0, 1, 1, None, 0, None, None, None, None, None) == -3
    assert lib.tmb_bincount(None, 9, 1, 1, None, None, None) == -1
    # rows f2 / f3 / f4: documented limits and dtype checks come first too
    E_DTYPE, E_SHAPE, E_NULL, E_LIMIT = -1, -2, -3, -6
    assert lib.tmb_softmax_rows(None, 0, 4, 1025, None, None, 1, None) == E_LIMIT
    assert lib.tmb_softmax_rows(None, 9, 4, 10, None, None, 1, None) == E_DTYPE
    assert lib.tmb_softmax_rows(None, 0, 4, 10, None, None, 0, None) == E_SHAPE  # generation counters start at 1
    assert lib.tmb_rows_outside_flag(None, 0, 4, 10, None, 0, 0, 0, None, 1, None) == E_NULL
    assert lib.tmb_mc_curve_update(None, 0, None, 0, 4, 1025, None, 5, 0, 0, 0, None, 0, None, None) == E_LIMIT
    assert lib.tmb_mc_curve_update(None, 0, None, 0, 4, 1, None, 5, 0, 0, 0, None, 0, None, None) == E_SHAPE
    assert lib.tmb_mc_topk_hist(None, 0, None, 0, 4, 1025, 5, 0, 0, None, None, None) == E_NULL  # wide rows are built
    assert lib.tmb_mc_topk_hist(None, 0, None, 0, 4, (1 << 24) + 1, 5, 0, 0, None, None, None) == E_LIMIT
    assert lib.tmb_mc_topk_hist_samplewise(None, 0, None, 0, 4, 0, 10, 2, 0, 0, None, None, None) == E_SHAPE
    assert lib.tmb_softmax_normalize(None, 0, 4, 1025, None, 0, 0, 0, None, None, 1, None) == E_LIMIT
    assert lib.tmb_clf_pack_keys_sharded(None, 1, 0, 0, 4, 3, 0, 0, 1, 0, None, None, 2, 1, 8, 0, None, None) == E_SHAPE
    assert lib.tmb_clf_sorted_scores(None, None, -1, 3, 1, 0, None, 0, None, None) == E_SHAPE
    assert lib.tmb_mc_topk_hist(None, 0, None, 0, 4, 10, 11, 0, 0, None, None, None) == E_SHAPE  # top_k > classes
    assert lib.tmb_mc_hist_samplewise(None, 9, 0, None, 0, 4, 10, 3, 0, 0, None, None, None) == E_DTYPE
    assert lib.tmb_mc_shared_confmat(None, 10, None, None) == E_NULL
    assert lib.tmb_mc_shared_stats(None, 10, 0, 1, None, None, None, None, None) == E_NULL
    assert lib.tmb_mc_scratch_bytes(10, 3) == (100 + 30) * 8  # TMB_MC_BOTH: batch confusion matrix + 3 histograms
    assert lib.tmb_clf_curves_chunked(None, 1, 0, 0, 5, 2, 0, 0, 1, 0, None, 0, None, None, None, None, None, None) == E_NULL


def test_no_cpu_fallback():
    import torch

    import torchmetrics_forked_b200.functional as F
    from torchmetrics_forked_b200.classification import BinaryAUROC, MulticlassConfusionMatrix

    with pytest.raises(RuntimeError, match="no CPU"):
        F.multiclass_confusion_matrix(torch.randn(8, 3), torch.randint(3, (8,)), 3)
    with pytest.raises(RuntimeError, match="no CPU"):
        MulticlassConfusionMatrix(3).update(torch.randn(8, 3), torch.randint(3, (8,)))
    with pytest.raises(RuntimeError, match="no CPU"):
        BinaryAUROC(thresholds=5).update(torch.rand(8), torch.randint(2, (8,)))


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    from torchmetrics_forked_b200 import _native

    monkeypatch.setattr(_native, "_lib", None)
    monkeypatch.setattr(_native, "LIB_PATH", tmp_path / "libtmb200.so")
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        _native.lib()
