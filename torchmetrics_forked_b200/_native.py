"""ctypes binding of libtmb200.so (the C ABI declared in include/tmb200.h).

This is the only place where Python meets the CUDA kernels.  There is deliberately no fallback: if the
shared object is missing or a tensor is not on a CUDA device, the call raises.  PyTorch is used for
device memory (`tensor.data_ptr()`), the current stream and nothing else.
"""
import ctypes
from typing import Optional

import torch
from torch import Tensor

from torchmetrics_forked_b200._build import LIB_PATH

c_void_p, c_int, c_int64 = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64

# dtype codes of include/tmb200.h
FLOAT_CODES = {torch.float32: 0, torch.float16: 1, torch.bfloat16: 2}
LABEL_CODES = {torch.int64: 0, torch.int32: 1, torch.int16: 2, torch.uint8: 3, torch.bool: 3, torch.int8: 4}

MC_CONFMAT, MC_STATS, MC_MICRO, MC_BOTH = 0, 1, 2, 3
ERR_TARGET_RANGE, ERR_PREDS_RANGE, ERR_TARGET_BINARY, ERR_SCORE_RANGE = 1, 2, 4, 8

_ARG_ERRORS = {
    -1: "unsupported dtype",
    -2: "invalid shape or size",
    -3: "null pointer",
    -4: "invalid mode",
    -5: "scratch buffer too small",
    -6: "size exceeds the documented limit of this entry point",
}

# name -> (restype, argtypes); must list every symbol the header declares (tests/test_cabi.py checks)
SIGNATURES = {
    "tmb_build_arch": (c_int, []),
    "tmb_mc_argmax_hist": (
        c_int,
        [c_void_p, c_int, c_void_p, c_int, c_int64, c_int64, c_int64, c_int, c_int64, c_int, c_void_p, c_void_p,
         c_void_p],
    ),
    "tmb_mc_hist_samplewise": (
        c_int,
        [c_void_p, c_int, c_int, c_void_p, c_int, c_int64, c_int64, c_int64, c_int, c_int64, c_void_p, c_void_p,
         c_void_p],
    ),
    "tmb_mc_stats_finalize_samplewise": (
        c_int, [c_void_p, c_int64, c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]
    ),
    "tmb_mc_shared_confmat": (c_int, [c_void_p, c_int64, c_void_p, c_void_p]),
    "tmb_mc_shared_stats": (
        c_int, [c_void_p, c_int64, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]
    ),
    "tmb_mc_topk_hist": (
        c_int,
        [c_void_p, c_int, c_void_p, c_int, c_int64, c_int64, c_int64, c_int, c_int64, c_void_p, c_void_p, c_void_p],
    ),
    "tmb_mc_topk_hist_samplewise": (
        c_int,
        [c_void_p, c_int, c_void_p, c_int, c_int64, c_int64, c_int64, c_int64, c_int, c_int64, c_void_p, c_void_p,
         c_void_p],
    ),
    "tmb_mc_label_hist": (
        c_int,
        [c_void_p, c_int, c_void_p, c_int, c_int64, c_int64, c_int, c_int64, c_int, c_void_p, c_void_p, c_void_p],
    ),
    "tmb_mc_stats_finalize": (c_int, [c_void_p, c_int64, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "tmb_mc_scratch_bytes": (c_int64, [c_int64, c_int]),
    "tmb_thresh_bin_scratch_elems": (c_int64, [c_int64, c_int64]),
    "tmb_thresh_bin_update": (
        c_int,
        [c_void_p, c_int, c_int, c_void_p, c_int, c_int64, c_int64, c_int64, c_void_p, c_int64, c_int, c_int,
         c_int64, c_int, c_void_p, c_int64, c_void_p, c_void_p],
    ),
    "tmb_thresh_bin_finalize": (
        c_int,
        [c_void_p, c_int64, c_int64, c_int64, c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p],
    ),
    "tmb_binary_curve_update_noalloc": (
        c_int, [c_void_p, c_int, c_void_p, c_int, c_int64, c_void_p, c_int64, c_int, c_int64, c_void_p, c_void_p, c_void_p]
    ),
    "tmb_sigmoid_normalize": (
        c_int, [c_void_p, c_int, c_int64, c_void_p, c_int, c_int, c_int64, c_void_p, c_void_p, c_int64, c_void_p]
    ),
    "tmb_clf_curve_scratch_bytes": (c_int64, [c_int64, c_int64]),
    "tmb_rows_outside_flag": (
        c_int,
        [c_void_p, c_int, c_int64, c_int64, c_void_p, c_int, c_int, c_int64, c_void_p, c_int64, c_void_p],
    ),
    "tmb_softmax_normalize": (
        c_int,
        [c_void_p, c_int, c_int64, c_int64, c_void_p, c_int, c_int, c_int64, c_void_p, c_void_p, c_int64, c_void_p],
    ),
    "tmb_softmax_rows": (c_int, [c_void_p, c_int, c_int64, c_int64, c_void_p, c_void_p, c_int64, c_void_p]),
    "tmb_mc_curve_update": (
        c_int,
        [c_void_p, c_int, c_void_p, c_int, c_int64, c_int64, c_void_p, c_int64, c_int, c_int64, c_int, c_void_p,
         c_int64, c_void_p, c_void_p],
    ),
    "tmb_clf_curves": (
        c_int,
        [c_void_p, c_int, c_void_p, c_int, c_int64, c_int64, c_int, c_int64, c_int64, c_void_p, c_int64, c_void_p,
         c_void_p, c_void_p, c_void_p, c_void_p, c_void_p],
    ),
    "tmb_clf_curves_chunked": (
        c_int,
        [c_void_p, c_int64, c_int, c_int, c_int64, c_int64, c_int, c_int64, c_int64, c_int, c_void_p, c_int64,
         c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p],
    ),
    "tmb_clf_curve_scores_chunked": (
        c_int,
        [c_void_p, c_int64, c_int, c_int, c_int64, c_int64, c_int, c_int64, c_int64, c_int, c_int, c_void_p,
         c_int64, c_void_p, c_void_p, c_void_p],
    ),
    "tmb_clf_pack_keys_sharded": (
        c_int,
        [c_void_p, c_int64, c_int, c_int, c_int64, c_int64, c_int, c_int64, c_int64, c_int, c_void_p, c_void_p, c_int64,
         c_int64, c_int64, c_int64, c_void_p, c_void_p],
    ),
    "tmb_clf_sorted_scores_scratch_bytes": (c_int64, [c_int64, c_int64]),
    "tmb_clf_sorted_scores": (
        c_int, [c_void_p, c_void_p, c_int64, c_int64, c_int, c_int, c_void_p, c_int64, c_void_p, c_void_p]
    ),
    "tmb_curve_reduce": (c_int, [c_void_p, c_void_p, c_int64, c_int64, c_void_p, c_void_p, c_void_p]),
    "tmb_compact_scratch_bytes": (c_int64, [c_int64]),
    "tmb_compact_not_ignored": (
        c_int,
        [c_void_p, c_int, c_void_p, c_int, c_int64, c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_int64,
         c_void_p],
    ),
    "tmb_bincount": (c_int, [c_void_p, c_int, c_int64, c_int64, c_void_p, c_void_p, c_void_p]),
    "tmb_label_max": (c_int, [c_void_p, c_int, c_int64, c_void_p, c_void_p]),
    "tmb_pack_states": (c_int, [c_void_p, c_int64, c_int64, c_void_p, c_void_p]),
    "tmb_unpack_states": (c_int, [c_void_p, c_int64, c_int64, c_void_p, c_void_p]),
    "tmb_sum_sequential_f32": (c_int, [c_void_p, c_int64, c_void_p, c_void_p]),
}

_lib: Optional[ctypes.CDLL] = None


def library_path() -> str:
    return str(LIB_PATH)


def lib() -> ctypes.CDLL:
    """Load libtmb200.so once; raise (never fall back) when it has not been built."""
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            raise RuntimeError(
                f"{LIB_PATH} is missing: the CUDA kernels have not been built and there is no CPU fallback. "
                "Run `python -c 'import __graft_entry__ as g; g.build()'` (needs nvcc) from the repo root."
            )
        handle = ctypes.CDLL(str(LIB_PATH))
        for name, (restype, argtypes) in SIGNATURES.items():
            fn = getattr(handle, name)
            fn.restype = restype
            fn.argtypes = argtypes
        _lib = handle
    return _lib


def check(rc: int, what: str) -> None:
    """Translate a C-ABI return code into a Python exception."""
    if rc == 0:
        return
    if rc < 0:
        raise ValueError(f"{what}: {_ARG_ERRORS.get(rc, 'argument error')} (code {rc})")
    raise RuntimeError(f"{what}: CUDA error {rc} while launching the kernel")


def require_cuda(t: Tensor, name: str) -> None:
    if not isinstance(t, Tensor):
        raise TypeError(f"`{name}` must be a torch.Tensor, got {type(t)}")
    if not t.is_cuda:
        raise RuntimeError(
            f"`{name}` lives on {t.device}: this package only implements the sm_100a CUDA path and has no CPU "
            "fallback. Move the metric and its inputs to a CUDA device."
        )


def float_code(t: Tensor, name: str = "preds") -> int:
    try:
        return FLOAT_CODES[t.dtype]
    except KeyError:
        raise ValueError(f"`{name}` has dtype {t.dtype}; supported floating dtypes: float32, float16, bfloat16") from None


def label_code(t: Tensor, name: str = "target") -> int:
    try:
        return LABEL_CODES[t.dtype]
    except KeyError:
        raise ValueError(
            f"`{name}` has dtype {t.dtype}; supported integer dtypes: int64, int32, int16, int8, uint8, bool"
        ) from None


_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)


def stream_ptr(device: torch.device) -> int:
    """cudaStream_t of torch's current stream on `device`, as an integer."""
    if _raw_stream is not None and device.index is not None:
        return _raw_stream(device.index)  # one C call; the Stream object route costs ~3 us per launch
    return torch.cuda.current_stream(device).cuda_stream


class _NoGuard:
    def __enter__(self) -> None:
        return None

    def __exit__(self, *exc: object) -> bool:
        return False


_NO_GUARD = _NoGuard()


def on_device(device: torch.device):
    """Context that makes `device` current for the launches inside it.  Small updates are launch / Python bound
    (tens of microseconds), and `torch.cuda.device(...)` alone costs several of them: when the device is current
    already — the usual case — nothing is switched."""
    if device.index is None or torch.cuda.current_device() == device.index:
        return _NO_GUARD
    return torch.cuda.device(device)


def ptr(t: Optional[Tensor]) -> int:
    return 0 if t is None else t.data_ptr()
