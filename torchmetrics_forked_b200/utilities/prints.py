"""Rank-aware warning/info helpers (reference: utilities/prints.py:22-55).

Only local rank 0 speaks, so an 8-GPU job prints every warning once.
"""
import logging
import os
import warnings
from functools import wraps
from typing import Any, Callable

log = logging.getLogger("torchmetrics_forked_b200")


def rank_zero_only(fn: Callable) -> Callable:
    """Decorator: run ``fn`` on local rank 0 only, return ``None`` elsewhere."""

    @wraps(fn)
    def guarded(*args: Any, **kwargs: Any) -> Any:
        return fn(*args, **kwargs) if rank_zero_only.rank == 0 else None

    return guarded


rank_zero_only.rank = getattr(rank_zero_only, "rank", int(os.environ.get("LOCAL_RANK", 0)))


@rank_zero_only
def rank_zero_warn(*args: Any, **kwargs: Any) -> None:
    warnings.warn(*args, **kwargs)  # noqa: B028


@rank_zero_only
def rank_zero_info(*args: Any, **kwargs: Any) -> None:
    log.info(*args, **kwargs)


@rank_zero_only
def rank_zero_debug(*args: Any, **kwargs: Any) -> None:
    log.debug(*args, **kwargs)
