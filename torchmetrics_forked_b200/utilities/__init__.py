"""Utilities the hot path needs (a small subset of the reference's `torchmetrics.utilities`)."""
from torchmetrics_forked_b200.utilities.checks import _check_same_shape  # noqa: F401
from torchmetrics_forked_b200.utilities.data import apply_to_collection, dim_zero_cat  # noqa: F401
from torchmetrics_forked_b200.utilities.distributed import gather_all_tensors  # noqa: F401
from torchmetrics_forked_b200.utilities.prints import rank_zero_debug, rank_zero_info, rank_zero_warn  # noqa: F401
