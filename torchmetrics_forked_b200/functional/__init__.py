"""Functional API (classification hot path only)."""
from torchmetrics_forked_b200.functional.classification import *  # noqa: F401,F403
from torchmetrics_forked_b200.functional.classification import __all__  # noqa: F401
