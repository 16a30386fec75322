"""B200-native classification-metric update path behind the torchmetrics `Metric` API.

`Metric` / `MetricCollection` and the classification classes keep the reference's interface
(torchmetrics 1.3.0dev); their state updates run as hand-written sm_100a CUDA kernels reached through the
C ABI in `include/tmb200.h`.  There is no CPU fallback for those kernels.
"""
from torchmetrics_forked_b200.collections import MetricCollection
from torchmetrics_forked_b200.metric import Metric

__version__ = "0.1.0"
__all__ = ["Metric", "MetricCollection"]
