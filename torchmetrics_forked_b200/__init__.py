"""B200-native classification-metric update path behind the torchmetrics `Metric` API.

`Metric` / `MetricCollection` and the classification classes keep the reference's interface
(torchmetrics 1.3.0dev); their state updates run as hand-written sm_100a CUDA kernels reached through the
C ABI in `include/tmb200.h`.  There is no CPU fallback for those kernels.
"""
from torchmetrics_forked_b200 import functional
from torchmetrics_forked_b200.classification import (
    AUROC,
    ROC,
    Accuracy,
    AveragePrecision,
    CohenKappa,
    ConfusionMatrix,
    F1Score,
    FBetaScore,
    HammingDistance,
    JaccardIndex,
    MatthewsCorrCoef,
    Precision,
    PrecisionAtFixedRecall,
    PrecisionRecallCurve,
    Recall,
    RecallAtFixedPrecision,
    Specificity,
    SpecificityAtSensitivity,
    StatScores,
)
from torchmetrics_forked_b200.collections import MetricCollection
from torchmetrics_forked_b200.metric import Metric

__version__ = "0.1.0"
# the task wrappers of the metrics on the path, under the names the reference exports them (torchmetrics/__init__.py:41-65)
__all__ = [
    "functional", "Metric", "MetricCollection", "AUROC", "ROC", "Accuracy", "AveragePrecision", "CohenKappa",
    "ConfusionMatrix", "F1Score", "FBetaScore", "HammingDistance", "JaccardIndex", "MatthewsCorrCoef", "Precision",
    "PrecisionAtFixedRecall", "PrecisionRecallCurve", "Recall", "RecallAtFixedPrecision", "Specificity",
    "SpecificityAtSensitivity", "StatScores",
]  # fmt: skip
