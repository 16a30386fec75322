"""Metric classes of the classification hot path (same names and constructor signatures as the reference)."""
from torchmetrics_forked_b200.classification.accuracy import (
    Accuracy,
    BinaryAccuracy,
    MulticlassAccuracy,
    MultilabelAccuracy,
)
from torchmetrics_forked_b200.classification.auroc import AUROC, BinaryAUROC, MulticlassAUROC, MultilabelAUROC
from torchmetrics_forked_b200.classification.average_precision import (
    AveragePrecision,
    BinaryAveragePrecision,
    MulticlassAveragePrecision,
    MultilabelAveragePrecision,
)
from torchmetrics_forked_b200.classification.confusion_matrix import (
    BinaryConfusionMatrix,
    ConfusionMatrix,
    MulticlassConfusionMatrix,
    MultilabelConfusionMatrix,
)
from torchmetrics_forked_b200.classification.f_beta import (
    BinaryF1Score,
    BinaryFBetaScore,
    F1Score,
    FBetaScore,
    MulticlassF1Score,
    MulticlassFBetaScore,
    MultilabelF1Score,
    MultilabelFBetaScore,
)
from torchmetrics_forked_b200.classification.jaccard import (
    BinaryJaccardIndex,
    JaccardIndex,
    MulticlassJaccardIndex,
    MultilabelJaccardIndex,
)
from torchmetrics_forked_b200.classification.precision_recall_curve import (
    BinaryPrecisionRecallCurve,
    MulticlassPrecisionRecallCurve,
    MultilabelPrecisionRecallCurve,
    PrecisionRecallCurve,
)
from torchmetrics_forked_b200.classification.roc import ROC, BinaryROC, MulticlassROC, MultilabelROC
from torchmetrics_forked_b200.classification.stat_scores import (
    BinaryStatScores,
    MulticlassStatScores,
    MultilabelStatScores,
    StatScores,
)

from torchmetrics_forked_b200.classification.cohen_kappa import BinaryCohenKappa, CohenKappa, MulticlassCohenKappa
from torchmetrics_forked_b200.classification.hamming import (
    BinaryHammingDistance,
    HammingDistance,
    MulticlassHammingDistance,
    MultilabelHammingDistance,
)
from torchmetrics_forked_b200.classification.matthews_corrcoef import (
    BinaryMatthewsCorrCoef,
    MatthewsCorrCoef,
    MulticlassMatthewsCorrCoef,
    MultilabelMatthewsCorrCoef,
)
from torchmetrics_forked_b200.classification.precision_fixed_recall import (
    BinaryPrecisionAtFixedRecall,
    MulticlassPrecisionAtFixedRecall,
    MultilabelPrecisionAtFixedRecall,
    PrecisionAtFixedRecall,
)
from torchmetrics_forked_b200.classification.precision_recall import (
    BinaryPrecision,
    BinaryRecall,
    MulticlassPrecision,
    MulticlassRecall,
    MultilabelPrecision,
    MultilabelRecall,
    Precision,
    Recall,
)
from torchmetrics_forked_b200.classification.recall_fixed_precision import (
    BinaryRecallAtFixedPrecision,
    MulticlassRecallAtFixedPrecision,
    MultilabelRecallAtFixedPrecision,
    RecallAtFixedPrecision,
)
from torchmetrics_forked_b200.classification.specificity import (
    BinarySpecificity,
    MulticlassSpecificity,
    MultilabelSpecificity,
    Specificity,
)
from torchmetrics_forked_b200.classification.specificity_sensitivity import (
    BinarySpecificityAtSensitivity,
    MulticlassSpecificityAtSensitivity,
    MultilabelSpecificityAtSensitivity,
    SpecificityAtSensitivity,
)

__all__ = [
    "Accuracy", "BinaryAccuracy", "MulticlassAccuracy", "MultilabelAccuracy",
    "AUROC", "BinaryAUROC", "MultilabelAUROC",
    "AveragePrecision", "BinaryAveragePrecision", "MultilabelAveragePrecision",
    "BinaryConfusionMatrix", "ConfusionMatrix", "MulticlassConfusionMatrix", "MultilabelConfusionMatrix",
    "BinaryF1Score", "BinaryFBetaScore", "F1Score", "FBetaScore", "MulticlassF1Score", "MulticlassFBetaScore",
    "MultilabelF1Score", "MultilabelFBetaScore",
    "BinaryJaccardIndex", "JaccardIndex", "MulticlassJaccardIndex", "MultilabelJaccardIndex",
    "BinaryPrecisionRecallCurve", "MultilabelPrecisionRecallCurve", "PrecisionRecallCurve",
    "ROC", "BinaryROC", "MultilabelROC",
    "BinaryStatScores", "MulticlassStatScores", "MultilabelStatScores", "StatScores",
    "BinaryCohenKappa", "CohenKappa", "MulticlassCohenKappa", "BinaryHammingDistance", "HammingDistance", "MulticlassHammingDistance", "MultilabelHammingDistance", "BinaryMatthewsCorrCoef", "MatthewsCorrCoef", "MulticlassMatthewsCorrCoef", "MultilabelMatthewsCorrCoef", "BinaryPrecisionAtFixedRecall", "MultilabelPrecisionAtFixedRecall", "PrecisionAtFixedRecall", "BinaryPrecision", "BinaryRecall", "MulticlassPrecision", "MulticlassRecall", "MultilabelPrecision", "MultilabelRecall", "Precision", "Recall", "BinaryRecallAtFixedPrecision", "MultilabelRecallAtFixedPrecision", "RecallAtFixedPrecision", "BinarySpecificity", "MulticlassSpecificity", "MultilabelSpecificity", "Specificity", "BinarySpecificityAtSensitivity", "MultilabelSpecificityAtSensitivity", "SpecificityAtSensitivity",
    "MulticlassAUROC", "MulticlassROC", "MulticlassAveragePrecision", "MulticlassPrecisionRecallCurve", "MulticlassPrecisionAtFixedRecall", "MulticlassRecallAtFixedPrecision", "MulticlassSpecificityAtSensitivity",
]  # fmt: skip
