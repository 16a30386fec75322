"""Precision / Recall metric classes (reference: classification/precision_recall.py:38-1023)."""
from torchmetrics_forked_b200.classification._stat_family import metric_family
from torchmetrics_forked_b200.functional.classification._stat_family import _precision_reduce, _recall_reduce

BinaryPrecision, MulticlassPrecision, MultilabelPrecision, Precision = metric_family(
    "Precision", _precision_reduce, True, __name__
)
BinaryRecall, MulticlassRecall, MultilabelRecall, Recall = metric_family("Recall", _recall_reduce, True, __name__)
