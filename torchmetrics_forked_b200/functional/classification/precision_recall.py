"""Precision and recall (reference: functional/classification/precision_recall.py)."""
from torchmetrics_forked_b200.functional.classification._stat_family import (  # noqa: F401
    _precision_recall_reduce,
    _precision_reduce,
    _recall_reduce,
    functional_family,
)

binary_precision, multiclass_precision, multilabel_precision, precision = functional_family(
    "precision", _precision_reduce
)
binary_recall, multiclass_recall, multilabel_recall, recall = functional_family("recall", _recall_reduce)
for _f in (binary_precision, multiclass_precision, multilabel_precision, precision, binary_recall,
           multiclass_recall, multilabel_recall, recall):
    _f.__module__ = __name__
