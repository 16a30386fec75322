"""``precision_at_fixed_recall`` (reference: functional/classification/precision_fixed_recall.py)."""
from torchmetrics_forked_b200.functional.classification._fixed_point import (  # noqa: F401
    _precision_at_recall,
    fixed_point_family,
)

(binary_precision_at_fixed_recall, multiclass_precision_at_fixed_recall, multilabel_precision_at_fixed_recall,
 precision_at_fixed_recall) = fixed_point_family("precision_at_fixed_recall", "min_recall", _precision_at_recall)
for _f in (binary_precision_at_fixed_recall, multiclass_precision_at_fixed_recall, multilabel_precision_at_fixed_recall, precision_at_fixed_recall):
    _f.__module__ = __name__
