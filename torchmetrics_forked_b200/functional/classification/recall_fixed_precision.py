"""``recall_at_fixed_precision`` (reference: functional/classification/recall_fixed_precision.py)."""
from torchmetrics_forked_b200.functional.classification._fixed_point import (  # noqa: F401
    _recall_at_precision,
    fixed_point_family,
)

(binary_recall_at_fixed_precision, multiclass_recall_at_fixed_precision, multilabel_recall_at_fixed_precision,
 recall_at_fixed_precision) = fixed_point_family("recall_at_fixed_precision", "min_precision", _recall_at_precision)
for _f in (binary_recall_at_fixed_precision, multiclass_recall_at_fixed_precision, multilabel_recall_at_fixed_precision, recall_at_fixed_precision):
    _f.__module__ = __name__
