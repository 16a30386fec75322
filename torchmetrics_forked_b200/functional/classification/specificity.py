"""Specificity (reference: functional/classification/specificity.py)."""
from torchmetrics_forked_b200.functional.classification._stat_family import (  # noqa: F401
    _specificity_reduce,
    functional_family,
)

binary_specificity, multiclass_specificity, multilabel_specificity, specificity = functional_family(
    "specificity", _specificity_reduce
)
for _f in (binary_specificity, multiclass_specificity, multilabel_specificity, specificity):
    _f.__module__ = __name__
