"""Specificity metric classes (reference: classification/specificity.py:31-509)."""
from torchmetrics_forked_b200.classification._stat_family import metric_family
from torchmetrics_forked_b200.functional.classification._stat_family import _specificity_reduce

BinarySpecificity, MulticlassSpecificity, MultilabelSpecificity, Specificity = metric_family(
    "Specificity", _specificity_reduce, True, __name__
)
