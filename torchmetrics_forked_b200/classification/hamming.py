"""Hamming distance metric classes (reference: classification/hamming.py:35-528)."""
from torchmetrics_forked_b200.classification._stat_family import metric_family
from torchmetrics_forked_b200.functional.classification._stat_family import _hamming_distance_reduce

BinaryHammingDistance, MulticlassHammingDistance, MultilabelHammingDistance, HammingDistance = metric_family(
    "HammingDistance", _hamming_distance_reduce, False, __name__
)
