"""Hamming distance (reference: functional/classification/hamming.py)."""
from torchmetrics_forked_b200.functional.classification._stat_family import (  # noqa: F401
    _hamming_distance_reduce,
    functional_family,
)

(binary_hamming_distance, multiclass_hamming_distance, multilabel_hamming_distance,
 hamming_distance) = functional_family("hamming_distance", _hamming_distance_reduce)
for _f in (binary_hamming_distance, multiclass_hamming_distance, multilabel_hamming_distance, hamming_distance):
    _f.__module__ = __name__
