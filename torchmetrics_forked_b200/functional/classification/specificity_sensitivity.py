"""``specificity_at_sensitivity`` (reference: functional/classification/specificity_sensitivity.py)."""
from torchmetrics_forked_b200.functional.classification._fixed_point import (  # noqa: F401
    _specificity_at_sensitivity,
    fixed_point_family,
)



def _convert_fpr_to_specificity(fpr):
    """``1 - fpr`` (reference :42-44; its own tests import the helper)."""
    return 1 - fpr


(binary_specificity_at_sensitivity, multiclass_specificity_at_sensitivity, multilabel_specificity_at_sensitivity,
 specificity_at_sensitivity) = fixed_point_family("specificity_at_sensitivity", "min_sensitivity", None)
for _f in (binary_specificity_at_sensitivity, multiclass_specificity_at_sensitivity, multilabel_specificity_at_sensitivity, specificity_at_sensitivity):
    _f.__module__ = __name__
specicity_at_sensitivity = specificity_at_sensitivity  # the reference exports the dispatcher under this spelling
