"""Exception and warning types of the public API (reference: utilities/exceptions.py:16-21)."""


class TorchMetricsUserError(Exception):
    """Raised when Metric API calls are combined in a way that cannot work (e.g. sync twice)."""


class TorchMetricsUserWarning(Warning):
    """Category for warnings that point at a questionable use of the API."""
