"""Stand-in for the un-vendored dependency `lightning-utilities` (pinned >=0.8.0,<0.10.0 by the reference,
requirements/base.txt:9).  TEST INFRASTRUCTURE ONLY: it exists so that the unmodified reference under
/root/reference can be imported in the build container to pin the oracle and to generate golden vectors.
Written from the documented behaviour of the four symbols the reference uses; no hot-path arithmetic.
"""
from lightning_utilities.core.apply_func import apply_to_collection  # noqa: F401
