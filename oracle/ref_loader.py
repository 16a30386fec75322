"""Import the UNMODIFIED reference (torchmetrics 1.3.0dev under /root/reference/src) next to the shim.

TEST INFRASTRUCTURE ONLY.  Used in the build container by `oracle/gen_golden.py` and by the CPU tests that
pin the oracle against the reference itself.  /root/reference does not exist on the GPU box, so nothing on
the `-m gpu` / smoke / bench paths may call this.
"""
import importlib
import sys
from pathlib import Path
from types import ModuleType
from typing import Optional

REFERENCE_SRC = Path("/root/reference/src")
SHIM = Path(__file__).resolve().parent / "shim"


def reference_available() -> bool:
    return (REFERENCE_SRC / "torchmetrics" / "metric.py").exists()


def load_reference() -> Optional[ModuleType]:
    """Return the reference `torchmetrics` module, or None when /root/reference is not mounted."""
    if not reference_available():
        return None
    for p in (str(SHIM), str(REFERENCE_SRC)):
        if p not in sys.path:
            sys.path.insert(0, p)
    return importlib.import_module("torchmetrics")
