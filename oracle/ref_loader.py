"""Import the UNMODIFIED reference (torchmetrics 1.3.0dev) next to the `lightning_utilities` stand-in.

TEST / MEASUREMENT INFRASTRUCTURE ONLY.  Two places can hold the reference:

* `/root/reference/src` — the build container.  Used by `oracle/gen_golden*.py` and by the CPU tests that pin the
  oracle against the reference itself.
* `oracle/_ref/reference.zip` — the staged, byte-identical copy made by `oracle/build_ref.py` (git-ignored, ships
  with the working tree; imported straight from the archive).  This is what exists on the GPU box; `bench.py --impl reference`, the `cpu_baseline` leg and the
  reference-on-CUDA column use it.

Nothing under `torchmetrics_forked_b200/` imports this module.
"""
import importlib
import sys
from pathlib import Path
from types import ModuleType
from typing import Optional

REFERENCE_SRC = Path("/root/reference/src")
HERE = Path(__file__).resolve().parent
SHIM = HERE / "shim"
STAGED = HERE / "_ref" / "reference.zip"


def reference_available() -> bool:
    """The live reference tree is mounted (build container only)."""
    return (REFERENCE_SRC / "torchmetrics" / "metric.py").exists()


def staged_available() -> bool:
    return STAGED.exists()


def load_reference() -> Optional[ModuleType]:
    """Return the reference `torchmetrics` module from /root/reference, or None when it is not mounted."""
    if not reference_available():
        return None
    for p in (str(SHIM), str(REFERENCE_SRC)):
        if p not in sys.path:
            sys.path.insert(0, p)
    return importlib.import_module("torchmetrics")


def load_staged_reference() -> Optional[ModuleType]:
    """Return the reference `torchmetrics` module from `oracle/_ref` (else from /root/reference), or None."""
    if staged_available():
        if str(STAGED) not in sys.path:
            sys.path.insert(0, str(STAGED))
        return importlib.import_module("torchmetrics")
    return load_reference()
