"""`package_available` / `compare_version` (stand-in, see package docstring)."""
import importlib
import importlib.util
import operator as _op
from typing import Callable

from packaging.version import Version


def package_available(package_name: str) -> bool:
    try:
        return importlib.util.find_spec(package_name) is not None
    except (ModuleNotFoundError, ValueError):
        return False


def compare_version(package: str, op: Callable, version: str, use_base_version: bool = False) -> bool:
    try:
        pkg = importlib.import_module(package)
    except (ImportError, ModuleNotFoundError):
        return False
    try:
        have = Version(getattr(pkg, "__version__"))
    except Exception:  # noqa: BLE001
        return True
    if use_base_version:
        have = Version(have.base_version)
    return op(have, Version(version))


__all__ = ["package_available", "compare_version", "_op"]
