"""Importable alias of the ``multimodal-reid_b200`` package (its directory name has a hyphen)."""
import importlib as _importlib
import sys as _sys

_pkg = _importlib.import_module("multimodal-reid_b200")
_sys.modules[__name__] = _pkg
