"""Loads the unmodified reference modules by path under private names (never copied).

Only available in the build container (``/root/reference`` is not mounted on the GPU box);
tests that need it are skipped elsewhere and fall back to ``tests/golden``."""
from __future__ import annotations

import importlib.util
import os
import sys
import warnings

REF_ROOT = os.environ.get("REID_REFERENCE_ROOT", "/root/reference")


def available() -> bool:
    return os.path.isfile(os.path.join(REF_ROOT, "reranking.py"))


def _load(name: str, fname: str):
    spec = importlib.util.spec_from_file_location(name, os.path.join(REF_ROOT, fname))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


_cache = {}


def reference():
    """Returns (evaluate_module, reranking_module) of the reference."""
    if "mods" in _cache:
        return _cache["mods"]
    warnings.filterwarnings("ignore", message=".*addmm_ is deprecated.*")
    warnings.filterwarnings("ignore", category=UserWarning, module="_ref_reranking")
    rr = _load("_ref_reranking", "reranking.py")
    saved = sys.modules.get("reranking")
    sys.modules["reranking"] = rr          # evaluate.py:4 does `from reranking import re_ranking`
    try:
        ev = _load("_ref_evaluate", "evaluate.py")
    finally:
        if saved is None:
            del sys.modules["reranking"]
        else:
            sys.modules["reranking"] = saved
    _cache["mods"] = (ev, rr)
    return ev, rr
