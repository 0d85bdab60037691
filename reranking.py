"""Top-level shim for ``from reranking import re_ranking`` (evaluate.py:4 of the reference)."""
import importlib as _importlib

_impl = _importlib.import_module("multimodal-reid_b200.reranking")

re_ranking = _impl.re_ranking
