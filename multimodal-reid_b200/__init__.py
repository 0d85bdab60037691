"""reid_b200 — B200-native retrieval-evaluation hot path of SuperbTUM/Multimodal-ReID.

The directory name follows the build contract (``multimodal-reid_b200``); because of the hyphen,
import it with ``importlib.import_module("multimodal-reid_b200")`` or through the top-level
aliases ``reid_b200`` / ``evaluate`` / ``reranking`` at the repository root.

Nothing here computes on the CPU: the modules bind libreid_b200.so (hand-written sm_100a CUDA)
and raise if it, or a B200, is missing."""
from . import _lib  # noqa: F401
from . import synth  # noqa: F401

__all__ = ["evaluate", "reranking", "engine", "synth"]


def __getattr__(name):
    if name in ("evaluate", "reranking", "engine", "build"):
        import importlib
        return importlib.import_module(f"{__name__}.{name}")
    raise AttributeError(name)
