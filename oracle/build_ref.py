"""Recipe for `oracle/_ref/`: the UNMODIFIED reference files of the hot path, taken from where they lie.

TEST / BASELINE INFRASTRUCTURE ONLY.  The reference is pure Python (no compiled sources), so "building" it means
placing byte-identical copies of the modules the path consists of -- `/root/reference/evaluate.py`,
`/root/reference/reranking.py` and (for the training-side row f4) `/root/reference/losses.py` -- under `oracle/_ref/`.  That directory is git-ignored (reference sources never enter
this repository's history) but NOT gpurun-ignored, so the files travel to the GPU box with the built `.so`, where
`bench.py --impl reference` / `cpu_baseline` time them on the host cores (`kind: "reference"`) and the `-m gpu` parity
tests use them as a second checker beside the oracle port.  `__graft_entry__.build()` runs this when `/root/reference`
is mounted; on the GPU box (no `/root/reference`) the prebuilt directory is used as it arrived.

    python oracle/build_ref.py            # writes oracle/_ref/{evaluate,reranking}.py + MANIFEST.json (sha256)
"""
from __future__ import annotations

import hashlib
import importlib.util
import json
import os
import shutil
import sys
import warnings

HERE = os.path.dirname(os.path.abspath(__file__))
REF_ROOT = os.environ.get("REID_REFERENCE_ROOT", "/root/reference")
OUT = os.path.join(HERE, "_ref")
FILES = ("evaluate.py", "reranking.py", "losses.py")      # losses.py: SURVEY 8(f4), live comparison on the GPU box


def build(verbose: bool = False) -> bool:
    """Returns True when oracle/_ref is complete afterwards."""
    if not all(os.path.isfile(os.path.join(REF_ROOT, f)) for f in FILES):
        return available()
    os.makedirs(OUT, exist_ok=True)
    manifest = {"source": REF_ROOT, "files": {}}
    for f in FILES:
        src, dst = os.path.join(REF_ROOT, f), os.path.join(OUT, f)
        shutil.copyfile(src, dst)
        with open(dst, "rb") as fh:
            manifest["files"][f] = hashlib.sha256(fh.read()).hexdigest()
    with open(os.path.join(OUT, "MANIFEST.json"), "w") as fh:
        json.dump(manifest, fh, indent=1)
    if verbose:
        print("oracle/_ref:", manifest["files"])
    return True


def available() -> bool:
    return all(os.path.isfile(os.path.join(OUT, f)) for f in FILES)


_cache = {}


def load():
    """(evaluate_module, reranking_module) of the unmodified reference, loaded from oracle/_ref under private names
    (the repository root holds drop-in shims called `evaluate` / `reranking`; the reference's own
    `from reranking import re_ranking` (evaluate.py:4) is pointed at its sibling while it loads)."""
    if "mods" in _cache:
        return _cache["mods"]
    if not available():
        raise FileNotFoundError("oracle/_ref is missing: run `python oracle/build_ref.py` where /root/reference is mounted")
    warnings.filterwarnings("ignore", message=".*addmm_.*")

    def _load(name, fname):
        spec = importlib.util.spec_from_file_location(name, os.path.join(OUT, fname))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        return mod

    rr = _load("_ref_reranking_shipped", "reranking.py")
    saved = sys.modules.get("reranking")
    sys.modules["reranking"] = rr
    try:
        ev = _load("_ref_evaluate_shipped", "evaluate.py")
    finally:
        if saved is None:
            del sys.modules["reranking"]
        else:
            sys.modules["reranking"] = saved
    _cache["mods"] = (ev, rr)
    return ev, rr


def load_losses():
    """The unmodified reference losses.py (pure torch; its XBM variant calls .cuda(), so it only runs on the GPU box)."""
    if "losses" not in _cache:
        if not os.path.isfile(os.path.join(OUT, "losses.py")):
            raise FileNotFoundError("oracle/_ref/losses.py is missing: run `python oracle/build_ref.py`")
        spec = importlib.util.spec_from_file_location("_ref_losses_shipped", os.path.join(OUT, "losses.py"))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        _cache["losses"] = mod
    return _cache["losses"]


if __name__ == "__main__":
    ok = build(verbose=True)
    print("oracle/_ref", "ready" if ok else "NOT available (no /root/reference and no prebuilt copy)")
