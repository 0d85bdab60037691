"""Builds libreid_b200.so (sm_100a only) in-tree with nvcc.  `python -m` friendly and importable."""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libreid_b200.so")
SOURCES = ["api.cu", "prep.cu", "dist_simt.cu", "dist_tc2.cu", "select.cu", "krecip.cu", "qe.cu",
           "jaccard.cu", "jaccard_score.cu", "jaccard_sparse.cu", "rank.cu", "score.cu", "train.cu", "stream_score.cu", "stream_rerank.cu", "stream_sym.cu", "no_alloc.cu", "fabric.cu", "eval_e2e.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden", "--threads", "0"]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found; reid_b200 has no prebuilt or CPU fallback")


def source_hash() -> str:
    """sha256 over the CUDA sources and the public header; baked into the library (reid_source_hash) so that a stale
    .so is recognised by content, not by file times (a snapshot copy does not keep them)."""
    import hashlib
    h = hashlib.sha256()
    files = sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cu", ".cuh", ".h")))
    files.append(os.path.join(HERE, "..", "include", "reid_b200.h"))
    for f in files:
        h.update(os.path.basename(f).encode())
        with open(f, "rb") as fh:
            h.update(fh.read())
    return h.hexdigest()[:32]


def built_hash() -> str | None:
    """Source hash recorded beside the library by the build that produced it."""
    try:
        with open(LIB + ".srchash") as f:
            return f.read().strip()
    except OSError:
        return None


def stale() -> bool:
    return not os.path.exists(LIB) or built_hash() != source_hash()


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not stale():
        return LIB
    objdir = os.path.join(HERE, "build")
    os.makedirs(objdir, exist_ok=True)
    nvcc = _nvcc()
    srchash = source_hash()
    objs = []
    procs = []
    for src in SOURCES:
        obj = os.path.join(objdir, src.replace(".cu", ".o"))
        objs.append(obj)
        cmd = [nvcc, *NVCC_FLAGS, *os.environ.get("REID_NVCC_EXTRA", "").split(), "-c", os.path.join(CSRC, src), "-o", obj]
        if src == "api.cu":
            cmd.insert(1, f'-DREID_SRC_HASH="{srchash}"')
        if verbose:
            cmd.insert(1, "-Xptxas")
            cmd.insert(2, "-v")
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    failed = False
    for src, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0 or verbose:
            sys.stderr.write(f"--- {src}\n{out}\n")
        failed |= p.returncode != 0
    if failed:
        raise RuntimeError("nvcc failed")
    cmd = [nvcc, "-shared", "-o", LIB, *objs, "-gencode", "arch=compute_100a,code=sm_100a",
           "-Xcompiler", "-fPIC", "-cudart", "static", "-ldl"]
    subprocess.check_call(cmd)
    with open(LIB + ".srchash", "w") as f:
        f.write(srchash)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
