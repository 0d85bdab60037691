"""ctypes binding of libreid_b200.so (the C-ABI in include/reid_b200.h).

There is deliberately no fallback: if the shared library is missing it is built with nvcc, and if
that is impossible, or the device is not an sm_100 part, the call raises."""
from __future__ import annotations

import ctypes as C
import os
import re

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libreid_b200.so")
HEADER_PATH = os.path.join(_HERE, "..", "include", "reid_b200.h")

ABI_VERSION = 3          # bumped whenever a signature in include/reid_b200.h changes
_p = C.c_void_p
_i = C.c_int
_l = C.c_int64
_f = C.c_float

def _parse_header():
    """Prototypes are read from the public header so the binding cannot drift from it."""
    with open(HEADER_PATH) as f:
        text = f.read()
    text = re.sub(r"/\*.*?\*/", " ", text, flags=re.S)
    sigs = {}
    for ret, name, args in re.findall(r"REID_API\s+([\w\s\*]+?)\b(reid_\w+)\s*\(([^)]*)\)\s*;", text):
        argtypes = []
        for a in [x.strip() for x in args.split(",")]:
            if a in ("", "void"):
                continue
            if "*" in a:
                argtypes.append(_p)
            elif re.match(r"(const\s+)?int64_t\b", a):
                argtypes.append(_l)
            elif re.match(r"(const\s+)?(int|int32_t)\b", a):
                argtypes.append(_i)
            elif re.match(r"(const\s+)?float\b", a):
                argtypes.append(_f)
            elif re.match(r"(const\s+)?double\b", a):
                argtypes.append(C.c_double)
            else:
                raise ReidError(f"unhandled parameter type in header: {a!r}")
        restype = C.c_char_p if "char" in ret else (_l if "int64_t" in ret else C.c_int)
        sigs[name] = (restype, argtypes)
    return sigs


class ReidError(RuntimeError):
    pass


def header_symbols() -> list[str]:
    """Every function the public header declares (used by the symbol-coverage test)."""
    return sorted(_parse_header())


_lib = None


def load(build_if_missing: bool = True) -> C.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    from . import build as _build
    if _build.stale():
        # missing, or built from other sources than the ones beside it (hash baked in at build time): rebuild, never
        # run an old binary silently
        if not build_if_missing:
            raise ReidError(f"{LIB_PATH} is missing or stale (run `python multimodal-reid_b200/build.py`); "
                            "there is no CPU fallback")
        _build.build()
    lib = C.CDLL(LIB_PATH)
    for name, (restype, args) in _parse_header().items():
        fn = getattr(lib, name)
        fn.argtypes = args
        fn.restype = restype
    if lib.reid_abi_version() != ABI_VERSION:
        raise ReidError("libreid_b200.so ABI version mismatch; rebuild")
    if lib.reid_source_hash().decode() != _build.source_hash():
        raise ReidError("libreid_b200.so was built from different sources than the ones beside it; rebuild "
                        "(python multimodal-reid_b200/build.py --force)")
    _lib = lib
    return lib


def call(name: str, *args) -> None:
    lib = load()
    rc = getattr(lib, name)(*args)
    if rc != 0:
        msg = lib.reid_last_error().decode("utf-8", "replace")
        raise ReidError(f"{name} failed ({rc}): {msg}")
