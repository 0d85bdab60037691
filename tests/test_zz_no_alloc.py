"""GPU: reid_topk_rows_scatter_static -- the scatter selection with its workspace in a device array built into the library
(no cudaMalloc by caller or library, no counterpart in the reference) --
against reid_topk_rows_scatter with a caller-provided workspace, through the C-ABI.  Integer / index work: bit-exact
(torch.equal on neighbour lists and values), as in tests/test_gpu_stages.py.  Several calls in a row on different problems
(the array keeps the last call's contents), and the size limit.  Collected last (file name) and independent of the rest of
the suite: runs alone in a fresh process as well as after the other tests."""
import ctypes as C
import importlib

import pytest
import torch

pytestmark = pytest.mark.gpu


def _problem(eng, engine_mod, n, d, dup, seed):
    """Upper-only symmetric distance matrix of n * dup rows (duplicates -> ties), its row maxima and chunk minima."""
    g = torch.Generator().manual_seed(seed)
    X = torch.randn(n, d, generator=g).repeat(dup, 1)
    X = X[torch.randperm(X.shape[0], generator=g)].to(eng.device)
    f = eng.prepare(X, True)
    saved = eng.symmetric
    eng.symmetric = True
    try:
        D, key, cmin = eng.dist(f, f, engine_mod.METRIC_SQEUCLIDEAN, want_rowmax=True, want_chunk_min=True,
                                upper_only=True)
    finally:
        eng.symmetric = saved
    return D, key, cmin, X.shape[0]


def _reference_entry(eng, lib_mod, engine_mod, D, key, cmin, N, K):
    nbr = torch.empty((N, K), dtype=torch.int32, device=eng.device)
    val = torch.empty((N, K), dtype=torch.float32, device=eng.device)
    nbytes = int(lib_mod.load().reid_topk_scatter_workspace_bytes(N, 0))
    ws = torch.empty(((nbytes + 15) // 16 * 2,), dtype=torch.int64, device=eng.device)
    lib_mod.call("reid_topk_rows_scatter", engine_mod._ptr(D), N, D.stride(0), engine_mod._ptr(key), engine_mod._ptr(cmin),
                 K, engine_mod._ptr(nbr), engine_mod._ptr(val), 0, engine_mod._ptr(ws), ws.numel() * 8,
                 engine_mod._stream())
    return nbr, val


def test_static_scratch_entry_equals_workspace_entry():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    engine_mod = importlib.import_module("multimodal-reid_b200.engine")
    lib_mod = importlib.import_module("multimodal-reid_b200._lib")
    eng = engine_mod.Engine(engine="auto")
    lib = lib_mod.load()
    assert lib.reid_topk_static_max_rows() == 4096
    # different sizes and K back to back: a larger problem first, so that the smaller ones run over its leftovers
    cases = [(700, 48, 2, 21, 1), (300, 32, 1, 7, 2), (700, 48, 2, 21, 1), (2048, 64, 2, 51, 3)]
    for it, (n, d, dup, K, seed) in enumerate(cases):
        D, key, cmin, N = _problem(eng, engine_mod, n, d, dup, seed)
        want_i, want_v = _reference_entry(eng, lib_mod, engine_mod, D, key, cmin, N, K)
        got_i = torch.full((N, K), -7, dtype=torch.int32, device=eng.device)
        got_v = torch.full((N, K), float("nan"), dtype=torch.float32, device=eng.device)
        for rep in range(2):                         # two calls in a row on the same inputs as well
            torch.cuda.synchronize()
            free0 = torch.cuda.mem_get_info()[0]
            lib_mod.call("reid_topk_rows_scatter_static", engine_mod._ptr(D), N, D.stride(0), engine_mod._ptr(key),
                         engine_mod._ptr(cmin), K, engine_mod._ptr(got_i), engine_mod._ptr(got_v), engine_mod._stream())
            torch.cuda.synchronize()
            free1 = torch.cuda.mem_get_info()[0]     # read before torch allocates anything for the comparison
            if it > 0 or rep > 0:                    # (the very first call may load the module that holds the array)
                assert free1 >= free0, "the no-allocation entry point consumed device memory"
            assert torch.equal(got_i, want_i), f"case {it} rep {rep}: neighbour lists differ"
            assert torch.equal(got_v, want_v), f"case {it} rep {rep}: values differ"
            got_i.fill_(-7)
            got_v.fill_(float("nan"))
    # beyond the built-in scratch: refused before anything is launched, with the limit in the message
    dummy = torch.zeros(64, dtype=torch.float32, device=eng.device)
    idummy = torch.zeros(64, dtype=torch.int32, device=eng.device)
    with pytest.raises(lib_mod.ReidError, match="4096"):
        lib_mod.call("reid_topk_rows_scatter_static", engine_mod._ptr(dummy), 4097, 4128, engine_mod._ptr(idummy),
                     engine_mod._ptr(dummy), 5, engine_mod._ptr(idummy), engine_mod._ptr(dummy), C.c_void_p(0))
    torch.cuda.synchronize()
