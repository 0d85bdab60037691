"""Selection on the upper-only matrix: scatter form vs pull form vs the fully stored matrix (CUDA events, L2 flushed)."""
import importlib, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
engine = importlib.import_module("multimodal-reid_b200.engine")
synth = importlib.import_module("multimodal-reid_b200.synth")
eng = engine.Engine(torch.device("cuda", 0))
flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
def timed(fn, n=5):
    ts = []
    for _ in range(n):
        flush.fill_(1)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); r = fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return sorted(ts)[len(ts) // 2], r
for name in os.environ.get("SHAPES", "duke,msmt17").split(","):
    sh = synth.SHAPES[name]
    qf, gf, *_ = synth.make(sh, 0)
    X = torch.cat([qf, gf]).cuda()
    N = X.shape[0]; K = max(sh.k1 + 1, sh.k2)
    f = eng.prepare(X, True)
    eng.symmetric = True
    td, (D, key, cmin) = timed(lambda: eng.dist(f, f, 0, want_rowmax=True, want_chunk_min=True))
    tf, want = timed(lambda: eng.topk_rows(D, N, key, K, chunk_min=cmin))
    del D
    tu, (D3, key3, cmin3) = timed(lambda: eng.dist(f, f, 0, want_rowmax=True, want_chunk_min=True, upper_only=True))
    eng.scatter_select = False
    tp, pull = timed(lambda: eng.topk_rows(D3, N, key3, K, chunk_min=cmin3, upper_only=True))
    eng.scatter_select = True
    ts, sc = timed(lambda: eng.topk_rows(D3, N, key3, K, chunk_min=cmin3, upper_only=True))
    nb = int(eng.lib.reid_topk_scatter_workspace_bytes(N, 0))
    ws = torch.empty(nb // 8 + 2, dtype=torch.int64, device="cuda")
    nbr = torch.empty((N, K), dtype=torch.int32, device="cuda")
    engine._lib.call("reid_topk_rows_scatter", D3.data_ptr(), N, D3.stride(0), key3.data_ptr(), cmin3.data_ptr(), K,
                     nbr.data_ptr(), None, 0, ws.data_ptr(), ws.numel() * 8, None)
    torch.cuda.synchronize()
    nch = (N + 31) // 32
    cnt = ws.view(torch.int32)[N: 2 * N]
    print(f"{name}: N={N} K={K} dist full {td:.3f} upper {tu:.3f} | select full {tf:.3f} pull {tp:.3f} scatter {ts:.3f} ms"
          f" | equal pull {torch.equal(pull, want)} scatter {torch.equal(sc, want)}"
          f" | candidates/row mean {cnt.float().mean().item():.1f} max {int(cnt.max())}", flush=True)
    del D3, ws
