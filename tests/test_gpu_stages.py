"""Stage-by-stage parity of the CUDA path against the CPU oracle on identical inputs.

GEMM rounding is taken out of the picture by handing the oracle the distance matrix the GPU
produced (the reference offers the same hook: re_ranking(..., local_distmat=D, only_local=True),
reranking.py:33-34), so every stage after the contraction must agree BIT FOR BIT."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

CASES = [("tiny", 20, 6, 0), ("tiny", 50, 15, 1), ("tiny", 5, 1, 2), ("tiny", 25, 3, 3),
         ("small", 20, 6, 0), ("small", 50, 15, 4), ("small", 7, 2, 5)]


def _features(synth, name, seed, device):
    sh = synth.SHAPES[name]
    qf, gf, qp, gp, qc, gc = synth.make(sh, seed)
    return sh, torch.cat([qf, gf]).to(device), qp, gp, qc, gc


def test_prepare_matches_normalize(eng, synth, oracle):
    sh, X, *_ = _features(synth, "small", 0, eng.device)
    f = eng.prepare(X, True)
    got = f.x32[:, : sh.d].cpu().numpy()
    want = oracle.l2_normalize(X.cpu().numpy())
    np.testing.assert_allclose(got, want, rtol=2e-6, atol=1e-8)
    assert f.dp % 64 == 0 and torch.all(f.x32[:, sh.d:] == 0)
    assert torch.all(f.hi[:, sh.d:] == 0) and torch.all(f.lo[:, sh.d:] == 0)
    np.testing.assert_allclose(f.sq.cpu().numpy(), (want.astype(np.float64) ** 2).sum(1), rtol=1e-5)
    # float16 split: x * 2^e = hi + lo to ~22 bits.  Normalised rows are bounded by 1, so the scale is fixed up front
    # (e = 14, independent of the other rows); raw rows take it from the measured maximum: max|x| * 2^e in [2^14, 2^15)
    def check_split(ff, e):
        hi = ff.hi.cpu().numpy().astype(np.float64)
        lo = ff.lo.cpu().numpy().astype(np.float64)
        x = ff.x32.cpu().numpy().astype(np.float64) * 2.0 ** e
        assert np.abs(hi + lo - x).max() <= 2.0 ** -11 * np.abs(lo).max() + 2.0 ** -25
        assert np.all(np.abs(lo) <= 2.0 ** -11 * np.abs(hi) + 2.0 ** -24)
        return np.abs(hi).max()
    assert f.amax.cpu().numpy().view(np.float32)[0] == 1.0
    assert check_split(f, 14) < 2.0 ** 15
    # raw (un-normalised) preparation reproduces the input exactly
    f2 = eng.prepare(X, False)
    assert torch.equal(f2.x32[:, : sh.d], X)
    amax = f2.amax.cpu().numpy().view(np.float32)[0]
    assert amax == np.abs(X.cpu().numpy()).max()
    assert 2.0 ** 14 <= check_split(f2, 14 - int(np.floor(np.log2(amax)))) < 2.0 ** 15


@pytest.mark.parametrize("engine_name", ["simt", "tcgen05"])
@pytest.mark.parametrize("name,seed", [("tiny", 0), ("small", 1), ("mid", 2)])
def test_sq_euclidean_and_arccos(eng, engine_mod, synth, oracle, name, seed, engine_name):
    sh, X, *_ = _features(synth, name, seed, eng.device)
    code = engine_mod.ENGINES[engine_name]
    f = eng.prepare(X, True)
    xn = f.x32[:, : sh.d].cpu().numpy()
    q = sh.q
    A, B = f.rows(slice(0, q)), f.rows(slice(q, None))
    D, key = eng.dist(A, B, engine_mod.METRIC_SQEUCLIDEAN, want_rowmax=True, engine=code)
    got = D[:, : sh.g].cpu().numpy()
    want = oracle.sq_euclidean(xn[:q], xn[q:])
    ref64 = ((xn[:q, None, :].astype(np.float64) - xn[None, q:, :].astype(np.float64)) ** 2).sum(-1) \
        if name != "mid" else None
    # north_star tolerance: distances within 1e-4 relative; fp32 accumulation is far tighter
    np.testing.assert_allclose(got, want, rtol=1e-4, atol=2e-6)
    assert np.abs(got - want).max() < 5e-6
    if ref64 is not None:
        assert np.abs(got - ref64).max() < 5e-6
    np.testing.assert_array_equal(eng.rowmax_values(key).cpu().numpy(), got.max(1))
    # cosine / arccos on the raw (un-normalised) features, as evaluate.py:16-26 is called
    fr = eng.prepare(X, False)
    A, B = fr.rows(slice(0, q)), fr.rows(slice(q, None))
    Dc, _ = eng.dist(A, B, engine_mod.METRIC_ARCCOS, engine=code)
    xr = X.cpu().numpy()
    np.testing.assert_allclose(Dc[:, : sh.g].cpu().numpy(), oracle.arccos_distance(xr[:q], xr[q:]),
                               rtol=1e-4, atol=1e-5)


@pytest.mark.parametrize("n,d,normalize", [(1000, 512, True), (777, 768, True), (1500, 1280, False), (300, 100, False)])
def test_tensor_core_contraction_accuracy(eng, engine_mod, n, d, normalize):
    """float16-split tcgen05 contraction vs float64.  The tensor core's float32 accumulator truncates
    (round-toward-zero) once per MMA step, which shows up as a smooth bias of up to ~4e-6 on large
    dot products (measured; 96 steps at d=512) on top of a ~1e-7 random part — the random part, which
    is what can reorder neighbours, is at the float32-sgemm level.  Bound: 1.2e-5 of the largest
    squared norm, i.e. ~8x inside the 1e-4 relative tolerance of north_star."""
    g = torch.Generator().manual_seed(n)
    X = (torch.randn(n, d, generator=g) * (1.0 if normalize else 3.0)).to(eng.device)
    X[::7] *= 0.01                                        # wide dynamic range across rows
    f = eng.prepare(X, normalize)
    x = f.x32[:, :d].double()
    ref = (x * x).sum(1)[:, None] + (x * x).sum(1)[None, :] - 2.0 * (x @ x.T)
    scale = float((x * x).sum(1).max())
    for name in ("tcgen05", "simt"):
        D, key = eng.dist(f, f, engine_mod.METRIC_SQEUCLIDEAN, want_rowmax=True, engine=engine_mod.ENGINES[name])
        got = D[:, :n].double()
        err = float((got - ref).abs().max()) / scale
        assert err < (1.0e-5 * max(1.0, f.dp / 512) if name == "tcgen05" else 4e-6), (name, err)
        np.testing.assert_array_equal(eng.rowmax_values(key).cpu().numpy(), D[:, :n].max(1).values.cpu().numpy())
    # strided row window (what a rank of a multi-GPU job passes): same numbers as the full tile schedule
    eng.symmetric = False
    Dfull, kfull, cfull = eng.dist(f, f, engine_mod.METRIC_SQEUCLIDEAN, want_rowmax=True, want_chunk_min=True)
    eng.symmetric = True
    Dwin, _ = eng.dist(f.rows(slice(1, n, 3)), f, engine_mod.METRIC_SQEUCLIDEAN)
    assert torch.equal(Dwin[:, :n], Dfull[1::3, :n])
    # symmetric schedule (upper-triangle tiles + mirrored stores): tiles of the upper triangle and the diagonal blocks
    # carry the bits of the full schedule, the rest is their exact mirror (within an accumulator ulp of the full
    # schedule's own lower triangle); row maxima and chunk minima describe ITS matrix
    Dsym, ksym, csym = eng.dist(f, f, engine_mod.METRIC_SQEUCLIDEAN, want_rowmax=True, want_chunk_min=True)
    S, Fm = Dsym[:, :n], Dfull[:, :n]
    blk = torch.arange(n, device=S.device) >> 8                      # 256-wide diagonal blocks are computed directly
    upper = blk[:, None] <= blk[None, :]
    assert torch.equal(S[upper], Fm[upper])                           # direct tiles: the same bits as the full schedule
    off = blk[:, None] != blk[None, :]
    assert torch.equal(S[off], S.T[off])                              # everything else is an exact mirror
    assert float((S - Fm).abs().max()) <= 4e-6 * scale
    np.testing.assert_array_equal(eng.rowmax_values(ksym).cpu().numpy(), S.max(1).values.cpu().numpy())
    pad = (-n) % 32
    Sp = torch.nn.functional.pad(S, (0, pad), value=float("inf")).reshape(n, -1, 32).min(2).values
    assert torch.equal(csym, Sp)


def _panel(eng, engine_mod, X, normalize=True):
    f = eng.prepare(X, normalize)
    D, key = eng.dist(f, f, engine_mod.METRIC_SQEUCLIDEAN, want_rowmax=True)
    return D, key


@pytest.mark.parametrize("name,k1,k2,seed", CASES)
def test_rerank_stages_bit_exact(eng, engine_mod, synth, oracle, name, k1, k2, seed):
    sh, X, qp, gp, qc, gc = _features(synth, name, seed, eng.device)
    N, q = sh.q + sh.g, sh.q
    D, key = _panel(eng, engine_mod, X)
    Dh = D[:, :N].cpu().numpy()
    det = {}
    final = eng.rerank_from_panel(D, key, q, N, k1, k2, 0.3, detail=det)[:, : sh.g].cpu().numpy()
    ref = oracle.rerank_from_dist(np.ascontiguousarray(Dh.T), q, k1, k2, 0.3, detail=True)

    # K4: row maxima / K5: neighbour lists (stable (value, index) order on both sides)
    np.testing.assert_array_equal(eng.rowmax_values(key).cpu().numpy(), Dh.max(1))
    K = max(k1 + 1, k2)
    np.testing.assert_array_equal(det["nbr"].cpu().numpy(), ref["nbr"][:, :K])
    # K6: expansion sets and float16 weights
    Vc, Vv, Vn = (det[k].cpu().numpy() for k in ("Vc", "Vv", "Vn"))
    for i in range(N):
        np.testing.assert_array_equal(Vc[i, : Vn[i]], ref["E"][i], err_msg=f"E[{i}]")
        np.testing.assert_array_equal(Vv[i, : Vn[i]].view(np.uint16), ref["V"][i].view(np.uint16),
                                      err_msg=f"V[{i}]")
    # K7: query expansion
    Qc, Qv, qoff, Qn = (det[k].cpu().numpy() for k in ("Qc", "Qv", "qoff", "Qn"))
    for i in range(N):
        a, n = qoff[i], Qn[i]
        np.testing.assert_array_equal(Qc[a:a + n], ref["Qc"][i], err_msg=f"Qc[{i}]")
        np.testing.assert_array_equal(Qv[a:a + n].view(np.uint16), ref["Qv"][i].view(np.uint16),
                                      err_msg=f"Qv[{i}]")
    # K8: inverted index (as sets per column; order inside a column is free)
    colptr, cent = (t.cpu().numpy() for t in det["csc"])
    crow = (cent & 0xffffffff).astype(np.int32)
    cval = ((cent >> 32) & 0xffff).astype(np.uint16).view(np.float16)
    rptr, rrows, rvals = ref["inv"]
    np.testing.assert_array_equal(colptr, rptr)
    for c in range(0, N, max(1, N // 50)):
        a, b = colptr[c], colptr[c + 1]
        o = np.argsort(crow[a:b], kind="stable")
        np.testing.assert_array_equal(crow[a:b][o], rrows[a:b])
        np.testing.assert_array_equal(cval[a:b][o].view(np.uint16), rvals[a:b].view(np.uint16))
    # K9: float16 Jaccard rows, K10: lambda mix — bit exact
    np.testing.assert_array_equal(det["J"].cpu().numpy().view(np.uint16), ref["J"].view(np.uint16))
    np.testing.assert_array_equal(final.view(np.uint32), ref["final"].view(np.uint32))

    # K11 / K12 on the re-ranked matrix
    cmc_o, map_o, aps_o, first_o = oracle.cmc_map(ref["final"], qp.numpy(), gp.numpy(), qc.numpy(), gc.numpy(),
                                                  detail=True)
    dev = eng.device
    fin = torch.from_numpy(ref["final"]).to(dev)
    ranked = eng.rank_rows(fin, sh.g).cpu().numpy()
    np.testing.assert_array_equal(ranked, np.argsort(ref["final"], axis=1, kind="stable"))
    ids = (qp.to(dev), gp.to(dev), qc.to(dev), gc.to(dev))
    slow = eng.evaluate_dist(fin, sh.g, *ids, min(50, sh.g), materialize_ranking=True)
    cmc, mAP, nvalid, ap, first = eng.evaluate_dist(fin, sh.g, *ids, min(50, sh.g))
    # the fused counting kernel and the explicit ranking + scan agree exactly
    assert torch.equal(cmc, slow[0]) and torch.equal(first, slow[4]) and torch.equal(nvalid, slow[2])
    np.testing.assert_allclose(ap.cpu().numpy(), slow[3].cpu().numpy(), rtol=0, atol=1e-15)
    np.testing.assert_array_equal(cmc.cpu().numpy(), cmc_o)
    assert abs(mAP.item() - map_o) < 1e-12
    np.testing.assert_array_equal(first.cpu().numpy(), first_o)
    ap = ap.cpu().numpy()
    valid = ~np.isnan(aps_o)
    np.testing.assert_allclose(ap[valid], aps_o[valid], rtol=0, atol=1e-14)
    assert np.all(ap[~valid] == -1.0)
    assert int(nvalid.item()) == int(valid.sum())

    # fused K9-K12 (the evaluator's path): same matrix bits when asked for it, same metrics either way
    mr = min(50, sh.g)
    saved_cap = eng.sparse_cap
    for want_final, cap in ((True, saved_cap), (False, saved_cap), (True, 0), (False, 0), (False, 64), (True, 64)):
        eng.sparse_cap = cap               # packed records / dense rows / records with most queries overflowing
        fin2, res = eng.rerank_from_panel(D, key, q, N, k1, k2, 0.3, score=(ids, mr), want_final=want_final)
        eng.sparse_cap = saved_cap
        if want_final:
            np.testing.assert_array_equal(fin2[:, : sh.g].cpu().numpy().view(np.uint32), ref["final"].view(np.uint32))
        else:
            assert fin2 is None
        cmc2, map2, nv2, ap2, first2 = res
        assert torch.equal(cmc2, cmc) and torch.equal(first2, first) and torch.equal(nv2, nvalid), want_final
        np.testing.assert_allclose(ap2.cpu().numpy(), ap, rtol=0, atol=1e-14)
        assert abs(map2.item() - map_o) < 1e-12


def test_fused_score_many_matches_junk_and_odd_ids(eng, engine_mod, synth, oracle):
    """Identity groups larger than one shared-memory slice (512), camera junk, absent identities, negative and
    INT64_MIN ids: the fused Jaccard+score kernel must agree with the explicit ranking on its own matrix."""
    sh, X, *_ = _features(synth, "mid", 5, eng.device)
    N, q, g = sh.q + sh.g, sh.q, sh.g
    rng = np.random.default_rng(11)
    odd = np.array([-(2 ** 63), -7, 0, 3, 2 ** 40], dtype=np.int64)
    g_pids = odd[rng.integers(0, 4, g)]                # ~1075 gallery entries per identity; 2**40 never appears
    q_pids = odd[rng.integers(0, 5, q)]                # a fifth of the queries have no match at all
    g_cams = rng.integers(0, 3, g).astype(np.int64)
    q_cams = rng.integers(0, 3, q).astype(np.int64)
    dev = eng.device
    ids = tuple(torch.from_numpy(a).to(dev) for a in (q_pids, g_pids, q_cams, g_cams))
    D, key = _panel(eng, engine_mod, X)
    saved_cap = eng.sparse_cap
    for k1, k2, cap in ((20, 6, saved_cap), (6, 1, saved_cap), (20, 6, 0), (20, 6, 1024)):
        eng.sparse_cap = 0
        final = eng.rerank_from_panel(D, key, q, N, k1, k2, 0.3)
        eng.sparse_cap = cap
        assert torch.equal(eng.rerank_from_panel(D, key, q, N, k1, k2, 0.3)[:, :g], final[:, :g])
        want = eng.evaluate_dist(final, g, *ids, 50, materialize_ranking=True)
        want_cmc, want_map, want_ap, want_first = oracle.cmc_map(final[:, :g].cpu().numpy(), q_pids, g_pids, q_cams,
                                                                 g_cams, detail=True)
        for want_final in (False, True):
            fin2, got = eng.rerank_from_panel(D, key, q, N, k1, k2, 0.3, score=(ids, 50), want_final=want_final)
            if want_final:
                assert torch.equal(fin2[:, :g], final[:, :g])
            assert torch.equal(got[0], want[0]) and torch.equal(got[2], want[2]) and torch.equal(got[4], want[4])
            np.testing.assert_allclose(got[3].cpu().numpy(), want[3].cpu().numpy(), rtol=0, atol=1e-13)
            np.testing.assert_array_equal(got[0].cpu().numpy(), want_cmc)
            np.testing.assert_array_equal(got[4].cpu().numpy(), want_first)
            assert abs(got[1].item() - want_map) < 1e-12
            assert int(got[2].item()) == int((~np.isnan(want_ap)).sum()) < q
    eng.sparse_cap = saved_cap


def test_select_fallback_and_ties(eng, engine_mod, oracle):
    """Massive ties (duplicated rows) and a capped candidate buffer force the exact arg-min path."""
    g = torch.Generator().manual_seed(7)
    base = torch.randn(40, 32, generator=g)
    X = base.repeat(12, 1).to(eng.device)            # 480 rows, every row duplicated 12 times
    N = X.shape[0]
    D, key = _panel(eng, engine_mod, X)
    Dh = D[:, :N].cpu().numpy()
    O = oracle.normalized_rows(np.ascontiguousarray(Dh.T))
    want = oracle.neighbour_lists(O, 21)
    f = eng.prepare(X, True)
    D2, key2, cmin = eng.dist(f, f, engine_mod.METRIC_SQEUCLIDEAN, want_rowmax=True, want_chunk_min=True)
    assert torch.equal(D2, D) and torch.equal(key2, key)
    np.testing.assert_array_equal(cmin.cpu().numpy(),
                                  np.pad(Dh, ((0, 0), (0, -N % 32)), constant_values=np.inf).reshape(N, -1, 32).min(2))
    for cap in (0, 16):
        for hint in (None, cmin):
            nbr, val = eng.topk_rows(D, N, key, 21, want_values=True, cap=cap, chunk_min=hint)
            np.testing.assert_array_equal(nbr.cpu().numpy(), want, err_msg=f"cap={cap} hinted={hint is not None}")
            np.testing.assert_array_equal(val.cpu().numpy(), np.take_along_axis(O, want.astype(np.int64), 1))


@pytest.mark.parametrize("n,dup,K", [(1500, 1, 21), (700, 7, 21), (1100, 1, 70), (300, 1, 9)])
def test_select_scatter_form_matches_pull_form(eng, engine_mod, n, dup, K):
    """Upper-only matrix: the scatter form (thresholds, push, merge) == the pull form with mirror reads == the
    selection on the fully stored matrix, bit for bit, including rows whose candidates overflow the slots."""
    g = torch.Generator().manual_seed(n)
    X = torch.randn(n, 48, generator=g).repeat(dup, 1)
    X = X[torch.randperm(X.shape[0], generator=g)].to(eng.device)      # duplicates spread over the diagonal blocks
    N = X.shape[0]
    f = eng.prepare(X, True)
    saved = eng.symmetric, eng.scatter_select
    eng.symmetric = True
    try:
        D, key, cmin = eng.dist(f, f, engine_mod.METRIC_SQEUCLIDEAN, want_rowmax=True, want_chunk_min=True)
        want, wval = eng.topk_rows(D, N, key, K, want_values=True, chunk_min=cmin)
        D3, key3, cmin3 = eng.dist(f, f, engine_mod.METRIC_SQEUCLIDEAN, want_rowmax=True, want_chunk_min=True,
                                   upper_only=True)
        assert torch.equal(key3, key) and torch.equal(cmin3, cmin)
        for cap in (0, 12):
            for scatter in (True, False):
                eng.scatter_select = scatter
                nbr, val = eng.topk_rows(D3, N, key3, K, want_values=True, cap=cap, chunk_min=cmin3, upper_only=True)
                assert torch.equal(nbr, want), f"cap={cap} scatter={scatter}"
                assert torch.equal(val, wval), f"cap={cap} scatter={scatter}"
    finally:
        eng.symmetric, eng.scatter_select = saved


def test_qe_merge_fallback(eng, engine_mod, synth, oracle):
    """A tiny per-warp sort buffer forces the k2-way merge path; results must not change."""
    sh, X, *_ = _features(synth, "tiny", 3, eng.device)
    N = sh.q + sh.g
    D, key = _panel(eng, engine_mod, X)
    nbr = eng.topk_rows(D, N, key, 21)
    sets = eng.krecip_sets(nbr, 20)
    Vc, Vv, Vn = eng.krecip_V(sets, 20, D, key)
    a = eng.query_expansion(nbr, 6, Vc, Vv, Vn, N)
    b = eng.query_expansion(nbr, 6, Vc, Vv, Vn, N, cap_sort=32)
    assert torch.equal(a[3], b[3])
    qoff, Qn = a[2].cpu().numpy(), a[3].cpu().numpy()
    for i in range(N):
        s = slice(qoff[i], qoff[i] + Qn[i])
        assert torch.equal(a[0][s], b[0][s]) and torch.equal(a[1][s].view(torch.int16), b[1][s].view(torch.int16))


def test_score_dist_many_matches_and_ties(eng, oracle):
    """More matches than one shared-memory chunk (1024), exact distance ties, junk, skipped queries."""
    rng = np.random.default_rng(3)
    q, g = 9, 6000
    dist = np.round(rng.random((q, g), dtype=np.float32) * 50) / np.float32(50)      # heavy ties
    g_pids = rng.integers(0, 3, g)                     # ~2000 matches per query identity
    g_cams = rng.integers(0, 4, g)
    q_pids = np.array([0, 1, 2, 0, 1, 2, 7, 0, 1])     # query 6: identity absent -> skipped
    q_cams = np.array([0, 1, 2, 3, 0, 1, 2, 3, 0])
    want_cmc, want_map, want_ap, want_first = oracle.cmc_map(dist, q_pids, g_pids, q_cams, g_cams, detail=True)
    dev = eng.device
    t = lambda a, dt: torch.from_numpy(np.ascontiguousarray(a)).to(dev, dt)
    args = (t(q_pids, torch.int64), t(g_pids, torch.int64), t(q_cams, torch.int64), t(g_cams, torch.int64))
    d = torch.zeros((q, 6016), dtype=torch.float32, device=dev)
    d[:, :g] = t(dist, torch.float32)
    for slow in (False, True):
        cmc, mAP, nvalid, ap, first = eng.evaluate_dist(d, g, *args, 50, materialize_ranking=slow)
        np.testing.assert_array_equal(cmc.cpu().numpy(), want_cmc)
        np.testing.assert_array_equal(first.cpu().numpy(), want_first)
        assert abs(mAP.item() - want_map) < 1e-12 and int(nvalid.item()) == 8
        ap = ap.cpu().numpy()
        ok = ~np.isnan(want_ap)
        np.testing.assert_allclose(ap[ok], want_ap[ok], rtol=0, atol=1e-13)
        assert ap[6] == -1.0


@pytest.mark.parametrize("n", [0, 1, 31, 1000, 16384, 16385, 100001, 2200003])
def test_exclusive_scan_matches_numpy(eng, n):
    """reid_exclusive_scan (CSR / CSC offsets, identity index): single CTA and the chained multi-CTA schedule."""
    import ctypes as C
    import importlib
    _lib = importlib.import_module("multimodal-reid_b200._lib")
    rng = np.random.default_rng(n)
    x = rng.integers(0, 1000, n, dtype=np.int32)
    # one unaligned view as well: the kernel vectorises only 16-byte aligned groups
    for off in (0, 1):
        buf = torch.zeros(n + 4, dtype=torch.int32, device=eng.device)
        buf[off: off + n] = torch.from_numpy(x).to(eng.device)
        src = buf[off: off + n]
        out = torch.full((n + 1,), -1, dtype=torch.int64, device=eng.device)
        for _ in range(2):      # twice: the chain's epoch flags must not leak between launches
            _lib.call("reid_exclusive_scan", C.c_void_p(src.data_ptr()), n, C.c_void_p(out.data_ptr()),
                      C.c_void_p(torch.cuda.current_stream().cuda_stream))
        want = np.concatenate([[0], np.cumsum(x.astype(np.int64))])
        assert np.array_equal(out.cpu().numpy(), want)


@pytest.mark.parametrize("name,batch", [("duke", 0), ("mid", 300), ("small", 64)])
def test_streamed_blocks_match_single_launch(eng, engine_mod, synth, name, batch):
    """Host features copied chunk by chunk under the contraction (diagonal blocks in upper-only form + the rectangles
    above them with mirrored statistics) == prepare() + dist(upper_only=True) on the whole matrix, bit for bit: the
    stored half of D, the row maxima, the chunk minima, and everything downstream (CMC, mAP, AP, first hits)."""
    sh = synth.SHAPES[name]
    qf, gf, qp, gp, qc, gc = synth.make(sh, 1)
    Xh = torch.cat([qf, gf]).pin_memory()
    N, q = Xh.shape[0], sh.q
    parts = [Xh] if batch == 0 else [Xh[i: i + batch] for i in range(0, N, batch)]
    dev = eng.device
    pids, cams = torch.cat([qp, gp]).to(dev), torch.cat([qc, gc]).to(dev)
    saved = eng.symmetric
    eng.symmetric = True
    try:
        f = eng.prepare(Xh.to(dev), True)
        D, key, cmin = eng.dist(f, f, engine_mod.METRIC_SQEUCLIDEAN, want_rowmax=True, want_chunk_min=True,
                                upper_only=True)
        old_bounds = engine_mod.Engine.stream_bounds
        if N < 4096:                                   # small shapes: force several (uneven) chunks anyway
            engine_mod.Engine.stream_bounds = staticmethod(
                lambda n: sorted({0, min(n, 256), min(n, 768), n}))
        try:
            D2, key2, cmin2, N2 = eng.prepare_dist_streamed(parts, True)
            res2 = eng.evaluate_reranked_host(parts, q, pids, cams, sh.k1, sh.k2, sh.lam, 50)
        finally:
            engine_mod.Engine.stream_bounds = staticmethod(old_bounds)
        assert N2 == N and torch.equal(key2, key) and torch.equal(cmin2, cmin)
        rows = torch.arange(N, device=dev)[:, None]
        cols = torch.arange(D.shape[1], device=dev)[None, :]
        stored = (cols >= (rows // 256) * 256) & (cols < N)
        assert torch.equal(D2[stored], D[stored])
        res = eng.evaluate_reranked(Xh.to(dev), q, pids, cams, sh.k1, sh.k2, sh.lam, normalize=True, max_rank=50)
        for u, v, what in zip(res, res2, ("cmc", "mAP", "n_valid", "ap", "first_hit")):
            assert torch.equal(u, v), f"{name}: {what} differs between the streamed and the resident evaluation"
    finally:
        eng.symmetric = saved
