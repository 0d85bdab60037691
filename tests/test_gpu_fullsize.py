"""Parity at BASELINE.json's FULL sizes (Duke, VeRi-776, Market-1501, MSMT17 shapes), where the CPU oracle would
take minutes to hours: size-independent properties of every stage, each checked against an independent
computation (float64 torch on sampled rows, torch's stable sort, the materialised path vs the fused one).

What must hold (tolerances are north_star's):
  * contraction: sampled rows within 1e-4 relative of a float64 evaluation; row maxima and 32-column chunk
    minima exact; the symmetric schedule's mirrored entries bit-equal to their originals; the full schedule
    within float32 rounding of the symmetric one;
  * selection: neighbour lists == a stable (value, index) sort of the normalised rows, exactly;
  * evaluator: the fused Jaccard + CMC/mAP kernel == scoring the materialised re-ranked matrix (CMC, first hits
    exact, AP to 1e-14), and the whole evaluation is deterministic bit for bit across runs.
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

FULL = ["duke", "veri", "market", "msmt17"]


def _sample_rows(N, q, rng, n=48):
    fixed = [0, 1, q - 1, q, q + 1, N - 1, N - 2, 255, 256, 257, (N // 256) * 256 - 1, min(N - 1, (N // 256) * 256)]
    rows = np.unique(np.concatenate([np.array(fixed), rng.integers(0, N, n)]))
    return rows[(rows >= 0) & (rows < N)]


@pytest.mark.parametrize("name", FULL)
def test_full_size_properties(eng, engine_mod, synth, name):
    sh = synth.SHAPES[name]
    dev = eng.device
    free, _ = torch.cuda.mem_get_info(dev)
    N, q, g, d = sh.q + sh.g, sh.q, sh.g, sh.d
    if free < 2.6 * 4 * N * N + (8 << 30):
        pytest.skip(f"{name}: needs {2.6 * 4 * N * N / 2**30:.0f} GiB of free HBM")
    qf, gf, qp, gp, qc, gc = synth.make(sh, 0)
    X = torch.cat([qf, gf]).to(dev)
    pids, cams = torch.cat([qp, gp]).to(dev), torch.cat([qc, gc]).to(dev)
    rng = np.random.default_rng(7)
    rows = _sample_rows(N, q, rng)
    rt = torch.from_numpy(rows).to(dev)

    # ---------------------------------------------------------------- contraction (symmetric schedule)
    f = eng.prepare(X, True)
    eng.symmetric = True
    D, key, cmin = eng.dist(f, f, engine_mod.METRIC_SQEUCLIDEAN, want_rowmax=True, want_chunk_min=True)
    x64 = f.x32[:, :d].double()
    sq = (x64 * x64).sum(1)
    ref = sq[rt, None] + sq[None, :] - 2.0 * (x64[rt] @ x64.T)            # float64, [rows, N]
    got = D[rt, :N].double()
    err = (got - ref).abs()
    assert bool((err <= 1e-4 * ref.abs() + 2e-5).all()), f"{name}: contraction off by {err.max().item():.3e}"
    # row maxima and chunk minima are exact functions of the stored matrix
    assert torch.equal(eng.rowmax_values(key)[rt], D[rt, :N].max(1).values)
    nch = (N + 31) // 32
    padded = torch.full((rows.size, nch * 32), float("inf"), device=dev)
    padded[:, :N] = D[rt, :N]
    assert torch.equal(cmin[rt], padded.view(rows.size, nch, 32).min(2).values)
    # mirrored entries are copies: D[i, j] == D[j, i] bit for bit outside the 256-wide diagonal blocks
    cols = torch.from_numpy(_sample_rows(N, q, np.random.default_rng(11))).to(dev)
    a = D[rt][:, cols]
    b = D[cols][:, rt].t()
    off_diag = (rt[:, None] // 256) != (cols[None, :] // 256)
    assert torch.equal(a[off_diag], b[off_diag])
    assert float((a - b).abs().max()) <= 4e-6          # inside a diagonal block both halves are computed
    # the full tile schedule (what a rank of a multi-GPU job runs) agrees to float32 rounding
    sym_rows = D[rt, :N].clone()
    sym_key = key.clone()
    # upper-only form (mirrored tiles not stored; selection and the V weights read the mirror image): same bits
    K = max(sh.k1 + 1, sh.k2)
    nbr1 = eng.topk_rows(D, N, key, K, chunk_min=cmin)
    sets1 = eng.krecip_sets(nbr1, sh.k1)
    V1 = eng.krecip_V(sets1, sh.k1, D, key)
    del D, cmin, a, b, padded, got, ref, err
    D3, key3, cmin3 = eng.dist(f, f, engine_mod.METRIC_SQEUCLIDEAN, want_rowmax=True, want_chunk_min=True,
                               upper_only=True)
    assert torch.equal(key3, key)
    nbr3 = eng.topk_rows(D3, N, key3, K, chunk_min=cmin3, upper_only=True)
    assert torch.equal(nbr3, nbr1), f"{name}: neighbour lists differ in upper-only form"
    V3 = eng.krecip_V(eng.krecip_sets(nbr3, sh.k1), sh.k1, D3, key3, upper_only=True)
    for u, v, what in zip(V1, V3, ("Vc", "Vv", "Vn")):
        used = torch.arange(u.shape[1], device=dev)[None, :] < V1[2][:, None] if u.dim() == 2 else None
        assert torch.equal(u[used], v[used]) if used is not None else torch.equal(u, v), f"{name}: {what} differs"
    del D3, cmin3, key3, nbr1, nbr3, V1, V3, sets1, key
    eng.symmetric = False
    D2, key2, cmin2 = eng.dist(f, f, engine_mod.METRIC_SQEUCLIDEAN, want_rowmax=True, want_chunk_min=True)
    eng.symmetric = True
    assert float((D2[rt, :N] - sym_rows).abs().max()) <= 4e-6
    assert float((eng.rowmax_values(key2) - eng.rowmax_values(sym_key)).abs().max()) <= 4e-6

    # ---------------------------------------------------------------- selection == stable sort of O = D / rowmax
    nbr = eng.topk_rows(D2, N, key2, K, chunk_min=cmin2)
    O = D2[rt, :N] / eng.rowmax_values(key2)[rt, None]
    O = O + 0.0                                                            # -0 ties with +0 like numpy's sort
    want = torch.sort(O, dim=1, stable=True).indices[:, :K]
    assert torch.equal(nbr[rt].long(), want), f"{name}: neighbour lists differ from a stable sort"
    assert float((nbr[:, 0] == torch.arange(N, device=dev)).float().mean()) > 0.999     # self is the nearest
    del D2, cmin2, O, want, nbr

    # ---------------------------------------------------------------- evaluator: fused == materialised, deterministic
    saved_min_rows, eng.upper_only_min_rows = eng.upper_only_min_rows, 0     # upper-only form at every size here
    r1 = eng.evaluate_reranked(X, q, pids, cams, sh.k1, sh.k2, sh.lam, normalize=True, max_rank=50)
    r1b = eng.evaluate_reranked(X, q, pids, cams, sh.k1, sh.k2, sh.lam, normalize=True, max_rank=50)
    for u, v, what in zip(r1, r1b, ("cmc", "mAP", "n_valid", "ap", "first_hit")):
        assert torch.equal(u, v), f"{name}: {what} differs between two runs"
    eng.upper_only = False                                                 # mirrored half stored: identical results
    r1c = eng.evaluate_reranked(X, q, pids, cams, sh.k1, sh.k2, sh.lam, normalize=True, max_rank=50)
    eng.upper_only = True
    eng.upper_only_min_rows = saved_min_rows
    for u, v, what in zip(r1, r1c, ("cmc", "mAP", "n_valid", "ap", "first_hit")):
        assert torch.equal(u, v), f"{name}: {what} differs between the upper-only and the full symmetric form"
    final = eng.rerank_features(X, q, sh.k1, sh.k2, sh.lam, normalize=True)
    assert final.shape[0] == q and bool(torch.isfinite(final[:, :g]).all())
    ids = (pids[:q].contiguous(), pids[q:].contiguous(), cams[:q].contiguous(), cams[q:].contiguous())
    r2 = eng.evaluate_dist(final, g, *ids, 50)
    assert torch.equal(r1[0], r2[0]) and torch.equal(r1[2], r2[2]) and torch.equal(r1[4], r2[4])
    np.testing.assert_allclose(r1[3].cpu().numpy(), r2[3].cpu().numpy(), rtol=0, atol=1e-14)
    assert abs(float(r1[1]) - float(r2[1])) < 1e-14
    if name != "msmt17":                                                   # explicit ranking: 4 radix passes over q x g
        r3 = eng.evaluate_dist(final, g, *ids, 50, materialize_ranking=True)
        assert torch.equal(r2[0], r3[0]) and torch.equal(r2[4], r3[4])
        np.testing.assert_allclose(r2[3].cpu().numpy(), r3[3].cpu().numpy(), rtol=0, atol=1e-14)
    # the re-ranked value of a pair the Jaccard term does not touch is fp16(1 - lambda) + fp32(lambda) * O exactly
    c1h, c2 = np.float32(np.float16(1 - sh.lam)), np.float32(sh.lam)
    assert float(final[:, :g].max()) <= float(c1h + c2) + 1e-6 and float(final[:, :g].min()) >= -1e-5
    cmc, mAP = r1[0].cpu().numpy(), float(r1[1])
    assert 0.0 < mAP <= 1.0 and np.all(np.diff(cmc) >= 0) and cmc[-1] <= 1.0 + 1e-6
    assert int(r1[2]) > 0.9 * q
