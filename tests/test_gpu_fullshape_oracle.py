"""GPU: END-TO-END parity at BASELINE.json's real configurations -- the CUDA path against the oracle (and, where
oracle/_ref arrived, against the unmodified reference itself) on the same synthetic features, the oracle running on
the GPU box's host cores (10-40 s per case).

north_star's tolerances, each written where it is applied:
  * final distances  <= 1e-4 relative, float16-boundary flips of the Jaccard chain counted and bounded (SURVEY 8a note 4)
  * CMC / mAP        <= 1e-5 absolute
  * index lists      equal modulo ties.  north_star says "tied within 1e-6"; the lambda-mixed matrix quantises the
                     Jaccard term to float16 (2^-11 spacing in [0.5, 1)), so an fp32-ulp difference in D moves a final
                     entry by a whole quantum and reorders it among neighbours closer than that.  The test therefore
                     (a) REPORTS the measured number of list positions that differ outside 1e-6 and (b) asserts the
                     relaxed rule this repository states in DESIGN 2: differing positions hold keys within 4 quanta.
"""
import contextlib
import dataclasses
import importlib
import io

import numpy as np
import pytest
import torch

from parity import METRIC_TOL, TIE_TOL, FP16_QUANTUM, assert_distances_close

pytestmark = [pytest.mark.gpu, pytest.mark.slow]

CASES = [("duke", 1, 20, 6), ("duke", 1, 50, 15), ("market", 1, 20, 6), ("veri", 1, 20, 6), ("msmt17", 3, 20, 6)]


@pytest.fixture(scope="module")
def api():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import evaluate
    import reranking
    return evaluate, reranking


def _shape(synth, name, scale, k1, k2):
    sh = synth.SHAPES[name]
    if scale > 1:
        sh = synth.scaled(sh, sh.q // scale, sh.g // scale)
    return dataclasses.replace(sh, k1=k1, k2=k2)


def _list_report(got, want, K, what):
    """Top-K lists of the re-ranked rows: positions that differ, split by the gap of the keys involved."""
    ig = np.argsort(got, axis=1, kind="stable")[:, :K]
    iw = np.argsort(want, axis=1, kind="stable")[:, :K]
    neq = ig != iw
    r = np.nonzero(neq)[0]
    gap = np.abs(want[r, ig[neq]].astype(np.float64) - want[r, iw[neq]].astype(np.float64))
    n_out = int((gap > TIE_TOL).sum())
    print(f"[parity] {what}: top-{K} positions {ig.size}, differing {int(neq.sum())}, outside 1e-6 ties {n_out} "
          f"({n_out / ig.size:.2e}), largest key gap {gap.max() if gap.size else 0.0:.3e}")
    assert np.all(gap <= 4 * FP16_QUANTUM), f"{what}: lists differ outside 4 float16 quanta (gap {gap.max():.3e})"
    assert n_out <= 2e-3 * ig.size, f"{what}: {n_out} of {ig.size} list positions differ outside 1e-6 ties"
    return n_out


@pytest.mark.parametrize("name,scale,k1,k2", CASES, ids=[f"{n}-1/{s}-k{a}-{b}" for n, s, a, b in CASES])
def test_full_shape_vs_oracle(api, synth, oracle, name, scale, k1, k2):
    """reranking.py:29-100 + evaluate.py:29-88 + evaluate.py:110-134 at the BASELINE shapes (configs[0..3])."""
    evaluate, reranking = api
    sh = _shape(synth, name, scale, k1, k2)
    qf, gf, qp, gp, qc, gc = synth.make(sh, 0)
    raw = torch.cat([qf, gf])
    f = torch.nn.functional.normalize(raw, dim=1, p=2)
    want = oracle.re_ranking(f[: sh.q].numpy(), f[sh.q:].numpy(), k1, k2, sh.lam)
    c_o, m_o = oracle.cmc_map(want, qp.numpy(), gp.numpy(), qc.numpy(), gc.numpy())
    # (1) the drop-in re_ranking: q x g float32 matrix
    got = reranking.re_ranking(f[: sh.q], f[sh.q:], k1, k2, sh.lam)
    assert got.shape == want.shape and got.dtype == np.float32
    n_flip, mx = assert_distances_close(got, want, max_flip_frac=5e-5, what=f"{sh.name} final")
    print(f"[parity] {sh.name} k1={k1} k2={k2}: {want.size} distances, max rel (non-flip) {mx:.2e}, fp16 flips {n_flip}")
    _list_report(got, want, 10, f"{sh.name} k1={k1} k2={k2}")
    # (2) eval_func on the GPU's own matrix, and the fused evaluator from raw features
    with contextlib.redirect_stdout(io.StringIO()):
        cmc, mAP = evaluate.eval_func(got, qp, gp, qc, gc)
        e = evaluate.R1_mAP_eval(sh.q, feat_norm=True, reranking=True, rerank_params=(k1, k2, sh.lam))
        e.reset()
        e.update((raw, torch.cat([qp, gp]), torch.cat([qc, gc])))
        cmc_e, map_e = e.compute()
    for c, m, what in ((cmc, mAP, "eval_func"), (cmc_e, map_e, "R1_mAP_eval")):
        assert np.abs(c - c_o).max() <= METRIC_TOL, f"{what}: CMC off by {np.abs(c - c_o).max():.2e}"       # 1e-5 abs
        assert abs(m - m_o) <= METRIC_TOL, f"{what}: mAP {m} vs oracle {m_o}"                               # 1e-5 abs


def test_cosine_duke_shape(api, synth, oracle):
    """evaluate.py:16-26 at the Duke shape (BASELINE configs[1] names both distances): distances within 1e-4 relative
    of the oracle, CMC / mAP of the arccos matrix within 1e-5."""
    evaluate, _ = api
    sh = synth.SHAPES["duke"]
    qf, gf, qp, gp, qc, gc = synth.make(sh, 0)
    got = evaluate.cosine_similarity(qf, gf)
    want = oracle.arccos_distance(qf.numpy(), gf.numpy())
    assert got.shape == want.shape and got.dtype == np.float32
    # arccos amplifies an ulp of the cosine near its clip points; the data sits around 1.2-1.6 rad
    np.testing.assert_allclose(got, want, rtol=1e-4, atol=1e-5)
    eu = evaluate.euclidean_distance(qf, gf)
    np.testing.assert_allclose(eu, oracle.sq_euclidean(qf.numpy(), gf.numpy()), rtol=1e-4, atol=1e-2 * 1e-4)
    with contextlib.redirect_stdout(io.StringIO()):
        cmc, mAP = evaluate.eval_func(got, qp, gp, qc, gc)
    c_o, m_o = oracle.cmc_map(want, qp.numpy(), gp.numpy(), qc.numpy(), gc.numpy())
    assert np.abs(cmc - c_o).max() <= METRIC_TOL and abs(mAP - m_o) <= METRIC_TOL


@pytest.mark.parametrize("max_rank", [1, 10, 100, 30000])
def test_eval_func_max_rank(api, synth, oracle, max_rank):
    """evaluate.py:29,37-39,68: general max_rank (the CMC is max_rank long; clipped to g with the note printed)."""
    evaluate, _ = api
    sh = synth.SHAPES["market"]
    qf, gf, qp, gp, qc, gc = synth.make(sh, 1)
    dist = oracle.sq_euclidean(qf[:300].numpy(), gf.numpy())
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf):
        cmc, mAP = evaluate.eval_func(dist, qp[:300], gp, qc[:300], gc, max_rank=max_rank)
    c_o, m_o = oracle.cmc_map(dist, qp[:300].numpy(), gp.numpy(), qc[:300].numpy(), gc.numpy(), max_rank=max_rank)
    assert cmc.shape == c_o.shape == (min(max_rank, sh.g),) and cmc.dtype == np.float32
    assert ("quite small" in buf.getvalue()) == (max_rank > sh.g)
    np.testing.assert_array_equal(cmc, c_o)          # same matrix on both sides: integer counts / n_valid
    assert abs(mAP - m_o) <= 1e-12


@pytest.mark.parametrize("d", [1519, 3072])
def test_wide_features_vs_float64(eng, engine_mod, d):
    """1 519-d (--mm) and 3 072-d (RN50) embeddings.  Squared distances of normalised rows against a
    float64 evaluation: within 1e-4 relative (north_star)."""
    gen = torch.Generator().manual_seed(d)
    X = torch.randn(3000, d, generator=gen).to(eng.device)
    f = eng.prepare(X, True)
    D, _ = eng.dist(f, f, engine_mod.METRIC_SQEUCLIDEAN)
    x64 = torch.nn.functional.normalize(X.double(), dim=1)
    ref = (x64 * x64).sum(1)[:, None] + (x64 * x64).sum(1)[None, :] - 2.0 * x64 @ x64.T
    got = D[:, :3000].double()
    off = ~torch.eye(3000, dtype=torch.bool, device=eng.device)
    rel = ((got - ref).abs() / ref.abs().clamp_min(1e-3))[off]
    print(f"[parity] d={d}: max relative error vs float64 {rel.max().item():.3e}")
    assert rel.max().item() <= 1e-4


def test_mid_vs_shipped_reference(api, synth):
    """The CUDA path against the UNMODIFIED reference (oracle/_ref: evaluate.py + reranking.py as shipped by
    oracle/build_ref.py), not the port: re_ranking matrix, eval_func on it, and R1_mAP_eval(reranking=True) with the
    reference's hard-coded 50/15/0.3."""
    build_ref = importlib.import_module("oracle.build_ref")
    if not build_ref.available():
        pytest.skip("oracle/_ref did not travel (built only where /root/reference is mounted)")
    evaluate, reranking = api
    ev, rr = build_ref.load()
    sh = synth.SHAPES["mid"]
    qf, gf, qp, gp, qc, gc = synth.make(sh, 3)
    raw = torch.cat([qf, gf])
    f = torch.nn.functional.normalize(raw, dim=1, p=2)
    with contextlib.redirect_stdout(io.StringIO()):
        want = np.ascontiguousarray(rr.re_ranking(f[: sh.q], f[sh.q:], sh.k1, sh.k2, sh.lam))
        c_r, m_r = ev.eval_func(want, qp.numpy(), gp.numpy(), qc.numpy(), gc.numpy())
        r = ev.R1_mAP_eval(sh.q, feat_norm=True, reranking=True)
        r.reset()
        r.update((raw, qp_gp := torch.cat([qp, gp]), torch.cat([qc, gc])))
        c_re, m_re = r.compute()
    got = reranking.re_ranking(f[: sh.q], f[sh.q:], sh.k1, sh.k2, sh.lam)
    assert_distances_close(got, want, what="mid vs reference")
    _list_report(got, want, 10, "mid vs reference")
    with contextlib.redirect_stdout(io.StringIO()):
        cmc, mAP = evaluate.eval_func(got, qp, gp, qc, gc)
        e = evaluate.R1_mAP_eval(sh.q, feat_norm=True, reranking=True)
        e.reset()
        e.update((raw, qp_gp, torch.cat([qc, gc])))
        cmc_e, map_e = e.compute()
    assert np.abs(cmc - c_r).max() <= METRIC_TOL and abs(mAP - m_r) <= METRIC_TOL
    assert np.abs(cmc_e - c_re).max() <= METRIC_TOL and abs(map_e - m_re) <= METRIC_TOL
