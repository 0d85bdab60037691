"""CPU: the oracle's numpy-primitive restatements vs numpy itself, and the oracle vs the golden
vectors frozen from the unmodified reference (oracle/make_golden.py)."""
import glob
import os

import numpy as np
import pytest

GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "*.npz")))


def test_golden_present():
    assert len(GOLDEN) >= 5


def test_exp_restatement_matches_numpy(oracle):
    rng = np.random.default_rng(0)
    x = rng.uniform(-1.03, 0.02, 400_000).astype(np.float32)
    np.testing.assert_array_equal(oracle.exp_f32_restated(x).view(np.uint32), np.exp(x).view(np.uint32))
    x = -rng.uniform(0, 1, 100_000).astype(np.float32) ** 3
    np.testing.assert_array_equal(oracle.exp_f32_restated(x).view(np.uint32), np.exp(x).view(np.uint32))


def test_pairwise_sum_restatement_matches_numpy(oracle):
    rng = np.random.default_rng(1)
    for n in list(range(1, 300)) + [511, 512, 513, 1000, 1377, 4099]:
        a = rng.uniform(0.2, 1.0, n).astype(np.float32)
        assert oracle.pairwise_sum_f32_restated(a).view(np.uint32) == np.sum(a).view(np.uint32), n


def test_half_k_is_bankers_rounding(oracle):
    assert [oracle.half_k(k) for k in (20, 50, 25, 5, 1, 2, 3, 6, 7)] == [11, 26, 13, 3, 1, 2, 3, 4, 5]


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[:-4] for p in GOLDEN])
def test_oracle_matches_golden(oracle, path):
    z = np.load(path)
    q, k1, k2, lam = int(z["q"]), int(z["k1"]), int(z["k2"]), float(z["lam"])
    f, pids, cams = z["feats"], z["pids"], z["cams"]
    if z["D"].size:
        # same N x N matrix as the reference used -> every later stage must agree bit for bit
        final = oracle.rerank_from_dist(z["D"], q, k1, k2, lam)
        np.testing.assert_array_equal(final.view(np.uint32), z["final"].view(np.uint32))
    # own contraction (OpenBLAS here, MKL in the reference): tolerance + fp16 flips
    from parity import assert_distances_close
    final2 = oracle.re_ranking(f[:q], f[q:], k1, k2, lam)
    assert_distances_close(final2, z["final"], what="oracle re_ranking")
    rows = z["eu"].shape[0]
    np.testing.assert_allclose(oracle.sq_euclidean(f[:q], f[q:])[:rows], z["eu"], rtol=1e-4, atol=2e-6)
    raw = z["raw"]
    np.testing.assert_allclose(oracle.arccos_distance(raw[:q], raw[q:])[:rows], z["co"], rtol=1e-4, atol=1e-5)
    cmc, mAP = oracle.cmc_map(z["final"], pids[:q], pids[q:], cams[:q], cams[q:])
    np.testing.assert_array_equal(cmc, z["cmc_rr"])
    assert abs(mAP - float(z["map_rr"])) < 1e-12
    cmc, mAP = oracle.evaluate_features(raw, pids, cams, q, feat_norm=True, reranking=False)
    np.testing.assert_allclose(cmc, z["cmc_eval"], atol=1e-5)
    assert abs(mAP - float(z["map_eval"])) < 1e-5
    cmc, mAP = oracle.evaluate_features(raw, pids, cams, q, feat_norm=True, reranking=True)
    np.testing.assert_allclose(cmc, z["cmc_eval_rr"], atol=1e-5)
    assert abs(mAP - float(z["map_eval_rr"])) < 1e-5


def test_cmc_map_edge_cases(oracle):
    # query 1 has only a same-camera match (junk) -> skipped; query 0 matches at rank 2 of the kept list
    dist = np.array([[0.1, 0.2, 0.3, 0.4], [0.4, 0.3, 0.2, 0.1]], np.float32)
    q_pids, q_cams = np.array([7, 9]), np.array([0, 1])
    g_pids, g_cams = np.array([7, 5, 7, 9]), np.array([0, 0, 1, 1])
    cmc, mAP, aps, first = oracle.cmc_map(dist, q_pids, g_pids, q_cams, g_cams, max_rank=50, detail=True)
    assert cmc.shape == (4,)                      # max_rank clipped to g (evaluate.py:37-39)
    np.testing.assert_array_equal(cmc, np.array([0, 1, 1], np.float32)[:3].tolist() + [1.0])
    assert first.tolist() == [1, -1] and np.isnan(aps[1]) and abs(mAP - 0.5) < 1e-15
    with pytest.raises(AssertionError):
        oracle.cmc_map(dist[1:], q_pids[1:], g_pids, q_cams[1:], g_cams)
