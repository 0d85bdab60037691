"""CPU: the counting formulation of CMC / mAP (oracle/stream_oracle.py = stream_score.cu's steps in numpy) equals the
ranking formulation (oracle.cmc_map, pinned to the reference's eval_func) — ties, junk, absent identities, any blocking."""
import numpy as np
import pytest

from oracle import stream_oracle as so


def _case(seed, q, g, n_id, n_cam, quant=None, absent=0):
    rng = np.random.default_rng(seed)
    dist = rng.random((q, g), dtype=np.float32) * 2
    if quant:
        dist = (np.round(dist * quant) / quant).astype(np.float32)          # many exact ties
    gp = rng.integers(0, n_id, g)
    qp = rng.integers(0, n_id, q)
    qp[:absent] = n_id + 3
    return dist, qp, gp, rng.integers(0, n_cam, q), rng.integers(0, n_cam, g)


@pytest.mark.parametrize("seed,q,g,n_id,n_cam,quant,blocks", [
    (0, 40, 500, 20, 4, None, None),
    (1, 40, 500, 20, 2, 16, [(0, 100), (100, 333), (333, 500)]),      # heavy ties, ragged blocks, half junk
    (2, 30, 400, 3, 3, 64, [(0, 7), (7, 400)]),                        # groups of ~130 columns
    (3, 25, 300, 60, 1, None, [(i, i + 50) for i in range(0, 300, 50)]),   # one camera: every same-id entry is junk
])
@pytest.mark.parametrize("use_cells", [True, False])
def test_counting_equals_ranking(oracle, seed, q, g, n_id, n_cam, quant, blocks, use_cells):
    dist, qp, gp, qc, gc = _case(seed, q, g, n_id, n_cam, quant, absent=3)
    if n_cam == 1:
        qc = qc + 1                      # otherwise no query has a match left (the reference asserts)
    cmc, mAP, ap, first = so.cmc_map_counting(dist, qp, gp, qc, gc, 50, blocks, use_cells)
    c2, m2, ap2, first2 = oracle.cmc_map(dist, qp, gp, qc, gc, 50, detail=True)
    np.testing.assert_array_equal(first, first2)
    np.testing.assert_array_equal(cmc, c2)
    v = ~np.isnan(ap2)
    np.testing.assert_allclose(ap[v], ap2[v], rtol=0, atol=1e-15)
    assert (ap[~v] == -1).all() and abs(mAP - m2) < 1e-15


def test_zero_and_negative_distances(oracle):
    dist, qp, gp, qc, gc = _case(7, 20, 200, 10, 3, quant=8)
    dist -= np.float32(0.25)             # slightly negative entries, exact zeros, -0.0
    dist[dist == 0] = np.float32(-0.0)
    cmc, mAP, ap, first = so.cmc_map_counting(dist, qp, gp, qc, gc, 50, [(0, 64), (64, 200)])
    c2, m2, ap2, first2 = oracle.cmc_map(dist, qp, gp, qc, gc, 50, detail=True)
    np.testing.assert_array_equal(first, first2)
    np.testing.assert_array_equal(cmc, c2)
    assert abs(mAP - m2) < 1e-15


@pytest.mark.parametrize("quant", [None, 8])
def test_blockwise_topk_equals_global_stable_topk(quant):
    """candidate position order == global column order among equal distances (Engine.retrieve_topk's claim)"""
    dist, *_ = _case(11, 30, 700, 5, 2, quant)
    K = 50
    want = np.argsort(dist, axis=1, kind="stable")[:, :K]
    for blocks in ([(0, 700)], [(0, 256), (256, 512), (512, 700)], [(i, min(700, i + 64)) for i in range(0, 700, 64)]):
        idx, val = so.topk_by_blocks(dist, K, blocks)
        np.testing.assert_array_equal(idx, want)
        np.testing.assert_array_equal(val, np.take_along_axis(dist, want, 1))
