"""CPU, build container only: the oracle restatement against the UNMODIFIED reference imported
from /root/reference (tests/_refload.py), on fresh seeds beyond the frozen golden vectors.
Skipped where the reference tree is not mounted (the GPU box)."""
import contextlib
import io

import numpy as np
import pytest
import torch

import _refload

pytestmark = pytest.mark.skipif(not _refload.available(), reason="/root/reference not mounted")


def _case(synth, name, seed):
    sh = synth.SHAPES[name]
    qf, gf, qp, gp, qc, gc = synth.make(sh, seed)
    f = torch.nn.functional.normalize(torch.cat([qf, gf]), dim=1, p=2)
    return sh, f, (qp.numpy(), gp.numpy(), qc.numpy(), gc.numpy()), torch.cat([qf, gf])


@pytest.mark.parametrize("name,k1,k2,lam,seed", [("tiny", 20, 6, 0.3, 10), ("tiny", 50, 15, 0.3, 11),
                                                  ("tiny", 7, 2, 0.1, 12), ("tiny", 6, 1, 0.9, 13),
                                                  ("small", 20, 6, 0.3, 14)])
def test_rerank_bit_exact_through_only_local(oracle, synth, name, k1, k2, lam, seed):
    ev, rr = _refload.reference()
    sh, f, ids, _ = _case(synth, name, seed)
    D = oracle.sq_euclidean(f.numpy(), f.numpy())
    D = D + np.float32(1e-3) * np.random.default_rng(seed).random(D.shape, dtype=np.float32)   # asymmetric on purpose
    with contextlib.redirect_stdout(io.StringIO()):
        want = np.ascontiguousarray(rr.re_ranking(f[: sh.q], f[sh.q:], k1, k2, lam, local_distmat=D.copy(),
                                                  only_local=True))
    got = oracle.rerank_from_dist(D, sh.q, k1, k2, lam)
    np.testing.assert_array_equal(got.view(np.uint32), want.view(np.uint32))


@pytest.mark.parametrize("seed", [20, 21])
def test_eval_func_and_distances(oracle, synth, seed):
    ev, rr = _refload.reference()
    sh, f, (qp, gp, qc, gc), raw = _case(synth, "small", seed)
    q = sh.q
    with contextlib.redirect_stdout(io.StringIO()):
        eu = ev.euclidean_distance(f[:q], f[q:])
        co = ev.cosine_similarity(raw[:q], raw[q:])
        cmc, mAP = ev.eval_func(eu, qp, gp, qc, gc)
    np.testing.assert_allclose(oracle.sq_euclidean(f[:q].numpy(), f[q:].numpy()), eu, rtol=1e-4, atol=2e-6)
    np.testing.assert_allclose(oracle.arccos_distance(raw[:q].numpy(), raw[q:].numpy()), co, rtol=1e-4, atol=1e-5)
    c2, m2 = oracle.cmc_map(eu, qp, gp, qc, gc)
    np.testing.assert_array_equal(c2, cmc)
    assert abs(m2 - mAP) < 1e-12


def test_evaluator_quirks(oracle, synth):
    """compute() ignores the constructor's max_rank (evaluate.py:132) and hard-codes 50/15/0.3."""
    ev, rr = _refload.reference()
    sh, f, (qp, gp, qc, gc), raw = _case(synth, "tiny", 30)
    pids, cams = np.concatenate([qp, gp]), np.concatenate([qc, gc])
    with contextlib.redirect_stdout(io.StringIO()):
        e = ev.R1_mAP_eval(sh.q, max_rank=10, feat_norm=True, reranking=True)
        e.reset()
        e.update((raw, torch.from_numpy(pids), torch.from_numpy(cams)))
        cmc, mAP = e.compute()
    assert cmc.shape == (50,)
    c2, m2 = oracle.evaluate_features(raw.numpy(), pids, cams, sh.q, True, True)
    np.testing.assert_allclose(c2, cmc, atol=1e-5)
    assert abs(m2 - mAP) < 1e-5
