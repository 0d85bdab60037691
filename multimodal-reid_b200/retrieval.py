"""Streamed top-k retrieval (first slice; BASELINE.json config 5: 100k query x 1M gallery x 512-d).

No function of the reference corresponds to this: its ``euclidean_distance`` + ``np.argsort`` (evaluate.py:7-13, :40)
would need the 400 GB q x g matrix.  The semantics are those two calls restricted to the first ``topk`` columns of every
row: squared L2 (or the arccos metric of evaluate.py:16-26), ascending, ties by gallery index.  With several ranks the
queries are sharded cyclically and the lists all-gathered over NVLink.  No CPU path."""
from __future__ import annotations

import numpy as np
import torch

from . import engine as _engine
from .evaluate import _ids
from .reranking import _to_device_f32


def topk_retrieval_device(qf, gf, topk: int = 100, metric: int = _engine.METRIC_SQEUCLIDEAN, feat_norm: bool = False,
                          eng: _engine.Engine | None = None, rows_per_pass: int = 16384, cols_per_pass: int = 65536):
    """Returns (indices int32 [q, topk], distances float32 [q, topk]) on the device, identical on every rank."""
    eng = eng or _engine.default_engine()
    cm = eng.comm
    q = int(qf.shape[0])
    # one prepare call: queries and gallery share the power-of-two operand scale, like one call of the evaluator
    allf = eng.prepare(torch.cat([_to_device_f32(qf, eng.device), _to_device_f32(gf, eng.device)]), feat_norm)
    Q, G = allf.rows(slice(cm.rank, q, cm.world)), allf.rows(slice(q, allf.n))
    idx, val = eng.retrieve_topk(Q, G, topk, metric, rows_per_pass, cols_per_pass)
    if cm.world > 1:
        with eng.span("x_topk"):
            idx = cm.gather_rows(idx, q)
            val = cm.gather_rows(val, q)
    return idx, val


def topk_retrieval(qf, gf, topk: int = 100, metric: str = "euclidean", feat_norm: bool = False):
    """numpy front end: (indices int32 [q, topk], distances float32 [q, topk])."""
    m = {"euclidean": _engine.METRIC_SQEUCLIDEAN, "cosine": _engine.METRIC_ARCCOS}[metric]
    idx, val = topk_retrieval_device(qf, gf, topk, m, feat_norm)
    return idx.cpu().numpy(), val.cpu().numpy()


def streamed_eval(qf, gf, q_pids, g_pids, q_camids, g_camids, max_rank: int = 50, topk: int = 0,
                  feat_norm: bool = True, rows_per_pass: int = 16384, cols_per_pass: int = 65536):
    """``R1_mAP_eval(reranking=False)`` semantics (evaluate.py:110-135: optional L2 normalisation, squared Euclidean
    distances, ``eval_func``) for a query / gallery pair whose distance matrix does not fit: returns
    ``(cmc float32[max_rank], mAP float64)`` and, with ``topk > 0``, also the exact top-k lists
    ``(indices int32 [q, topk], distances float32 [q, topk])`` computed on the same panels."""
    eng = _engine.default_engine()
    cm = eng.comm
    q, g = int(qf.shape[0]), int(gf.shape[0])
    if g < max_rank:
        max_rank = g
        print("Note: number of gallery samples is quite small, got {}".format(g))
    allf = eng.prepare(torch.cat([_to_device_f32(qf, eng.device), _to_device_f32(gf, eng.device)]), feat_norm)
    mine = slice(cm.rank, q, cm.world)
    Q, G = allf.rows(mine), allf.rows(slice(q, q + g))
    qp, gp = _ids(q_pids, eng.device), _ids(g_pids, eng.device)
    qc, gc = _ids(q_camids, eng.device), _ids(g_camids, eng.device)
    (cmc, mAP, nvalid, _, _), top = eng.evaluate_streamed(Q, G, qp[mine].contiguous(), gp, qc[mine].contiguous(), gc,
                                                          max_rank, topk, _engine.METRIC_SQEUCLIDEAN, rows_per_pass,
                                                          cols_per_pass, q_total=q)
    assert int(nvalid.item()) > 0, "Error: all query identities do not appear in gallery"
    out = (cmc.cpu().numpy(), np.float64(mAP.item()))
    if top is None:
        return out
    idx, val = top
    if cm.world > 1:
        idx, val = cm.gather_rows(idx, q), cm.gather_rows(val, q)
    return out + (idx.cpu().numpy(), val.cpu().numpy())


def streamed_rerank_eval(qf, gf, q_pids, g_pids, q_camids, g_camids, k1: int = 20, k2: int = 6, lambda_value: float = 0.3,
                         topk: int = 100, max_rank: int = 50, feat_norm: bool = True, rows_per_pass: int = 16384,
                         cols_per_pass: int = 65536, wide: int = 64, cap_o: int = 32768, return_diag: bool = False):
    """``R1_mAP_eval(reranking=True)`` semantics (evaluate.py:110-135 with re_ranking's k1 / k2 / lambda_value,
    reranking.py:29-100) for a query / gallery pair whose N x N matrix cannot exist (BASELINE config 5: 100k x 1M):
    returns ``(cmc float32[max_rank], mAP float64, indices int32 [q, topk], final float32 [q, topk])`` -- CMC / mAP of the
    k-reciprocal re-ranked distances and the exact top-k of every re-ranked row.  With several ranks the rows of the
    N x N sweep and the queries are sharded cyclically; neighbour lists and results are all-gathered."""
    eng = _engine.default_engine()
    cm = eng.comm
    q, g = int(qf.shape[0]), int(gf.shape[0])
    if g < max_rank:
        max_rank = g
        print("Note: number of gallery samples is quite small, got {}".format(g))
    allf = eng.prepare(torch.cat([_to_device_f32(qf, eng.device), _to_device_f32(gf, eng.device)]), feat_norm)
    pids = torch.cat([_ids(q_pids, eng.device), _ids(g_pids, eng.device)])
    cams = torch.cat([_ids(q_camids, eng.device), _ids(g_camids, eng.device)])
    (cmc, mAP, nvalid, _, _), (idx, val), diag = eng.evaluate_streamed_reranked(
        allf, q, pids, cams, k1, k2, lambda_value, topk=topk, max_rank=max_rank, Kw=wide, cap_o=cap_o,
        rows_per_pass=rows_per_pass, cols_per_pass=cols_per_pass)
    assert int(nvalid.item()) > 0, "Error: all query identities do not appear in gallery"
    if cm.world > 1:
        idx, val = cm.gather_rows(idx.contiguous(), q), cm.gather_rows(val.contiguous(), q)
    out = (cmc.cpu().numpy(), np.float64(mAP.item()), idx.cpu().numpy(), val.cpu().numpy())
    return out + (diag,) if return_diag else out
