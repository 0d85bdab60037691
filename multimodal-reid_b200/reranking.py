"""Drop-in for the reference's ``reranking.py`` (k-reciprocal re-ranking, Zhong et al. CVPR'17).

Same call surface and semantics as ``re_ranking`` (reranking.py:29-100) — k1 / k2 / lambda_value,
``local_distmat`` added to the N x N matrix (:43-44) or used verbatim with ``only_local=True``
(:33-34), float16 rounding chain, float32 q x g result — executed by the sm_100a kernels in
libreid_b200.so.  No CPU path: without a B200 the call raises."""
from __future__ import annotations

import numpy as np
import torch

from . import engine as _engine


def _to_device_f32(x, device) -> torch.Tensor:
    if isinstance(x, np.ndarray):
        x = torch.from_numpy(x)
    if x.dtype != torch.float32:
        x = x.float()          # fp16 embeddings are up-cast once; the reference would compute in fp16
    return x.to(device, non_blocking=True).contiguous()


_single = {}


def _single_rank_engine(eng: _engine.Engine) -> _engine.Engine:
    key = str(eng.device)
    if key not in _single:
        _single[key] = _engine.Engine(eng.device, comm=_engine.Comm(single=True))
    return _single[key]


def re_ranking_device(probFea, galFea, k1, k2, lambda_value, local_distmat=None, only_local=False,
                      eng: _engine.Engine | None = None) -> torch.Tensor:
    """Same as :func:`re_ranking` but returns the device tensor [q_local, g] (float32).  On several ranks the rows are
    this rank's queries: `Engine.own_queries(N, q, k1, k2)` names them (snake blocks on the peer-memory path, rank +
    l * world on the NCCL path and for a caller-supplied local_distmat)."""
    eng = eng or _engine.default_engine()
    q = int(probFea.shape[0])
    g = int(galFea.shape[0])
    N = q + g
    dev = eng.device
    cm = eng.comm
    if only_local or local_distmat is not None:
        L = _to_device_f32(local_distmat, dev)
        if L.shape != (N, N):
            raise ValueError(f"local_distmat must be {N}x{N}, got {tuple(L.shape)}")
        if cm.world != 1:
            # The caller's N x N matrix is replicated on every rank already, and nothing about it is sharded: every rank
            # runs the one-rank path on it ("replicas") and keeps the rows of its own queries (rank, rank + world, ...),
            # the same rows re_ranking returns on the NCCL path.
            eng = _single_rank_engine(eng)
    if only_local:
        # the reference reads column i of the matrix for probe i (reranking.py:46)
        D = eng.transposed(L)
        key = eng.rowmax(D, N)
        final = eng.rerank_from_panel(D, key, q, N, k1, k2, lambda_value)
    else:
        feats = torch.cat([_to_device_f32(probFea, dev), _to_device_f32(galFea, dev)])
        if local_distmat is None:
            final = eng.rerank_features(feats, q, k1, k2, lambda_value, normalize=False)
        else:
            allf = eng.prepare(feats, False)
            D, _ = eng.dist(allf, allf, _engine.METRIC_SQEUCLIDEAN, want_rowmax=False)
            eng.transposed(L, D, accumulate=True)
            key = eng.rowmax(D, N)
            final = eng.rerank_from_panel(D, key, q, N, k1, k2, lambda_value)
    if (only_local or local_distmat is not None) and cm.world != 1:
        final = final[cm.rank:: cm.world]
    return final[:, :g]


def re_ranking(probFea, galFea, k1, k2, lambda_value, local_distmat=None, only_local=False):
    """reranking.py:29 — returns the re-ranked q x g distance matrix as a float32 numpy array."""
    out = re_ranking_device(probFea, galFea, k1, k2, lambda_value, local_distmat, only_local)
    return out.cpu().numpy()
