"""Drop-in for the reference's ``evaluate.py``: same names, arguments, prints, return types and
quirks (squared distances; ``compute()`` ignores the constructor's ``max_rank`` and hard-codes
k1=50, k2=15, lambda=0.3 when ``reranking=True``; AssertionError when no query has a match) with
every array operation executed on a B200 by libreid_b200.so.  No CPU path."""
from __future__ import annotations

import numpy as np
import torch

from . import engine as _engine
from .reranking import re_ranking, re_ranking_device, _to_device_f32


def _ids(x, device) -> torch.Tensor:
    if isinstance(x, torch.Tensor):
        t = x
    else:
        t = torch.from_numpy(np.ascontiguousarray(np.asarray(x)))
    return t.to(device=device, dtype=torch.int64, non_blocking=True).contiguous()


def _dist_matrix(qf, gf, metric) -> np.ndarray:
    eng = _engine.default_engine()
    out_dtype = qf.dtype if isinstance(qf, torch.Tensor) else torch.float32
    A = eng.prepare(_to_device_f32(qf, eng.device), False)
    B = eng.prepare(_to_device_f32(gf, eng.device), False)
    D, _ = eng.dist(A, B, metric)
    D = D[:, : B.n]
    if metric == _engine.METRIC_SQEUCLIDEAN and out_dtype == torch.float16:
        D = D.half()           # evaluate.py:7-13 returns the input dtype
    return D.cpu().numpy()


def euclidean_distance(qf, gf):
    """evaluate.py:7-13 — squared L2, float32 numpy [m, n]."""
    return _dist_matrix(qf, gf, _engine.METRIC_SQEUCLIDEAN)


def cosine_similarity(qf, gf):
    """evaluate.py:16-26 — arccos of the clipped cosine, float32 numpy [m, n]."""
    return _dist_matrix(qf, gf, _engine.METRIC_ARCCOS)


def eval_func_device(dist: torch.Tensor, g: int, q_pids, g_pids, q_camids, g_camids, max_rank=50,
                     eng: _engine.Engine | None = None, q_total: int | None = None):
    """dist: device tensor [q_local, >= g] float32 (row pitch free).  Returns (cmc, mAP) like
    eval_func; only cmc[max_rank] and mAP leave the device."""
    eng = eng or _engine.default_engine()
    if dist.dtype != torch.float32 or dist.stride(1) != 1 or dist.stride(0) % 4 != 0:
        buf = torch.empty((dist.shape[0], _engine.Engine.pitch(g)), dtype=torch.float32, device=dist.device)
        buf[:, :g].copy_(dist[:, :g])
        dist = buf
    if g < max_rank:
        max_rank = g
        print("Note: number of gallery samples is quite small, got {}".format(g))
    cmc, mAP, nvalid, _, _ = eng.evaluate_dist(dist, g, q_pids, g_pids, q_camids, g_camids, max_rank, q_total)
    nv = int(nvalid.item())
    assert nv > 0, "Error: all query identities do not appear in gallery"
    return cmc.cpu().numpy(), np.float64(mAP.item())


def eval_func(distmat, q_pids, g_pids, q_camids, g_camids, max_rank=50):
    """evaluate.py:29-88 — Market-1501 CMC / mAP for a q x g distance matrix."""
    eng = _engine.default_engine()
    dev = eng.device
    if isinstance(distmat, np.ndarray):
        d = torch.from_numpy(np.ascontiguousarray(distmat))
    else:
        d = distmat
    num_q, num_g = d.shape
    cm = eng.comm
    mine = slice(cm.rank, num_q, cm.world)          # one process per GPU: every rank scores its share of the queries
    if cm.world > 1:
        d = d[mine]
    # the scoring kernel streams rows with 16-byte loads: rows are placed at a pitch that is a multiple of 4 floats (one
    # copy either way -- the matrix has to reach the device)
    buf = torch.empty((d.shape[0], _engine.Engine.pitch(num_g)), dtype=torch.float32, device=dev)
    buf[:, :num_g].copy_(d, non_blocking=True)
    d = buf
    return eval_func_device(d, num_g, _ids(q_pids, dev)[mine].contiguous(), _ids(g_pids, dev),
                            _ids(q_camids, dev)[mine].contiguous(), _ids(g_camids, dev), max_rank, eng, q_total=num_q)


class _IdStore:
    """The evaluator's ``pids`` / ``camids`` (evaluate.py:101-102, 107-108).  The reference extends a Python list with
    one numpy scalar per sample and turns it back into an array in compute() -- milliseconds of interpreter time per
    evaluation at Duke scale, more than the whole GPU pipeline.  Batches are kept as they arrive instead; the list is
    only built when somebody reads the attribute, and from then on it is the source of truth (it may have been edited)."""

    def __init__(self, initial=None):
        self.chunks = []
        self.listed = None
        if initial is not None and len(initial):
            self.listed = initial if isinstance(initial, list) else list(initial)

    @staticmethod
    def _as_numpy(x):
        return np.asarray(x.cpu() if isinstance(x, torch.Tensor) else x)

    def extend(self, x):
        if self.listed is not None:
            self.listed.extend(self._as_numpy(x))
        else:
            self.chunks.append(x.detach() if isinstance(x, torch.Tensor) else np.asarray(x))

    def as_list(self):
        if self.listed is None:
            self.listed = []
            for c in self.chunks:
                self.listed.extend(self._as_numpy(c))
            self.chunks = []
        return self.listed

    def to_device(self, device) -> torch.Tensor:
        if self.listed is not None:
            return _ids(np.asarray(self.listed), device)
        parts = [_ids(c, device).reshape(-1) for c in self.chunks]
        if not parts:
            return torch.empty((0,), dtype=torch.int64, device=device)
        return parts[0] if len(parts) == 1 else torch.cat(parts)


class R1_mAP_eval():
    """evaluate.py:91-135.  ``update`` keeps tensors where they are (CPU or CUDA); ``compute``
    moves them to the B200 once and returns (cmc float32[50], mAP float64)."""

    def __init__(self, num_query, max_rank=50, feat_norm=True, reranking=False, rerank_params=(50, 15, 0.3)):
        super(R1_mAP_eval, self).__init__()
        self.num_query = num_query
        self.max_rank = max_rank
        self.feat_norm = feat_norm
        self.reranking = reranking
        # extension (keyword, default = the values the reference hard-codes at evaluate.py:127)
        self.rerank_params = tuple(rerank_params)

    def reset(self):
        self.feats = []
        self.feats_aug = []          # extension: second (flipped) embedding pass, see update()
        self.pids = []
        self.camids = []

    # public list attributes like the reference's; backed by _IdStore so that update()/compute() stay off the
    # interpreter's per-element path unless a caller actually looks at (or assigns) the lists
    @property
    def pids(self):
        return self._pid_store.as_list()

    @pids.setter
    def pids(self, value):
        self._pid_store = _IdStore(value)

    @property
    def camids(self):
        return self._cam_store.as_list()

    @camids.setter
    def camids(self, value):
        self._cam_store = _IdStore(value)

    def update(self, output, feat_aug=None):  # called once for each batch
        """evaluate.py:104-108.  Extension (keyword, default = the reference's behaviour): `feat_aug`, the embeddings
        of the horizontally flipped batch.  The reference's scripts average the two passes themselves before calling
        the evaluator (prompt_learning.py:775-776: `(embeddings + embeddings_augmented) / 2`); given both, compute()
        forms the average inside the prologue kernel instead (same float32 arithmetic, one pass less over N x d)."""
        feat, pid, camid = output
        self.feats.append(feat.detach())
        if feat_aug is not None:
            assert tuple(feat_aug.shape) == tuple(feat.shape), "feat_aug must match feat"
            self.feats_aug.append(feat_aug.detach())
        self._pid_store.extend(pid)
        self._cam_store.extend(camid)

    def compute(self):  # called after each epoch
        eng = _engine.default_engine()
        dev = eng.device
        # Features still in host memory + re-ranking on one rank: the copy is streamed under the contraction
        # (engine.prepare_dist_streamed) instead of being finished first.  Same bits either way.
        tta = len(getattr(self, "feats_aug", [])) > 0
        if tta:
            assert len(self.feats_aug) == len(self.feats), "feat_aug must be given for every batch or for none"
        streamed = (not tta and self.reranking and self.feat_norm and eng.stream_h2d and eng.comm.world == 1
                    and eng.engine != _engine.ENGINES["simt"] and eng.symmetric and eng.upper_only
                    and len(self.feats) > 0
                    and all(isinstance(f, torch.Tensor) and not f.is_cuda and f.dtype == torch.float32 and f.dim() == 2
                            and f.shape[1] == self.feats[0].shape[1] for f in self.feats))
        host_f32 = (len(self.feats) > 0
                    and all(isinstance(f, torch.Tensor) and not f.is_cuda and f.dtype == torch.float32 and f.dim() == 2
                            and f.shape[1] == self.feats[0].shape[1] for f in self.feats))
        if streamed:
            feats = None
            N = sum(int(f.shape[0]) for f in self.feats)
        elif host_f32 and eng.comm.world > 1 and eng.shard_h2d and not tta:
            # several ranks, features in host memory: every rank uploads 1/world of the rows, NCCL all-gather does the rest
            feats = eng.comm.upload_sharded(self.feats, dev)
            N = feats.shape[0]
        else:
            feats = torch.cat([_to_device_f32(f, dev) for f in self.feats], dim=0)
            N = feats.shape[0]
        feats_aug = torch.cat([_to_device_f32(f, dev) for f in self.feats_aug], dim=0) if tta else None
        if self.feat_norm:
            print("The test feature is normalized")
        q = self.num_query
        g = N - q
        pids = self._pid_store.to_device(dev)
        camids = self._cam_store.to_device(dev)
        if self.reranking:
            print('=> Enter reranking')
            k1, k2, lam = self.rerank_params                                          # evaluate.py:127
            # evaluate.py:132 does not forward self.max_rank: always 50
            max_rank = 50
            if g < max_rank:
                max_rank = g
                print("Note: number of gallery samples is quite small, got {}".format(g))
            if streamed:
                cmc, mAP, nvalid, _, _ = eng.evaluate_reranked_host(self.feats, q, pids, camids, k1, k2, lam, max_rank)
            else:
                cmc, mAP, nvalid, _, _ = eng.evaluate_reranked(feats, q, pids, camids, k1, k2, lam, self.feat_norm,
                                                               max_rank, feats_aug=feats_aug)
            assert int(nvalid.item()) > 0, "Error: all query identities do not appear in gallery"
            cmc, mAP = cmc.cpu().numpy(), np.float64(mAP.item())
            print("Rank@{:d}:{:.1%}, Rank@{:d}:{:.1%}, Rank@{:d}:{:.1%}, mAP:{:.1%}".format(1, cmc[0], 5, cmc[4], 10, cmc[9], mAP))
            return cmc, mAP
        else:
            print('=> Computing DistMat with euclidean_distance')
            allf = eng.prepare(feats, self.feat_norm, feats_aug)
            dist, _ = eng.dist(allf.rows(slice(eng.comm.rank, q, eng.comm.world)), allf.rows(slice(q, N)),
                               _engine.METRIC_SQEUCLIDEAN)
        # evaluate.py:132 does not forward self.max_rank: always 50
        cm = eng.comm
        mine = slice(cm.rank, q, cm.world)          # `dist` holds this rank's queries (rank, rank + world, ...)
        cmc, mAP = eval_func_device(dist, g, pids[:q][mine].contiguous(), pids[q:].contiguous(),
                                    camids[:q][mine].contiguous(), camids[q:].contiguous(), 50, eng, q_total=q)
        print("Rank@{:d}:{:.1%}, Rank@{:d}:{:.1%}, Rank@{:d}:{:.1%}, mAP:{:.1%}".format(1, cmc[0], 5, cmc[4], 10, cmc[9], mAP))
        return cmc, mAP
