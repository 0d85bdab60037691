"""CPU oracle for the training-side distance ops.  TEST INFRASTRUCTURE ONLY.

NumPy restatement of ``euclidean_dist`` (losses.py:14-28), ``hard_example_mining`` (losses.py:31-83), their
gradients as torch autograd derives them, and the triplet loss on top (losses.py:86-108) of
SuperbTUM/Multimodal-ReID @ 8dc4505d.  Only ``tests/`` may import it.  Parity pin: the reference ships no
tests; ``oracle/make_golden_losses.py`` freezes outputs AND autograd gradients of the unmodified reference
(imported from /root/reference in the build container) into ``tests/golden/losses/*.npz``, and
``tests/test_oracle.py`` checks this file against them.
"""
from __future__ import annotations

import numpy as np


def euclidean_dist(x: np.ndarray, y: np.ndarray) -> np.ndarray:
    """losses.py:22-27: xx + yy, minus 2 x y^T, clamp(min=1e-12), sqrt — all float32."""
    x = x.astype(np.float32)
    y = y.astype(np.float32)
    xx = (x * x).sum(1, dtype=np.float32)[:, None]
    yy = (y * y).sum(1, dtype=np.float32)[None, :]
    v = (xx + yy) - np.float32(2) * (x @ y.T)
    return np.sqrt(np.maximum(v, np.float32(1e-12))).astype(np.float32)


def hard_example_mining(dist: np.ndarray, lab_r: np.ndarray, lab_c: np.ndarray | None = None, skip_diag=False):
    """losses.py:51-79: per row, max over same-label columns / min over other-label columns, first index on ties
    (torch.max / torch.min on CPU keep the first extremum); indices are absolute columns (:70-77).
    lab_c / skip_diag restate the masks of losses.py:126-128 (XBM variant)."""
    lab_c = lab_r if lab_c is None else lab_c
    m, n = dist.shape
    pos = lab_r[:, None] == lab_c[None, :]
    neg = ~pos
    if skip_diag:
        pos = pos.copy()
        pos[np.arange(m), np.arange(m)] = False
    ap = np.empty(m, np.float32)
    an = np.empty(m, np.float32)
    pi = np.empty(m, np.int64)
    ni = np.empty(m, np.int64)
    for i in range(m):
        pc = np.flatnonzero(pos[i])
        nc = np.flatnonzero(neg[i])
        if len(pc):
            j = int(np.argmax(dist[i, pc]))
            ap[i], pi[i] = dist[i, pc[j]], pc[j]
        else:
            ap[i], pi[i] = -np.inf, -1
        if len(nc):
            j = int(np.argmin(dist[i, nc]))
            an[i], ni[i] = dist[i, nc[j]], nc[j]
        else:
            an[i], ni[i] = np.inf, -1
    return ap, an, pi, ni


def euclidean_dist_grad(x: np.ndarray, y: np.ndarray, dist: np.ndarray, G: np.ndarray):
    """What autograd returns for losses.py:22-27 given dL/d dist = G: sqrt gives G / (2 dist), the clamp passes it
    where the squared form was >= 1e-12, and the squared form has d/dx_i = 2 (x_i - y_j).  float64 accumulation."""
    x64, y64 = x.astype(np.float64), y.astype(np.float64)
    W = np.where(dist > np.float32(1e-6), G.astype(np.float64) / dist.astype(np.float64), 0.0)
    gx = W.sum(1)[:, None] * x64 - W @ y64
    gy = W.sum(0)[:, None] * y64 - W.T @ x64
    return gx, gy


def triplet(x: np.ndarray, labels: np.ndarray, margin):
    """losses.py:96-108: loss value and d loss / d x (x used as both operands)."""
    dist = euclidean_dist(x, x)
    ap, an, pi, ni = hard_example_mining(dist, labels)
    m = len(ap)
    ap64, an64 = ap.astype(np.float64), an.astype(np.float64)
    if margin is not None:          # MarginRankingLoss(an, ap, 1) = mean(max(0, ap - an + margin))
        t = ap64 - an64 + margin
        loss = np.maximum(t, 0).mean()
        g_ap = (t > 0) / m
        g_an = -g_ap
    else:                           # SoftMarginLoss(an - ap, 1) = mean(log(1 + exp(ap - an)))
        t = ap64 - an64
        loss = np.log1p(np.exp(t)).mean()
        g_ap = 1.0 / (1.0 + np.exp(-t)) / m
        g_an = -g_ap
    G = np.zeros(dist.shape, np.float64)
    np.add.at(G, (np.arange(m), pi), g_ap)
    np.add.at(G, (np.arange(m), ni), g_an)
    gx, gy = euclidean_dist_grad(x, x, dist, G)
    return loss, gx + gy, (dist, ap, an, pi, ni)
