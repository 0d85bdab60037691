"""CPU restatement of the COUNTING formulation of CMC / mAP that stream_score.cu implements.  TEST INFRASTRUCTURE ONLY.

evaluate.py:40-88 ranks every gallery row and reads the matches' positions off the ranking; the streamed scorer never
ranks: per query it sorts the matches' (distance, column) keys, bins every entry of every column block by the number of
match keys below it (cells of a distance histogram first, exact search only in cells that hold a match), and turns prefix
sums minus the junk in front into kept ranks.  This file restates those steps one for one in numpy so that the identity
    counting over column blocks  ==  oracle.reid_oracle.cmc_map (stable ranking; itself pinned to the reference)
is checked on the CPU (tests/test_stream_oracle.py), ties, junk and absent identities included."""
from __future__ import annotations

import numpy as np

NB = 2048      # SG_NB of stream_score.cu


def _f2key(d: np.ndarray) -> np.ndarray:
    """order-preserving uint32 image of float32 (common.cuh f2key), -0 canonicalised"""
    b = (np.asarray(d, np.float32) + np.float32(0)).view(np.uint32)
    return np.where(b >> 31 != 0, ~b, b | np.uint32(0x80000000)).astype(np.uint64)


def dist_key(d, col) -> np.ndarray:
    return (_f2key(d) << np.uint64(32)) | np.asarray(col, np.uint64)


def _cell(d: np.ndarray, scale: np.float32) -> np.ndarray:
    x = np.asarray(d, np.float32) * scale                      # __fmul_rn
    x = np.where(x > 0, np.minimum(x, np.float32(4e9)), 0)     # __float2uint_rz: negative -> 0, saturating
    return np.minimum(np.trunc(x).astype(np.int64), NB - 1)


def cmc_map_counting(dist, q_pids, g_pids, q_cams, g_cams, max_rank=50, blocks=None, use_cells=True):
    """Returns (cmc float32[max_rank], mAP float64, ap[q] (-1 = skipped), first_hit[q])."""
    dist = np.asarray(dist, np.float32)
    nq, g = dist.shape
    blocks = blocks or [(0, g)]
    max_rank = min(max_rank, g)
    ap = np.full(nq, -1.0)
    first = np.full(nq, -1, np.int64)
    hist_cmc = np.zeros(max_rank, np.int64)
    for i in range(nq):
        grp = np.flatnonzero(g_pids == q_pids[i])                           # reid_stream_group_lookup
        junk = g_cams[grp] == q_cams[i]
        mk = np.sort(dist_key(dist[i, grp[~junk]], grp[~junk]))              # reid_stream_group_sort
        jk = dist_key(dist[i, grp[junk]], grp[junk])
        P = len(mk)
        if P == 0:
            continue
        dmatch = dist[i, (mk & np.uint64(0xffffffff)).astype(np.int64)]
        dmax = dmatch[-1]
        bins = np.zeros(P, np.int64)
        fast = use_cells and dmax > 0 and np.isfinite(dmax)
        scale = np.float32(NB) / np.float32(dmax) if fast else np.float32(0)
        if fast and not np.isfinite(scale):
            scale = np.float32(0)
        mcell = _cell(dmatch, scale) if fast else None
        for c0, c1 in blocks:                                               # reid_stream_count_panel, one call per block
            d = dist[i, c0:c1]
            cols = np.arange(c0, c1)
            if fast:
                near = d <= dmax
                cell = _cell(d[near], scale)
                marked = np.isin(cell, mcell)
                ex_d, ex_c = d[near][marked], cols[near][marked]
                cnt = np.bincount(cell[~marked], minlength=NB)
                for c in np.flatnonzero(cnt):                               # a cell without a match
                    lo = int(np.searchsorted(mcell, c, side="left"))        # #{matches in lower cells}
                    if lo < P:
                        bins[lo] += cnt[c]
            else:
                near = ~(d > dmax)
                ex_d, ex_c = d[near], cols[near]
            k = dist_key(ex_d, ex_c)
            k = k[k <= mk[-1]]
            np.add.at(bins, np.searchsorted(mk, k, side="left"), 1)          # #{match keys < k}
        prefix = np.cumsum(bins)                                            # reid_stream_score
        junk_before = np.array([(jk < m).sum() for m in mk])
        rank = prefix - junk_before                                         # 1-based among the kept entries
        ap[i] = float(np.sum(np.arange(1, P + 1, dtype=np.float64) / rank.astype(np.float64)) / P)
        first[i] = rank[0] - 1
        if first[i] < max_rank:
            hist_cmc[first[i]] += 1
    valid = ap >= 0
    assert valid.any(), "Error: all query identities do not appear in gallery"
    cmc = (np.cumsum(hist_cmc).astype(np.float32) / np.float32(valid.sum())).astype(np.float32)
    return cmc, float(np.mean(ap[valid])), ap, first


def topk_by_blocks(dist, K, blocks):
    """Engine.retrieve_topk's merge argument in numpy: stable top-K inside every column block, winners laid out block
    after block, stable top-K over that candidate matrix, candidates mapped back to global columns."""
    dist = np.asarray(dist, np.float32) + np.float32(0)
    cv, ci = [], []
    for c0, c1 in blocks:
        o = np.argsort(dist[:, c0:c1], axis=1, kind="stable")[:, :K]
        ci.append(o + c0)
        cv.append(np.take_along_axis(dist[:, c0:c1], o, 1))
    cv, ci = np.concatenate(cv, 1), np.concatenate(ci, 1)
    pos = np.argsort(cv, axis=1, kind="stable")[:, :K]
    return np.take_along_axis(ci, pos, 1), np.take_along_axis(cv, pos, 1)
