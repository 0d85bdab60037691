"""TEST INFRASTRUCTURE ONLY (imported by tests/; never by the product path).

CPU restatement of the upper-half neighbour-list sweep (multimodal-reid_b200/csrc/stream_sym.cu +
Engine.neighbour_lists_symmetric): the Kw nearest columns of EVERY row of a symmetric matrix, in stable (value, column)
order, while only the diagonal blocks and the rectangles right of them are ever looked at.  What it follows in the
reference: reranking.py:46-48 needs `initial_rank[:, :k]` = argsort of every row of the N x N matrix; at the 100k x 1M
scale that matrix cannot exist, and half of it need not even be computed.

The point of the restatement is the exactness argument, checked against a plain full-row sort in tests/test_sym_oracle.py
(ties included): a column c of a rectangle receives every element <= tau[c], where tau[c] is the Kw-th smallest value
among entries row c has ALREADY been given -- an upper bound on its final Kw-th value -- so nothing that belongs in the
final list is ever filtered (non-strict comparison: an equal value with a smaller column index can still displace).
"""
from __future__ import annotations

import numpy as np


def _topk_keys(vals: np.ndarray, cols: np.ndarray, k: int):
    """k smallest of (value, column) pairs in stable (value, column) order."""
    order = np.lexsort((cols, vals))[:k]
    return vals[order], cols[order]


def panel_bounds(n: int, rows_per_pass: int, kw: int) -> list[int]:
    """Engine.sym_panel_bounds without the multiple-of-256 rounding (any block size: the CPU test uses small ones)."""
    b = max(1, rows_per_pass)
    starts = list(range(0, n, b))
    if len(starts) > 1 and n - starts[-1] < max(kw, b // 4):
        starts.pop()
    return starts + [n]


def upper_half_lists(D: np.ndarray, kw: int, rows_per_pass: int, cols_per_pass: int, slab: int | None = None,
                     world: int = 1):
    """D: symmetric [n, n].  Only D[r, c] with c >= (start of r's panel) is read.  Returns (idx [n, kw], val [n, kw],
    pushes) -- pushes = number of candidate records the column side produced (the work the thresholds save).
    world > 1: the panels are dealt out in snake order and every rank keeps its own copy of the lists (own rows: complete
    for the entries it has seen; other rows: the pushes of its panels); the copies are merged at the end."""
    n = D.shape[0]
    bounds = panel_bounds(n, rows_per_pass, kw)
    npan = len(bounds) - 1

    def owner(p):
        pos = p % (2 * world)
        return pos if pos < world else 2 * world - 1 - pos

    inf = np.float32(np.inf)
    Mv = np.full((world, n, kw), inf, dtype=D.dtype)
    Mi = np.full((world, n, kw), np.iinfo(np.int32).max, dtype=np.int64)
    tau = np.full((n,), inf, dtype=D.dtype)
    # diagonal blocks: initial lists and thresholds of the own rows (every rank learns all thresholds: all-reduce min)
    for p in range(npan):
        r0, r1 = bounds[p], bounds[p + 1]
        w = owner(p)
        for r in range(r0, r1):
            v, c = _topk_keys(D[r, r0:r1], np.arange(r0, r1), kw)
            Mv[w, r, : len(v)], Mi[w, r, : len(c)] = v, c
            tau[r] = Mv[w, r, kw - 1]
    taus = np.tile(tau, (world, 1))
    pushes = 0
    for w in range(world):
        for p in range(npan):
            if owner(p) != w:
                continue
            r0, r1 = bounds[p], bounds[p + 1]
            m = r1 - r0
            row_lists = [(Mv[w, r0:r1].copy(), Mi[w, r0:r1].copy())]
            step = max(kw, cols_per_pass)
            starts = list(range(r1, n, step))
            if len(starts) > 1 and n - starts[-1] < kw:
                starts.pop()
            for b, c0 in enumerate(starts):
                c1 = starts[b + 1] if b + 1 < len(starts) else n
                R = D[r0:r1, c0:c1]
                # rows of the rectangle: plain per-panel selection
                bv = np.empty((m, min(kw, c1 - c0)), dtype=D.dtype)
                bi = np.empty((m, min(kw, c1 - c0)), dtype=np.int64)
                for i in range(m):
                    bv[i], bi[i] = _topk_keys(R[i], np.arange(c0, c1), kw)
                row_lists.append((bv, bi))
                # columns of the rectangle: threshold-filtered pushes, folded slab by slab
                sl = m if slab is None else slab
                for s0 in range(0, m, sl):
                    s1 = min(m, s0 + sl)
                    for c in range(c0, c1):
                        take = np.nonzero(R[s0:s1, c - c0] <= taus[w, c])[0]
                        if take.size == 0:
                            continue
                        pushes += take.size
                        v = np.concatenate([Mv[w, c], R[s0 + take, c - c0]])
                        i = np.concatenate([Mi[w, c], r0 + s0 + take])
                        Mv[w, c], Mi[w, c] = _topk_keys(v, i, kw)
                        taus[w, c] = min(taus[w, c], Mv[w, c, kw - 1])
            # the panel's rows: own list so far (columns below r1) + the rectangles' winners
            for i in range(m):
                v = np.concatenate([x[0][i] for x in row_lists])
                c = np.concatenate([x[1][i] for x in row_lists])
                Mv[w, r0 + i], Mi[w, r0 + i] = _topk_keys(v, c, kw)
    # merge of the ranks' copies
    out_v = np.empty((n, kw), dtype=D.dtype)
    out_i = np.empty((n, kw), dtype=np.int64)
    for r in range(n):
        out_v[r], out_i[r] = _topk_keys(Mv[:, r].reshape(-1), Mi[:, r].reshape(-1), kw)
    return out_i, out_v, pushes


def full_row_lists(D: np.ndarray, kw: int):
    """The reference semantics: stable argsort of every full row, first kw (reranking.py:48 on raw distances)."""
    n = D.shape[0]
    idx = np.empty((n, kw), dtype=np.int64)
    val = np.empty((n, kw), dtype=D.dtype)
    for r in range(n):
        val[r], idx[r] = _topk_keys(D[r], np.arange(n), kw)
    return idx, val
