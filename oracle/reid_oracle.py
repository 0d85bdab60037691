"""CPU oracle for the retrieval-evaluation hot path.  TEST INFRASTRUCTURE ONLY.

This module is a NumPy restatement of the algorithm implemented by the reference's
``evaluate.py`` and ``reranking.py`` (SuperbTUM/Multimodal-ReID @ 8dc4505d).  It exists so that
the CUDA path can be checked on machines where ``/root/reference`` is not mounted (the GPU box).
Only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference`` legs
of ``bench.py`` may import it; the product package never does.

Parity pin: the reference ships no tests or golden vectors for this path, so the
pin is the unmodified reference itself, imported from ``/root/reference`` in the build container:
``tests/test_oracle_vs_reference.py`` checks every function below bit-for-bit (index lists,
fp16 chains, fp32 outputs) against it, and ``oracle/make_golden.py`` freezes reference outputs
into ``tests/golden/*.npz`` that travel to the GPU box.

The restatement is *sparse*: the reference materialises a dense N x N float16 ``V`` and a full
argsort; only the first max(k1+1, k2) columns of the argsort and the non-zeros of ``V`` are ever
read, so the lists below carry exactly the same information (verified bit-identical).

Citations are ``file:line`` in the reference tree.
"""
from __future__ import annotations

import numpy as np

# --------------------------------------------------------------------------------------------
# distance matrices                                                          evaluate.py:7-26
# --------------------------------------------------------------------------------------------


def sq_euclidean(qf: np.ndarray, gf: np.ndarray) -> np.ndarray:
    """‖q‖² + ‖g‖² − 2 q·gᵀ in float32, no sqrt, no clamp (evaluate.py:7-13, reranking.py:36-41)."""
    qf = np.asarray(qf, dtype=np.float32)
    gf = np.asarray(gf, dtype=np.float32)
    qq = np.sum(qf * qf, axis=1, dtype=np.float32)[:, None]
    gg = np.sum(gf * gf, axis=1, dtype=np.float32)[None, :]
    base = (qq + gg).astype(np.float32)
    return (base + np.float32(-2.0) * (qf @ gf.T)).astype(np.float32)


def arccos_distance(qf: np.ndarray, gf: np.ndarray) -> np.ndarray:
    """arccos(clip(q·g · (1 / (‖q‖‖g‖)), ±(1 − 1e-5)))  (evaluate.py:16-26).

    The reference multiplies by the reciprocal of the norm product rather than dividing."""
    qf = np.asarray(qf, dtype=np.float32)
    gf = np.asarray(gf, dtype=np.float32)
    eps = 0.00001
    dots = qf @ gf.T
    qn = np.sqrt(np.sum(qf * qf, axis=1, dtype=np.float32))[:, None]
    gn = np.sqrt(np.sum(gf * gf, axis=1, dtype=np.float32))[None, :]
    recip = (np.float32(1.0) / (qn * gn)).astype(np.float32)
    c = np.clip(dots * recip, -1 + eps, 1 - eps)
    return np.arccos(c)


def l2_normalize(x: np.ndarray, eps: float = 1e-12) -> np.ndarray:
    """Row L2 normalisation with the denominator clamped at eps (evaluate.py:114)."""
    x = np.asarray(x, dtype=np.float32)
    n = np.sqrt(np.sum(x * x, axis=1, dtype=np.float32))[:, None]
    return (x / np.maximum(n, np.float32(eps))).astype(np.float32)


# --------------------------------------------------------------------------------------------
# numpy float32 exp / pairwise sum, restated so the CUDA kernels can follow them bit-for-bit
# --------------------------------------------------------------------------------------------

_EXP_P = [np.float32(v) for v in (
    9.999999999980870924916e-01, 7.257664613233124478488e-01, 2.473615434895520810817e-01,
    5.114512081637298353406e-02, 6.757896990527504603057e-03, 5.082762527590693718096e-04)]
_EXP_Q = [np.float32(v) for v in (
    1.0, -2.742335390411667452936e-01, 2.159509375685829852307e-02)]
_LOG2E = np.float32(1.44269504088896340736)
_RINT_MAGIC = np.float32(12582912.0)  # 0x1.8p23
_LN2_HI = np.float32(-6.93145752e-1)
_LN2_LO = np.float32(-1.42860677e-6)


def _fma32(a, b, c):
    """float32 fused multiply-add via float64 (exact product, single rounding of the sum is
    emulated: the double sum of a 48-bit product and a 24-bit addend can double-round only in
    half-way cases that do not occur for the polynomial ranges used here; verified against
    np.exp in tests)."""
    return (a.astype(np.float64) * b.astype(np.float64) + c.astype(np.float64)).astype(np.float32)


def exp_f32_restated(x: np.ndarray) -> np.ndarray:
    """numpy's AVX2/AVX512 float32 exp: range reduction by round-to-nearest multiples of ln2
    (Cody–Waite split), a (5,2) rational minimax, then scaling by 2^q.  The CUDA kernel
    ``np_expf`` in csrc/np_exact.cuh executes the same sequence with __fmaf_rn / __fdiv_rn."""
    x = np.asarray(x, dtype=np.float32)
    q = (x * _LOG2E).astype(np.float32)
    q = ((q + _RINT_MAGIC).astype(np.float32) - _RINT_MAGIC).astype(np.float32)
    r = _fma32(q, np.full_like(x, _LN2_HI), x)
    r = _fma32(q, np.full_like(x, _LN2_LO), r)
    num = np.full_like(x, _EXP_P[5])
    for k in (4, 3, 2, 1, 0):
        num = _fma32(num, r, np.full_like(x, _EXP_P[k]))
    den = _fma32(np.full_like(x, _EXP_Q[2]), r, np.full_like(x, _EXP_Q[1]))
    den = _fma32(den, r, np.full_like(x, _EXP_Q[0]))
    y = (num / den).astype(np.float32)
    return np.ldexp(y, q.astype(np.int32)).astype(np.float32)


def pairwise_sum_f32_restated(a: np.ndarray) -> np.float32:
    """numpy's contiguous float32 add-reduce: <8 sequential; <=128 eight strided accumulators
    combined as ((r0+r1)+(r2+r3))+((r4+r5)+(r6+r7)) plus a sequential tail; above that split at
    n/2 rounded down to a multiple of 8 and recurse."""
    a = np.asarray(a, dtype=np.float32)
    n = a.shape[0]
    if n < 8:
        s = np.float32(0.0)
        # numpy starts from the first element, not from zero (matters only for -0.0)
        if n == 0:
            return s
        s = a[0]
        for v in a[1:]:
            s = np.float32(s + v)
        return s
    if n <= 128:
        r = [a[k] for k in range(8)]
        i = 8
        while i < n - (n % 8):
            for k in range(8):
                r[k] = np.float32(r[k] + a[i + k])
            i += 8
        res = np.float32(np.float32(np.float32(r[0] + r[1]) + np.float32(r[2] + r[3]))
                         + np.float32(np.float32(r[4] + r[5]) + np.float32(r[6] + r[7])))
        while i < n:
            res = np.float32(res + a[i])
            i += 1
        return res
    n2 = n // 2
    n2 -= n2 % 8
    return np.float32(pairwise_sum_f32_restated(a[:n2]) + pairwise_sum_f32_restated(a[n2:]))


# --------------------------------------------------------------------------------------------
# k-reciprocal re-ranking                                                  reranking.py:29-100
# --------------------------------------------------------------------------------------------


def half_k(k1: int) -> int:
    """int(np.around(k1 / 2)) + 1 — banker's rounding (reranking.py:60)."""
    return int(np.around(k1 / 2)) + 1


def normalized_rows(D: np.ndarray) -> np.ndarray:
    """O[i, j] = D[j, i] / max_k D[k, i]  (reranking.py:46), returned C-contiguous."""
    D = np.asarray(D, dtype=np.float32)
    colmax = np.max(D, axis=0)
    return np.ascontiguousarray((D / colmax).T)


def neighbour_lists(O: np.ndarray, K: int) -> np.ndarray:
    """First K columns of a *stable* argsort of every row (reranking.py:48 reads only those).

    The reference's argsort is unstable; lists agree with it except inside groups of equal
    keys, which is the documented tolerance."""
    N = O.shape[0]
    K = min(K, O.shape[1])
    out = np.empty((N, K), dtype=np.int32)
    for i in range(N):
        row = O[i]
        if K < row.shape[0]:
            kth = np.partition(row, K - 1)[K - 1]
            cand = np.nonzero(row <= kth)[0]
        else:
            cand = np.arange(row.shape[0])
        order = np.argsort(row[cand], kind="stable")
        out[i] = cand[order[:K]]
    return out


def reciprocal_sets(nbr: np.ndarray, depth: int):
    """R_depth(i) = { nbr[i, r] : r < depth and i in nbr[nbr[i, r], :depth] } in list order
    (reranking.py:53-56 with depth = k1+1; :60-64 with depth = round(k1/2)+1).
    Returns a boolean mask [N, depth] over nbr[:, :depth]."""
    N = nbr.shape[0]
    fwd = nbr[:, :depth]
    mask = np.zeros((N, depth), dtype=bool)
    me = np.arange(N, dtype=nbr.dtype)[:, None]
    for r in range(depth):
        back = nbr[fwd[:, r], :depth]
        mask[:, r] = (back == me).any(axis=1)
    return mask


def expansion_sets(nbr: np.ndarray, k1: int):
    """E_i = unique-sorted( R_i ∪ ⋃{ Rc : c ∈ R_i, |Rc ∩ R_i| > 2/3 |Rc| } )  (reranking.py:57-69).
    Returns a list of int32 arrays."""
    N = nbr.shape[0]
    h = half_k(k1)
    full = reciprocal_sets(nbr, k1 + 1)
    half = reciprocal_sets(nbr, h)
    half_lists = [nbr[c, :h][half[c]] for c in range(N)]
    out = []
    for i in range(N):
        R = nbr[i, :k1 + 1][full[i]]
        Rset = set(R.tolist())
        parts = [R]
        for c in R.tolist():
            Rc = half_lists[c]
            inter = sum(1 for v in Rc.tolist() if v in Rset)
            if inter > 2 / 3 * len(Rc):
                parts.append(Rc)
        out.append(np.unique(np.concatenate(parts)).astype(np.int32))
    return out


def gaussian_rows(O: np.ndarray, E: list) -> list:
    """V[i, E_i] = float16( exp(−O[i, E_i]) / Σ exp(−O[i, E_i]) )  (reranking.py:70-71)."""
    vals = []
    for i, idx in enumerate(E):
        w = np.exp(-O[i, idx])
        vals.append((w / np.sum(w)).astype(np.float16))
    return vals


def query_expansion(cols: list, vals: list, nbr: np.ndarray, k2: int):
    """V_qe[i] = float16( float32-sum_r V[nbr[i, r]] / k2 ), r < k2  (reranking.py:73-78).
    Sparse: the union of the k2 supports, summed in neighbour order."""
    if k2 == 1:
        return cols, vals
    N = len(cols)
    ocols, ovals = [], []
    kk = np.float32(k2)
    for i in range(N):
        src = nbr[i, :k2]
        u = np.unique(np.concatenate([cols[s] for s in src]))
        acc = np.zeros(u.shape[0], dtype=np.float32)
        for s in src:
            pos = np.searchsorted(u, cols[s])
            acc[pos] = acc[pos] + vals[s].astype(np.float32)
        ocols.append(u.astype(np.int32))
        ovals.append((acc / kk).astype(np.float16))
    return ocols, ovals


def inverted_index(cols: list, vals: list, N: int):
    """For every column c: ascending rows r with V[r, c] != 0 and those values (reranking.py:80-82)."""
    rows = np.concatenate([np.full(c.shape[0], i, dtype=np.int32) for i, c in enumerate(cols)])
    cc = np.concatenate(cols)
    vv = np.concatenate(vals)
    keep = vv != 0
    rows, cc, vv = rows[keep], cc[keep], vv[keep]
    order = np.lexsort((rows, cc))
    rows, cc, vv = rows[order], cc[order], vv[order]
    ptr = np.zeros(N + 1, dtype=np.int64)
    np.add.at(ptr, cc.astype(np.int64) + 1, 1)
    ptr = np.cumsum(ptr)
    return ptr, rows, vv


def jaccard_rows(cols: list, vals: list, inv, q: int, N: int) -> np.ndarray:
    """float16 chain of reranking.py:84-93: per query, columns in ascending order, every
    ``t + min`` rounded to float16, then 1 − t / (2 − t) with each op rounded to float16."""
    ptr, rows, vv = inv
    J = np.empty((q, N), dtype=np.float16)
    for i in range(q):
        t = np.zeros(N, dtype=np.float16)
        for c, v in zip(cols[i].tolist(), vals[i]):
            if v == 0:
                continue
            a, b = ptr[c], ptr[c + 1]
            tgt = rows[a:b]
            t[tgt] = t[tgt] + np.minimum(v, vv[a:b])
        J[i] = 1 - t / (2 - t)
    return J


def rerank_from_dist(D: np.ndarray, q: int, k1: int, k2: int, lam: float, detail: bool = False):
    """Everything after the N x N distance matrix (reranking.py:45-100).  ``D`` is what the
    reference calls ``original_dist`` before line 46 (it reads *column* i for probe i)."""
    N = D.shape[0]
    O = normalized_rows(D)
    K = max(k1 + 1, k2)
    nbr = neighbour_lists(O, K)
    E = expansion_sets(nbr, k1)
    Vv = gaussian_rows(O, E)
    Qc, Qv = query_expansion(E, Vv, nbr, k2)
    inv = inverted_index(Qc, Qv, N)
    J = jaccard_rows(Qc, Qv, inv, q, N)
    final = J * (1 - lam) + O[:q] * lam          # float16*pyfloat -> float16 ; float32*pyfloat -> float32
    final = final[:q, q:]
    if detail:
        return dict(O=O, nbr=nbr, E=E, V=Vv, Qc=Qc, Qv=Qv, inv=inv, J=J, final=final)
    return final


def re_ranking(qf, gf, k1, k2, lam):
    """Feature entry (reranking.py:29-44 + the rest)."""
    feat = np.concatenate([np.asarray(qf, np.float32), np.asarray(gf, np.float32)])
    D = sq_euclidean(feat, feat)
    return rerank_from_dist(D, qf.shape[0], k1, k2, lam)


# --------------------------------------------------------------------------------------------
# CMC / mAP                                                                  evaluate.py:29-88
# --------------------------------------------------------------------------------------------


def cmc_map(distmat, q_pids, g_pids, q_camids, g_camids, max_rank=50, detail=False):
    """Market-1501 protocol: gallery entries with the query's pid *and* camera are discarded,
    queries without a remaining match are skipped (evaluate.py:55-63), CMC is the clipped
    cumulative match curve truncated at max_rank (:65-68), AP the mean precision at the
    matches in float64 (:73-80).  Rows are ranked with a stable sort."""
    distmat = np.asarray(distmat)
    q_pids, g_pids = np.asarray(q_pids), np.asarray(g_pids)
    q_camids, g_camids = np.asarray(q_camids), np.asarray(g_camids)
    nq, ng = distmat.shape
    max_rank = min(max_rank, ng)
    hits = np.zeros(max_rank, dtype=np.int64)
    aps, first = [], []
    for i in range(nq):
        order = np.argsort(distmat[i], kind="stable")
        pid_o, cam_o = g_pids[order], g_camids[order]
        junk = (pid_o == q_pids[i]) & (cam_o == q_camids[i])
        rel = (pid_o == q_pids[i])[~junk]
        if not rel.any():
            first.append(-1)
            aps.append(np.nan)
            continue
        pos = np.nonzero(rel)[0]
        first.append(int(pos[0]))
        if pos[0] < max_rank:
            hits[pos[0]:] += 1
        prec = np.arange(1, pos.shape[0] + 1, dtype=np.float64) / (pos + 1.0)
        aps.append(prec.sum() / pos.shape[0])
    aps = np.asarray(aps, dtype=np.float64)
    valid = ~np.isnan(aps)
    assert valid.any(), "Error: all query identities do not appear in gallery"
    nvalid = float(valid.sum())
    cmc = hits.astype(np.float32) / nvalid
    mAP = np.mean(aps[valid])
    if detail:
        return cmc.astype(np.float32), mAP, aps, np.asarray(first)
    return cmc.astype(np.float32), mAP


def evaluate_features(feats, pids, camids, num_query, feat_norm=True, reranking=False,
                      k1=50, k2=15, lam=0.3, max_rank=50):
    """R1_mAP_eval.compute() (evaluate.py:110-135): normalise, split, distance or re-rank, score.
    The reference hard-codes k1=50,k2=15,λ=0.3 (evaluate.py:127) and never forwards max_rank."""
    feats = np.asarray(feats, dtype=np.float32)
    if feat_norm:
        feats = l2_normalize(feats)
    qf, gf = feats[:num_query], feats[num_query:]
    pids, camids = np.asarray(pids), np.asarray(camids)
    if reranking:
        dist = re_ranking(qf, gf, k1, k2, lam)
    else:
        dist = sq_euclidean(qf, gf)
    return cmc_map(dist, pids[:num_query], pids[num_query:], camids[:num_query], camids[num_query:],
                   max_rank=max_rank)
