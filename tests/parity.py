"""Comparison helpers implementing north_star's tolerances (SURVEY.md §8a parity notes 4).

* distances: <= 1e-4 relative, except entries classified as float16-boundary flips — the
  reference's Jaccard chain is float16, so an fp32-ulp difference in the N x N matrix can move a
  final entry by 1..3 float16 quanta (multiples of 2^-12 * 0.7 .. 2^-11); those are counted and
  bounded (reference-vs-itself noise floor measured in the survey: ~1e-6 of entries).
* index lists: exact modulo groups whose keys are within 1e-6.
* Rank-k / mAP: <= 1e-5 absolute.
"""
import numpy as np

REL_TOL = 1e-4
TIE_TOL = 1e-6
METRIC_TOL = 1e-5
FP16_QUANTUM = 2.0 ** -11          # spacing of float16 in [0.5, 1): the Jaccard term lives there


def classify_distances(got: np.ndarray, want: np.ndarray):
    """Returns (n_bad, n_flips, max_rel_excluding_flips).  A flip = |diff| is (close to) a small
    multiple (<= 4) of half a float16 quantum times 0.7 (the lambda-mix weight), or of the quantum
    itself."""
    got = np.asarray(got, np.float64)
    want = np.asarray(want, np.float64)
    diff = np.abs(got - want)
    rel = diff / np.maximum(np.abs(want), 1e-12)
    over = rel > REL_TOL
    flips = over & (diff <= 4 * FP16_QUANTUM + 1e-6)
    bad = over & ~flips
    rest = rel[~over]
    return int(bad.sum()), int(flips.sum()), float(rest.max() if rest.size else 0.0)


def assert_distances_close(got, want, max_flip_frac=2e-5, what="distances"):
    n_bad, n_flip, mx = classify_distances(got, want)
    assert n_bad == 0, f"{what}: {n_bad} entries differ by more than {REL_TOL} relative (not fp16 flips)"
    assert n_flip <= max(3, max_flip_frac * np.asarray(want).size), \
        f"{what}: {n_flip} fp16-boundary flips of {np.asarray(want).size} entries"
    return n_flip, mx


def assert_ranking_equal_mod_ties(idx_got, idx_want, keys, tol=TIE_TOL, what="ranking"):
    """idx_*: [rows, K] index lists; keys: [rows, cols] the values being ranked.  Positions where
    the lists differ must hold keys within tol of each other."""
    idx_got = np.asarray(idx_got)
    idx_want = np.asarray(idx_want)
    assert idx_got.shape == idx_want.shape
    neq = idx_got != idx_want
    if not neq.any():
        return 0
    r = np.nonzero(neq)[0]
    kg = keys[r, idx_got[neq]]
    kw = keys[r, idx_want[neq]]
    assert np.all(np.abs(kg.astype(np.float64) - kw.astype(np.float64)) <= tol), \
        f"{what}: lists differ outside tie groups (max key gap {np.abs(kg - kw).max()})"
    return int(neq.sum())
