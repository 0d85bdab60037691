"""CPU: the upper-half neighbour-list sweep (oracle/sym_oracle.py restates csrc/stream_sym.cu + the host sequencing)
gives exactly the lists of a stable full-row sort -- ties, slabs, short tail panels and several ranks included."""
import numpy as np
import pytest

from oracle import sym_oracle as so


def _sym(n, seed, levels=None):
    rng = np.random.default_rng(seed)
    x = rng.standard_normal((n, 16)).astype(np.float32)
    x /= np.linalg.norm(x, axis=1, keepdims=True)
    D = (2 - 2 * x @ x.T).astype(np.float32)
    D = np.maximum(D, D.T)                       # exactly symmetric
    if levels:                                   # heavy ties: a few distinct values only
        D = (np.round(D * levels) / levels).astype(np.float32)
    return D


@pytest.mark.parametrize("n,kw,rpp,cpp,slab,world,levels", [
    (97, 8, 16, 32, None, 1, None),
    (97, 8, 16, 32, 4, 1, 6),                    # ties everywhere, pushes in slabs of 4 rows
    (130, 12, 32, 48, None, 3, None),            # three ranks, snake-dealt panels
    (130, 12, 32, 48, 8, 2, 5),
    (75, 10, 64, 64, None, 1, None),             # short tail joins the last panel
    (40, 16, 64, 64, None, 2, 4),                # one panel only: nothing but the diagonal block
])
def test_upper_half_lists_equal_full_row_sort(n, kw, rpp, cpp, slab, world, levels):
    D = _sym(n, n * 7 + kw, levels)
    want_i, want_v = so.full_row_lists(D, kw)
    got_i, got_v, pushes = so.upper_half_lists(D, kw, rpp, cpp, slab, world)
    np.testing.assert_array_equal(got_i, want_i)
    np.testing.assert_array_equal(got_v, want_v)
    # the thresholds do filter: far fewer candidate records than mirrored elements
    bounds = so.panel_bounds(n, rpp, kw)
    mirrored = sum((b1 - b0) * (n - b1) for b0, b1 in zip(bounds, bounds[1:]))
    if levels is None and mirrored:
        assert pushes < mirrored


def test_strict_threshold_would_lose_ties():
    """Why the push compares with <=: an element equal to the running Kw-th value but with a smaller column index must
    still displace it."""
    n, kw = 24, 2
    D = np.full((n, n), 5.0, dtype=np.float32)
    np.fill_diagonal(D, 0.0)
    D[20, 21] = D[21, 20] = 1.0                  # row 20's block list: (0, col 20), (1, col 21) -> tau = 1
    D[3, 20] = D[20, 3] = 1.0                    # equal to tau, smaller column: belongs in front of col 21
    want_i, _ = so.full_row_lists(D, kw)
    got_i, _, _ = so.upper_half_lists(D, kw, 8, 8)
    assert want_i[20].tolist() == [20, 3]
    np.testing.assert_array_equal(got_i, want_i)
