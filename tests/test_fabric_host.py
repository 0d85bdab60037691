"""CPU: host arithmetic of the peer-memory path (multimodal-reid_b200/fabric.py) -- snake-block ownership, the tile list
of a rank's share of the symmetric contraction, the arena layout.  The kernels mirror these functions (csrc/common.cuh);
the GPU side is checked end to end by tests/multi_rank_check.py (sharded result == one-rank result, bit for bit)."""
import importlib

import numpy as np
import pytest

fab = importlib.import_module("multimodal-reid_b200.fabric")


@pytest.mark.parametrize("P", [1, 2, 3, 4, 8])
def test_snake_ownership_is_a_partition(P):
    for T in (1, 2, 7, 16, 33, 367):
        seen = {}
        for t in range(T):
            r, lt = fab.snake_owner(t, P), fab.snake_local_tile(t, P)
            assert 0 <= r < P and fab.snake_global_tile(lt, r, P) == t
            assert (r, lt) not in seen
            seen[(r, lt)] = t
        for r in range(P):
            mine = sorted(t for t in range(T) if fab.snake_owner(t, P) == r)
            # local order is ascending global order, and tiles_below counts exactly the owned prefix
            assert [fab.snake_global_tile(lt, r, P) for lt in range(len(mine))] == mine
            for t_end in range(T + 1):
                assert fab.snake_tiles_below(t_end, r, P) == sum(1 for t in mine if t < t_end)


@pytest.mark.parametrize("P", [1, 2, 4, 8])
@pytest.mark.parametrize("n", [1, 255, 256, 257, 1400, 19889])
def test_rows_and_tiles_cover_the_matrix_once(P, n):
    rows = [fab.own_row_index(n, r, P) for r in range(P)]
    assert sorted(int(x) for a in rows for x in a) == list(range(n))
    assert [a.size for a in rows] == [fab.owned_rows(n, r, P) for r in range(P)]
    T = fab.n_tiles(n)
    tiles = set()
    for r in range(P):
        tl = fab.tile_list(n, r, P)
        assert tl.dtype == np.int32 and tl.shape[1] == 2
        for mb, nb in tl:
            assert mb <= nb < T and fab.snake_owner(int(mb), P) == r
            assert (int(mb), int(nb)) not in tiles
            tiles.add((int(mb), int(nb)))
    assert len(tiles) == T * (T + 1) // 2          # every upper-triangle tile exactly once over all ranks


def test_upper_triangle_work_is_balanced():
    # MSMT17 shape: 367 tile rows; the snake order balances tiles to < 0.1 %, cyclic tile rows would be off by ~4 % at 8 ranks
    for P in (2, 4, 8):
        cnt = [len(fab.tile_list(93820, r, P)) for r in range(P)]
        assert max(cnt) - min(cnt) <= 1 + 367 // (2 * P) // 20, cnt
    cnt = [len(fab.tile_list(19889, r, 8)) for r in range(8)]       # Duke: 78 tile rows
    assert max(cnt) <= 1.03 * (sum(cnt) / 8), cnt


def test_layout_is_deterministic_and_aligned():
    a = fab.make_layout(93820, 11659, 21, 252, 8, 93820 * 6 * 252)
    b = fab.make_layout(93820, 11659, 21, 252, 8, 93820 * 6 * 252)
    assert a == b
    offs = [a.flags, a.stats, a.inbox, a.nbr, a.vn, a.vc, a.vv, a.qn, a.qc, a.qv, a.scores, a.panel, a.total]
    assert offs == sorted(offs) and all(o % 256 == 0 for o in offs)
    assert a.stats_slot % 16 == 0 and a.stats_slot >= 2 * 93820 * 4 and a.score_slot % 16 == 0
    assert a.nloc_max % 256 == 0 and a.nloc_max >= fab.owned_rows_max(93820, 8)
    assert a.total - a.panel == a.nloc_max * a.ld * 4


def test_upper_half_sweep_panels_and_slabs():
    """Host arithmetic of the upper-half N x N sweep (engine.neighbour_lists_symmetric): panel boundaries are multiples
    of 256 with short tails joined, every rectangle right of a diagonal block splits into column blocks that tile
    [r1, N) exactly, slabs are multiples of 256 and only appear where a column's own block is short."""
    import importlib
    E = importlib.import_module("multimodal-reid_b200.engine").Engine
    for N, rpp, Kw in ((1_100_000, 16384, 64), (3204, 256, 64), (530, 256, 64), (200, 256, 64), (16384 * 3 + 5000, 16384, 64),
                       (16384 * 3 + 100, 16384, 128), (70_000, 1000, 32)):
        b = E.sym_panel_bounds(N, rpp, Kw)
        assert b[0] == 0 and b[-1] == N and all(x < y for x, y in zip(b, b[1:]))
        assert all(x % 256 == 0 for x in b[:-1])
        B = max(256, rpp // 256 * 256)
        sizes = [y - x for x, y in zip(b, b[1:])]
        assert all(s == B for s in sizes[:-1]) and (len(sizes) == 1 or max(Kw, B // 4) <= sizes[-1] < B + max(Kw, B // 4))
        for p in range(len(sizes)):
            r1 = b[p + 1]
            if r1 == N:
                continue
            blocks = [(r1 + c0, r1 + c1) for c0, c1 in E.column_blocks(N - r1, 4 * B, Kw)]
            assert blocks[0][0] == r1 and blocks[-1][1] == N
            assert all(a[1] == c[0] for a, c in zip(blocks, blocks[1:])) and all(c1 - c0 >= Kw for c0, c1 in blocks)
            for c0, c1 in blocks:
                slab = E.sym_slab_rows(b, sizes[p], c0, c1, Kw)
                assert 0 < slab <= max(sizes[p], 256) and (slab == sizes[p] or slab % 256 == 0)
                t = min(y - x for x, y in zip(b, b[1:]) if x < c1 and y > c0)
                if Kw * sizes[p] <= 96 * t:
                    assert slab == sizes[p]                      # regular rectangle at Kw = 64: pushed whole
                else:
                    assert Kw * min(slab, sizes[p]) / t <= 96 + Kw * 256 / t
    # the 100k x 1M case: 67 regular panels + a tail of 2 272 rows joined (2 272 < 16 384 / 4)
    b = E.sym_panel_bounds(1_100_000, 16384, 64)
    assert len(b) - 1 == 67 and b[-1] - b[-2] == 16384 + 2272
