"""CPU: the cost-balanced tile plan of the streaming Poisson kernel (csrc/poisson_stream_plan.h), run on the host through
b200cfd_stream_plan_host with the very functions the device kernel k_stream_plan runs, and the chunk-flag arithmetic of
k_stream_make_rhs restated and checked against a brute-force walk of the pipeline's row indices.

The plan has no counterpart in the reference (v5.py:202-226 / :336-346 sweep the whole array); what is checked here are
the invariants the kernel relies on: every strip's rows are owned exactly once, cuts fall on multiples of 3 (stores are
whole 3-row TMA boxes), no more tiles than warps resident at once, and the tiles' costs are balanced."""
import ctypes as C

import numpy as np
import pytest

from cfd_simulations_b200 import _lib

SW, SKR = 128, 3


def geometry(L):
    H = (L + 3) // 4 * 4
    E = (3 - (2 * L) % 3) % 3
    return H, SW - 2 * H, E


def chunk_range(i, L, E, NC):
    """chunks whose levels read the rhs of row i (the formula of k_stream_make_rhs)"""
    return max((i - 1 + L + E + 2) // 3, 0), min((i + 2 * L + E) // 3, NC - 1)


def make_flags(ny, nx, L, special):
    """special: boolean (ny, nx) array of masked / never-updated cells -> flags [n_strips][NC] as k_stream_make_rhs marks them"""
    H, OW, E = geometry(L)
    n_strips = (nx + OW - 1) // OW
    NC = (ny + 2 * L + E) // SKR + 2
    flags = np.zeros((n_strips, NC), np.uint8)
    ii, jj = np.nonzero(special)
    for i, j in zip(ii.tolist(), jj.tolist()):
        s0, s1 = max((j - H) // OW, 0), min((j + H) // OW, n_strips - 1)
        c0, c1 = chunk_range(i, L, E, NC)
        flags[s0:s1 + 1, c0:c1 + 1] = 1
    return flags, n_strips, NC, (2 * L + E) // SKR


def run_plan(flags, n_strips, NC, ny, shift, ntile, alpha=35, min_groups=4):
    lib = _lib.load()
    out = np.zeros((ntile, 4), np.int32)
    f = np.ascontiguousarray(flags)
    rc = lib.b200cfd_stream_plan_host(f.ctypes.data_as(C.c_void_p), n_strips, NC, ny, shift, ntile, alpha, min_groups,
                                      out.ctypes.data_as(C.c_void_p))
    assert rc == 0
    return out


def check_partition(tiles, n_strips, ny, min_groups=4):
    live = tiles[tiles[:, 2] > tiles[:, 1]]
    for s in range(n_strips):
        t = live[live[:, 0] == s]
        assert len(t) >= 1, f"strip {s} has no tile"
        t = t[np.argsort(t[:, 1])]
        assert t[0, 1] == 0 and t[-1, 2] == ny
        assert np.all(t[1:, 1] == t[:-1, 2]), f"strip {s}: gap or overlap"
        assert np.all(t[:, 1] % 3 == 0)
        assert np.all(t[:-1, 2] % 3 == 0)
        if len(t) > 1:
            assert np.all(t[:, 2] - t[:, 1] >= 3 * (min_groups - 1))
    return live


def tile_costs(live, flags, shift, alpha):
    NC = flags.shape[1]
    cost = []
    for s, r0, r1, _ in live.tolist():
        g = np.arange(r0 // 3, (r1 + 2) // 3)
        c = np.minimum(g + shift, NC - 1)
        cost.append(100 * len(g) + alpha * int(flags[s, c].sum()))
    return np.array(cost, float)


def test_chunk_flag_range_matches_the_pipeline_walk():
    # chunk c holds phi rows 3c - (L + E) + j, j = 0..2; level lv (0-based) computes, and reads the rhs of, row q + j - 1 - lv
    for L in (2, 6):
        _, _, E = geometry(L)
        NC = 200
        reads = {}
        for c in range(NC):
            q = 3 * c - (L + E)
            for j in range(3):
                for lv in range(L):
                    reads.setdefault(q + j - 1 - lv, set()).add(c)
        for i in range(0, 400):
            c0, c1 = chunk_range(i, L, E, NC)
            want = {c for c in reads.get(i, set())}
            assert set(range(c0, c1 + 1)) == want, (L, i)


@pytest.mark.parametrize("alpha", [0, 20, 35, 60, 100, 150])
def test_plan_8192_square_is_a_balanced_partition(alpha):
    ny = nx = 8192
    L = 6
    special = np.zeros((ny, nx), bool)
    special[0, :] = special[-1, :] = True
    special[:, 0] = special[:, -1] = True
    y, x = np.ogrid[:ny, :nx]
    disc = (x - nx // 2) ** 2 + (y - ny // 2) ** 2 <= (nx // 40) ** 2
    special |= disc
    # thin the point set: per row keep first/last special column of each 8-column group (same flags, far fewer points)
    keep = np.zeros_like(special)
    ii, jj = np.nonzero(special)
    sel = (jj % 8 == 0) | (jj == nx - 1)
    keep[ii[sel], jj[sel]] = True
    flags, n_strips, NC, shift = make_flags(ny, nx, L, keep)
    ntile = 8 * 148
    tiles = run_plan(flags, n_strips, NC, ny, shift, ntile, alpha)
    live = check_partition(tiles, n_strips, ny)
    # (tiles the minimax allocation has no use for stay empty: one more tile for SOME of the equal strips would not lower the maximum)
    assert ntile - n_strips < len(live) <= ntile
    cost = tile_costs(live, flags, shift, alpha)
    # whole tiles per strip: 1184 tiles over 74 strips of slightly different cost leave some strips with 15 and some with 16
    assert cost.max() / (cost.sum() / ntile) < 1.08, cost.max() / (cost.sum() / ntile)
    # ... and no allocation does better: with one tile fewer in total the most expensive tile could not get cheaper
    strip_cost = np.array([cost[live[:, 0] == s].sum() for s in range(n_strips)])
    n_s = np.array([np.sum(live[:, 0] == s) for s in range(n_strips)])
    T = (strip_cost / n_s).max()
    assert np.ceil(strip_cost / (T * 0.999)).sum() > ntile - n_strips
    # table order: row-major over (tile-in-strip, strip): the first n_strips entries are the first tiles of strips 0, 1, 2, ...
    assert np.array_equal(tiles[:n_strips, 0], np.arange(n_strips)) and np.all(tiles[:n_strips, 1] == 0)
    assert np.all(tiles[len(live):, 2] == 0)                      # unused entries trail
    if alpha == 0:          # no cost difference: equal row ranges (to within one group)
        h = live[:, 2] - live[:, 1]
        assert h.max() - h.min() <= 6
    else:                   # the all-flagged edge strips get more, shorter tiles
        n_edge = np.sum(live[:, 0] == 0)
        n_mid = np.sum(live[:, 0] == 5)
        assert n_edge > n_mid


@pytest.mark.parametrize("ny,nx,ntile", [(16, 64, 1), (16, 64, 8), (130, 258, 1184), (1200, 160, 1184), (512, 512, 1184),
                                         (37, 4096, 40), (8192, 2048 + 112, 1184)])
def test_plan_small_and_ragged_shapes(ny, nx, ntile):
    for L in (2, 6):
        H, OW, E = geometry(L)
        special = np.zeros((ny, nx), bool)
        special[0, :] = special[-1, :] = True
        special[:, 0] = special[:, -1] = True
        special[ny // 3:ny // 2, nx // 4:nx // 3] = True
        flags, n_strips, NC, shift = make_flags(ny, nx, L, special)
        G = (ny + 2) // 3
        nt = max(n_strips, min(ntile, n_strips * max(1, G // 4)))      # what poisson_stream_run asks for
        tiles = run_plan(flags, n_strips, NC, ny, shift, nt)
        live = check_partition(tiles, n_strips, ny)
        assert len(live) <= nt


def test_plan_rejects_bad_arguments():
    lib = _lib.load()
    out = np.zeros((4, 4), np.int32)
    f = np.zeros((2, 10), np.uint8)
    assert lib.b200cfd_stream_plan_host(f.ctypes.data_as(C.c_void_p), 2, 10, 20, 4, 1, 35, 4, out.ctypes.data_as(C.c_void_p)) == -1
    assert lib.b200cfd_stream_plan_host(None, 2, 10, 20, 4, 4, 35, 4, out.ctypes.data_as(C.c_void_p)) == -1
