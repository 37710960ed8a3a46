"""CPU tests of the host logic of libgprb200_dist.so (gprb200_dist_plan: no GPU, no NCCL call): every lower block has exactly
one owner on every grid and map, local offsets tile the local matrix, the staircase table matches the ownership, and the
flop balance the mirrored row map was introduced for holds."""
import numpy as np
import pytest

from gpr_b200 import capi_dist


def _plans(N, nb, Pr, Pc, **kw):
    return [capi_dist.plan(N, nb, Pr, Pc, r, **kw) for r in range(Pr * Pc)]


@pytest.mark.parametrize("N,nb,Pr,Pc", [(5000, 512, 1, 1), (5000, 512, 2, 1), (5000, 512, 1, 2), (9000, 256, 2, 2), (16384, 512, 4, 2),
                                        (10000, 128, 2, 4), (7777, 384, 3, 2), (4096, 1024, 1, 8)])
@pytest.mark.parametrize("mr,mc", [(True, False), (False, False), (True, True)])
def test_every_block_has_one_owner_and_offsets_tile_the_local_matrix(N, nb, Pr, Pc, mr, mc):
    nblk = (N + nb - 1) // nb
    hgt = [min(nb, N - i * nb) for i in range(nblk)]
    plans = _plans(N, nb, Pr, Pc, mirror_rows=mr, mirror_cols=mc)
    owners = np.zeros((nblk, nblk), dtype=int)
    for r, p in enumerate(plans):
        rows = [i for i in range(nblk) if p["row_offset"][i] >= 0]
        cols = [j for j in range(nblk) if p["col_offset"][j] >= 0]
        # offsets are the running sums of the owned block heights / widths, in increasing block order
        assert [p["row_offset"][i] for i in rows] == list(np.cumsum([0] + [hgt[i] for i in rows])[:-1])
        assert [p["col_offset"][j] for j in cols] == list(np.cumsum([0] + [hgt[j] for j in cols])[:-1])
        assert p["H"] == sum(hgt[i] for i in rows) and p["Wtot"] == sum(hgt[j] for j in cols)
        assert p["bytes"] == 8.0 * max(p["H"], 1) * max((p["Wtot"] + 1) // 2 * 2, 2)
        for i in rows:
            for j in cols:
                owners[i, j] += 1
        # grid position: rank r sits at row r // Pc, column r % Pc, and ranks of one grid row own the same block rows
        same_row = plans[(r // Pc) * Pc]
        assert np.array_equal(p["row_offset"], same_row["row_offset"])
        assert np.array_equal(p["col_offset"], plans[r % Pc]["col_offset"])
        # staircase: the first local row a column needs is the offset of its first owned block row at or below the diagonal
        for j in cols:
            below = [i for i in rows if i >= j]
            first = p["row_offset"][below[0]] if below else p["H"]
            c0 = p["col_offset"][j]
            for c in range(0, hgt[j], 64):
                want = first + (c if (below and below[0] == j) else 0)
                assert p["stair"][(c0 + c) // 64] == want
    assert np.all(owners == 1)


def test_mirrored_row_map_balances_the_lower_triangle():
    """N = 131072, nb = 1024 (BASELINE config C5) on 8 GPUs: flop of the busiest rank over N^3 / (3 g)."""
    N, nb = 131072, 1024
    ideal = N ** 3 / 3.0 / 8
    worst = {}
    for name, (Pr, Pc, mr) in {"2x4 plain": (2, 4, False), "4x2 plain": (4, 2, False), "4x2 mirrored": (4, 2, True)}.items():
        f = [p["flops"] for p in _plans(N, nb, Pr, Pc, mirror_rows=mr)]
        worst[name] = max(f) / ideal
        assert sum(f) >= 8 * ideal                       # nothing is left out: full diagonal tiles and the panel solves add work
        assert sum(f) <= 8 * ideal * 1.03
    assert worst["4x2 mirrored"] < 1.03 < worst["2x4 plain"] < worst["4x2 plain"]


def test_plan_rejects_bad_arguments():
    lib = capi_dist.load_dist_library()
    out = np.zeros(6)
    assert lib.gprb200_dist_plan(1000, 200, 1, 1, 0, 1, 0, out.ctypes.data, None, None, None, 0) == -2   # nb not a multiple of 128
    assert lib.gprb200_dist_plan(1000, 256, 2, 2, 4, 1, 0, out.ctypes.data, None, None, None, 0) == -2   # rank outside the grid
