"""T0: RowsDivision / neighbour tables (mpi/mpi_convolution.c:350-364,142-149)."""
import pytest

import b200conv
from b200conv import B2CError, block_of, decompose, neighbours, rows_division

# (W, H) -> {n: (P, Q) or None}
APP_B = {
    (1920, 630): {1: (1, 1), 2: (1, 2), 4: (1, 4), 8: (2, 4), 9: (3, 3), 16: (2, 8), 25: (5, 5)},
    (1920, 1260): {1: (1, 1), 2: (1, 2), 4: (2, 2), 8: (2, 4), 9: (3, 3), 16: (4, 4), 25: (5, 5)},
    (1920, 2520): {1: (1, 1), 2: (2, 1), 4: (2, 2), 8: (4, 2), 9: (3, 3), 16: (4, 4), 25: (5, 5)},
    (1920, 5040): {1: (1, 1), 2: (2, 1), 4: (4, 1), 8: (4, 2), 9: (3, 3), 16: (8, 2), 25: (5, 5)},
    (16384, 16384): {1: (1, 1), 2: (1, 2), 4: (2, 2), 8: (2, 4), 9: None, 16: (4, 4), 25: None},
    (32768, 32768): {1: (1, 1), 2: (1, 2), 4: (2, 2), 8: (2, 4), 9: None, 16: (4, 4), 25: None},
    (32768, 16384): {2: (1, 2)},
    (65536, 32768): {8: (2, 4)},
}


@pytest.mark.parametrize("wh", sorted(APP_B))
def test_rows_division_table(wh):
    w, h = wh
    for n, want in APP_B[wh].items():
        p = rows_division(n, h, w)
        if want is None:
            assert p == 0
            with pytest.raises(B2CError, match="Cannot divide to processes"):
                decompose(n, h, w)
        else:
            assert (p, n // p) == want
            assert decompose(n, h, w) == want


def test_rows_division_bruteforce_rule():
    # smallest semi-perimeter among exact divisors, ties -> smallest row count
    for n in (1, 2, 3, 4, 6, 8, 12):
        for h in (6, 12, 24, 30):
            for w in (4, 12, 18, 48):
                best, cost = 0, h + w + 1
                for p in range(1, n + 1):
                    if n % p or h % p or w % (n // p):
                        continue
                    c = h // p + w // (n // p)
                    if c < cost:
                        best, cost = p, c
                assert rows_division(n, h, w) == best


def test_neighbours_zero_matches_reference_rule():
    # N = rank-Q unless first row, S = rank+Q unless last row, W/E = rank-/+1 (:142-149)
    for P, Q in ((1, 1), (1, 2), (2, 1), (2, 2), (2, 4), (4, 2), (3, 3)):
        for rank in range(P * Q):
            nb = neighbours(rank, (P, Q), "zero")
            pr, pc = divmod(rank, Q)
            assert nb["N"] == (rank - Q if pr > 0 else -1)
            assert nb["S"] == (rank + Q if pr < P - 1 else -1)
            assert nb["W"] == (rank - 1 if pc > 0 else -1)
            assert nb["E"] == (rank + 1 if pc < Q - 1 else -1)
            assert nb["NW"] == (rank - Q - 1 if pr > 0 and pc > 0 else -1)
            assert nb["SE"] == (rank + Q + 1 if pr < P - 1 and pc < Q - 1 else -1)


def test_neighbours_torus_wraps_and_self():
    assert set(neighbours(0, (1, 1), "torus").values()) == {0}
    nb = neighbours(0, (1, 2), "torus")
    assert nb["N"] == 0 and nb["S"] == 0 and nb["W"] == 1 and nb["E"] == 1 and nb["NW"] == 1
    nb = neighbours(5, (2, 4), "torus")  # row 1, col 1
    assert nb == {"N": 1, "S": 1, "W": 4, "E": 6, "NW": 0, "NE": 2, "SW": 0, "SE": 2}
    nb = neighbours(0, (4, 2), "torus")
    assert nb["N"] == 6 and nb["W"] == 1 and nb["NW"] == 7


def test_block_of_row_major():
    assert block_of(0, (2, 2), 2520, 1920) == (0, 0, 1260, 960)
    assert block_of(3, (2, 2), 2520, 1920) == (1260, 960, 1260, 960)
    assert block_of(5, (2, 4), 16384, 16384) == (8192, 4096, 8192, 4096)
