"""Process-grid decomposition (layer L2).  Thin wrappers over the C++ core so Python and
the launcher share one implementation of ``RowsDivision`` (mpi/mpi_convolution.c:350-364),
the divisibility check (:53-58), block origins (:72-73,85-86) and the neighbour table
(:142-149; extended to 8 neighbours and an optional torus)."""
from __future__ import annotations

import ctypes as C

from .._lib import BOUNDARY, DIRS, B2CError, check, last_error, lib


def rows_division(n_processes: int, rows: int, columns: int) -> int:
    """Number of process-grid rows; 0 when the image cannot be divided."""
    return int(lib.b2c_rows_division(int(n_processes), int(rows), int(columns)))


def decompose(n: int, height: int, width: int):
    """-> (P, Q).  Raises ``B2CError('Cannot divide to processes')`` like the reference aborts."""
    p, q = C.c_int(), C.c_int()
    rc = lib.b2c_decompose(int(n), int(height), int(width), C.byref(p), C.byref(q))
    if rc < 0:
        raise B2CError(last_error())
    return p.value, q.value


def neighbours(rank: int, grid, boundary: str = "zero") -> dict:
    """-> {"N": rank|-1, "S": ..., "W", "E", "NW", "NE", "SW", "SE"}."""
    out = (C.c_int * 8)()
    check(lib.b2c_neighbours(int(rank), int(grid[0]), int(grid[1]), BOUNDARY[boundary], C.byref(out)), "neighbours")
    return {d: int(out[i]) for i, d in enumerate(DIRS)}


def block_of(rank: int, grid, height: int, width: int):
    """-> (row0, col0, rows, cols) of ``rank`` (row-major rank grid)."""
    p, q = int(grid[0]), int(grid[1])
    r, c = height // p, width // q
    return (rank // q) * r, (rank % q) * c, r, c
