"""Property tests (hypothesis) of the pure CPU layer: decomposition, neighbour tables, raw block I/O
and the oracles.  They restate the contracts of SURVEY §2.2 (A6, A7, A12, A11/A25) as invariants
instead of hand-picked examples."""
import numpy as np
from hypothesis import HealthCheck, given, settings
from hypothesis import strategies as st

import b200conv
from b200conv import numpy_oracle, reference_oracle, sharded_oracle
from b200conv.parallel import block_of, neighbours, rows_division
from b200conv.utils.rawio import read_block as read_raw_block, write_block as write_raw_block

FAST = settings(max_examples=60, deadline=None, suppress_health_check=[HealthCheck.function_scoped_fixture])
OPP = {"N": "S", "S": "N", "W": "E", "E": "W", "NW": "SE", "SE": "NW", "NE": "SW", "SW": "NE"}


@FAST
@given(n=st.integers(1, 64), h=st.integers(1, 400), w=st.integers(1, 400))
def test_rows_division_is_the_cheapest_exact_grid(n, h, w):
    r = rows_division(n, h, w)
    grids = [(p, n // p) for p in range(1, n + 1) if n % p == 0 and h % p == 0 and w % (n // p) == 0]
    if not grids:
        assert r == 0
        return
    cost = lambda g: h // g[0] + w // g[1]       # block semi-perimeter, mpi_convolution.c:350-364
    best = min(cost(g) for g in grids)
    assert r == min(p for p, q in grids if cost((p, q)) == best)   # ties keep the smallest row count


@FAST
@given(p=st.integers(1, 6), q=st.integers(1, 6), boundary=st.sampled_from(["zero", "torus"]))
def test_neighbour_tables_are_symmetric(p, q, boundary):
    for rank in range(p * q):
        nb = neighbours(rank, (p, q), boundary)
        for d, peer in nb.items():
            if boundary == "torus":
                assert 0 <= peer < p * q
            if peer >= 0:
                assert neighbours(peer, (p, q), boundary)[OPP[d]] == rank
            else:
                r, c = divmod(rank, q)
                assert (("N" in d and r == 0) or ("S" in d and r == p - 1) or
                        ("W" in d and c == 0) or ("E" in d and c == q - 1))


@FAST
@given(p=st.integers(1, 5), q=st.integers(1, 5), bh=st.integers(1, 9), bw=st.integers(1, 9))
def test_blocks_tile_the_image_exactly_once(p, q, bh, bw):
    h, w = p * bh, q * bw
    seen = np.zeros((h, w), dtype=np.int32)
    for rank in range(p * q):
        r0, c0, r, c = block_of(rank, (p, q), h, w)
        assert (r, c) == (bh, bw)
        seen[r0:r0 + r, c0:c0 + c] += 1
    assert (seen == 1).all()


@FAST
@given(data=st.data(), ch=st.sampled_from([1, 3]), h=st.integers(1, 24), w=st.integers(1, 24))
def test_raw_block_io_round_trip(tmp_path, data, ch, h, w):
    img = np.frombuffer(data.draw(st.binary(min_size=h * w * ch, max_size=h * w * ch)), dtype=np.uint8)
    img = img.reshape((h, w) if ch == 1 else (h, w, ch))
    path = tmp_path / f"img_{h}_{w}_{ch}.raw"
    img.tofile(path)
    r0 = data.draw(st.integers(0, h - 1))
    c0 = data.draw(st.integers(0, w - 1))
    r = data.draw(st.integers(1, h - r0))
    c = data.draw(st.integers(1, w - c0))
    blk = read_raw_block(path, w, h, ch, r0, c0, r, c)
    assert (blk == img[r0:r0 + r, c0:c0 + c]).all()
    out = tmp_path / f"out_{h}_{w}_{ch}.raw"
    np.zeros_like(img).tofile(out)
    write_raw_block(out, blk, w, h, ch, r0, c0)
    back = np.fromfile(out, dtype=np.uint8).reshape(img.shape)
    want = np.zeros_like(img)
    want[r0:r0 + r, c0:c0 + c] = blk
    assert (back == want).all()


@settings(max_examples=40, deadline=None)
@given(data=st.data(), ch=st.sampled_from([1, 3]), h=st.integers(1, 14), w=st.integers(1, 14),
       reps=st.integers(0, 6), boundary=st.sampled_from(["zero", "torus"]))
def test_cpp_oracle_equals_numpy_restatement(data, ch, h, w, reps, boundary):
    img = np.frombuffer(data.draw(st.binary(min_size=h * w * ch, max_size=h * w * ch)), dtype=np.uint8)
    img = img.reshape((h, w) if ch == 1 else (h, w, ch))
    assert (reference_oracle(img, reps, boundary) == numpy_oracle(img, reps, boundary)).all()


@settings(max_examples=40, deadline=None)
@given(data=st.data(), ch=st.sampled_from([1, 3]), p=st.integers(1, 3), q=st.integers(1, 3), bh=st.integers(1, 6),
       bw=st.integers(1, 6), reps=st.integers(0, 9), tb=st.sampled_from([1, 2, 4, 8]),
       boundary=st.sampled_from(["zero", "torus"]))
def test_every_decomposition_gives_the_one_rank_result(data, ch, p, q, bh, bw, reps, tb, boundary):
    h, w = p * bh, q * bw
    img = np.frombuffer(data.draw(st.binary(min_size=h * w * ch, max_size=h * w * ch)), dtype=np.uint8)
    img = img.reshape((h, w) if ch == 1 else (h, w, ch))
    assert (sharded_oracle(img, reps, (p, q), boundary, tb=tb) == reference_oracle(img, reps, boundary)).all()


def test_iterations_compose_in_zero_mode():
    # "re-running on blur_x.raw continues blurring" (checkpoint/resume): f^a(f^b(x)) = f^(a+b)(x)
    img = np.random.default_rng(0).integers(0, 256, (19, 23, 3), dtype=np.uint8)
    for boundary in ("zero", "torus"):
        assert (reference_oracle(reference_oracle(img, 3, boundary), 4, boundary) == reference_oracle(img, 7, boundary)).all()
    assert b200conv.__version__
