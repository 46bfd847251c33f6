"""T3 on one GPU: every rank of a grid as its own engine in this process, stepped pass by
pass (parallel/local.py).  The result must equal the 1-rank oracle byte for byte on every
grid — decomposition independence, which the reference does not have."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from b200conv import LocalGrid, reference_oracle


def rnd(h, w, ch, seed=0):
    return np.random.default_rng(seed).integers(0, 256, (h, w) if ch == 1 else (h, w, ch), dtype=np.uint8)


def run_grid(img, grid, reps, **kw):
    ch = 1 if img.ndim == 2 else 3
    g = LocalGrid(img.shape[1], img.shape[0], ch, grid, **kw)
    try:
        g.load(torch.from_numpy(img).cuda())
        g.run(reps)
        return g.gather().numpy()
    finally:
        g.close()


GRIDS = [(1, 2), (2, 1), (2, 2), (2, 4), (4, 2)]


@pytest.mark.parametrize("boundary", ["zero", "torus"])
@pytest.mark.parametrize("grid", GRIDS)
@pytest.mark.parametrize("ch", [1, 3])
def test_fused_push_aligned(boundary, grid, ch):
    # block byte width is a multiple of 16 -> ring is pushed from inside the stencil kernel
    h, w = 64 * grid[0], 128 * grid[1] * (1 if ch == 1 else 1)
    img = rnd(h, w * 16 // 16, ch, seed=grid[0] * 7 + grid[1])
    assert (img.shape[1] // grid[1] * ch) % 16 == 0
    want = reference_oracle(img, 11, boundary)
    for tb in (1, 4):
        got = run_grid(img, grid, 11, boundary=boundary, tb=tb)
        assert (got == want).all(), f"tb={tb}: {np.argwhere(got != want)[:4]}"


@pytest.mark.parametrize("boundary", ["zero", "torus"])
@pytest.mark.parametrize("grid", [(2, 2), (2, 4)])
def test_unaligned_fallback_push(boundary, grid):
    img = rnd(36 * grid[0], 25 * grid[1], 3, seed=5)  # 75-byte block rows
    want = reference_oracle(img, 9, boundary)
    for tb in (1, 2, 4):
        assert (run_grid(img, grid, 9, boundary=boundary, tb=tb) == want).all()


@pytest.mark.parametrize("grid", [(2, 2), (4, 2)])
def test_naive_kernel_with_push(grid):
    img = rnd(32 * grid[0], 48 * grid[1], 3, seed=6)
    for boundary in ("zero", "torus"):
        assert (run_grid(img, grid, 5, boundary=boundary, kernel="naive") == reference_oracle(img, 5, boundary)).all()


def test_plumbing_config_shape():
    # BASELINE.json config 1: 1920x2520 rgb, 40 reps, 4 ranks (2x2 grid, 1260x960 blocks)
    from b200conv.utils import waterfall_image
    img = waterfall_image(2520, 1920, 3, seed=0)
    want = reference_oracle(img, 40, "zero")
    assert (run_grid(img, (2, 2), 40, boundary="zero") == want).all()


def test_long_run_many_epochs():
    img = rnd(128, 256, 1, seed=12)
    want = reference_oracle(img, 203, "torus")
    assert (run_grid(img, (2, 2), 203, boundary="torus", tb=4, strip_bytes=4) == want).all()


@pytest.mark.parametrize("boundary", ["zero", "torus"])
@pytest.mark.parametrize("grid,hw", [((2, 2), (8, 8)), ((4, 2), (8, 6)), ((2, 4), (4, 16)), ((1, 2), (5, 32)),
                                     ((2, 1), (6, 7)), ((3, 3), (9, 9))])
def test_tiny_blocks_clamp_the_halo_depth(boundary, grid, hw):
    """Blocks smaller than the requested fusion depth: a k-deep halo can only come from the
    adjacent rank, so the engine lowers k (mpi_convolution.c has k=1 throughout)."""
    for ch in (1, 3):
        img = rnd(hw[0], hw[1], ch, seed=hw[0] + hw[1])
        want = reference_oracle(img, 7, boundary)
        assert (run_grid(img, grid, 7, boundary=boundary, tb=8) == want).all()


def test_torus_one_rank_wraps_onto_itself():
    import b200conv
    for hw in ((4, 4), (9, 33), (64, 200)):
        img = rnd(hw[0], hw[1], 3, seed=hw[1])
        out = b200conv.blur(torch.from_numpy(img).cuda(), 9, boundary="torus", tb=4).cpu().numpy()
        assert (out == reference_oracle(img, 9, "torus")).all()
