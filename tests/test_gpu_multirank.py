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
