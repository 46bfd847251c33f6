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


# ---- one halo exchange per group of passes (halo_depth > tb), csrc/cuda/engine.cu make_passes

@pytest.mark.parametrize("boundary", ["zero", "torus"])
@pytest.mark.parametrize("grid", [(1, 2), (2, 2), (2, 4), (3, 3)])
@pytest.mark.parametrize("tb,halo_depth,reps", [(4, 8, 21), (4, 16, 37), (4, 32, 70), (2, 6, 13), (8, 16, 41), (1, 3, 7)])
def test_grouped_exchange(boundary, grid, tb, halo_depth, reps):
    ch = 3 if grid[1] != 3 else 1
    img = rnd(48 * grid[0], 64 * grid[1], ch, seed=halo_depth + reps)
    want = reference_oracle(img, reps, boundary)
    got = run_grid(img, grid, reps, boundary=boundary, tb=tb, halo_depth=halo_depth)
    assert (got == want).all(), f"{np.argwhere(got != want)[:4]}"


@pytest.mark.parametrize("kernel", ["stream", "tma"])
def test_grouped_exchange_unaligned_and_tma(kernel):
    # 75-byte block rows: the ring is pushed by halo_push_kernel after the last pass of a group
    img = rnd(96, 100, 3, seed=21)
    for boundary in ("zero", "torus"):
        want = reference_oracle(img, 29, boundary)
        assert (run_grid(img, (2, 4), 29, boundary=boundary, tb=4, halo_depth=12, kernel=kernel) == want).all()
    img = rnd(128, 512, 1, seed=22)
    want = reference_oracle(img, 50, "torus")
    assert (run_grid(img, (2, 2), 50, boundary="torus", tb=4, halo_depth=16, kernel=kernel, strip_bytes=16) == want).all()


def test_grouped_exchange_on_a_self_wrapping_torus():
    import b200conv
    for hw, hd in (((64, 200), 8), ((150, 704), 32), ((96, 512), 16)):
        img = rnd(hw[0], hw[1], 3, seed=hw[1])
        out = b200conv.blur(torch.from_numpy(img).cuda(), 45, boundary="torus", tb=4, halo_depth=hd).cpu().numpy()
        assert (out == reference_oracle(img, 45, "torus")).all(), (hw, hd)


@pytest.mark.parametrize("case", [
    # (H, W, ch, grid, boundary, opts): last chunk shorter than the halo depth unless the plan merges it
    (57, 512, 1, (1, 1), "torus", dict(tb=4, chunk_rows=8)),
    (57 * 2, 512, 1, (2, 1), "zero", dict(tb=4, chunk_rows=8)),
    (2520, 1920, 1, (8, 1), "zero", dict(tb=4)),             # 315 x 1920 blocks
    (315 * 2, 960 * 2, 1, (2, 2), "torus", dict(tb=4, chunk_rows=24)),
    (130, 512, 3, (2, 1), "torus", dict(tb=4, chunk_rows=16, halo_depth=8)),
])
def test_short_last_chunk_keeps_the_whole_ring(case):
    h, w, ch, grid, boundary, opts = case
    img = rnd(h, w, ch, seed=h)
    want = reference_oracle(img, 13, boundary)
    if grid == (1, 1):
        import b200conv
        got = b200conv.blur(torch.from_numpy(img).cuda(), 13, boundary=boundary, **opts).cpu().numpy()
    else:
        got = run_grid(img, grid, 13, boundary=boundary, **opts)
    assert (got == want).all(), f"{np.argwhere(got != want)[:4]}"


def test_runs_with_neighbours_replay_as_one_graph():
    """Epochs come from a device word, so a run with neighbours is the same launch sequence every
    time: the second run of the same length is ONE graph launch (plus nothing else)."""
    import b200conv
    img = rnd(256, 1024, 3, seed=31)
    eng = b200conv.Engine(1024, 256, 3, boundary="torus", tb=4, halo_depth=8, use_graph=1)
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        eng.load(torch.from_numpy(img).cuda())
        eng.exchange()
        eng.run(40)
        l0 = eng.launches
        eng.run(40)
        assert eng.launches - l0 == 1
        eng.run(40)
        out = torch.empty(img.shape, dtype=torch.uint8, device="cuda")
        eng.store(out)
    st.synchronize()
    assert (out.cpu().numpy() == reference_oracle(img, 120, "torus")).all()
    eng.close()


def test_halo_profile_counts_waits():
    import b200conv
    img = rnd(128, 512, 3, seed=32)
    eng = b200conv.Engine(512, 128, 3, boundary="torus", tb=4, kernel="stream")
    eng.set_profile(True)
    eng.load(torch.from_numpy(img).cuda())
    eng.exchange()
    eng.run(16)
    prof = eng.halo_profile()
    assert prof["waits"] > 0 and prof["publishes"] > 0
    eng.close()
