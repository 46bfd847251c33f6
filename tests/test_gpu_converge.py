"""T4: fused convergence check — early exit must give the same bytes as the full run."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from b200conv import blur, reference_oracle


def rnd(h, w, ch, seed=0):
    return np.random.default_rng(seed).integers(0, 256, (h, w) if ch == 1 else (h, w, ch), dtype=np.uint8)


def test_constant_torus_converges_at_first_check():
    img = np.full((64, 64, 3), 95, dtype=np.uint8)
    out, iters = blur(torch.from_numpy(img).cuda(), 1000, boundary="torus", check_every=20, return_iterations=True)
    assert iters == 20
    assert (out.cpu().numpy() == 95).all()


@pytest.mark.parametrize("boundary", ["zero", "torus"])
@pytest.mark.parametrize("tb", [1, 4])
def test_random_image_early_exit_equals_full_run(boundary, tb):
    img = rnd(128, 128, 1, seed=3)
    want = reference_oracle(img, 1000, boundary)
    out, iters = blur(torch.from_numpy(img).cuda(), 1000, boundary=boundary, check_every=20, tb=tb,
                      return_iterations=True)
    assert iters < 1000 and iters % 20 == 0
    assert (out.cpu().numpy() == want).all()
    # the detected fixed point really is one
    assert (reference_oracle(want, 1, boundary) == want).all()


def test_check_without_convergence_runs_everything():
    img = rnd(96, 200, 3, seed=4)
    out, iters = blur(torch.from_numpy(img).cuda(), 30, check_every=10, return_iterations=True)
    assert iters == 30
    assert (out.cpu().numpy() == reference_oracle(img, 30, "zero")).all()


def test_naive_kernel_ignores_check_every():
    # the naive kernel has no change detector: check_every must be ignored, not read as "nothing changed"
    img = rnd(64, 96, 1, seed=5)
    out, iters = blur(torch.from_numpy(img).cuda(), 12, kernel="naive", check_every=4, return_iterations=True)
    assert iters == 12
    assert (out.cpu().numpy() == reference_oracle(img, 12, "zero")).all()


def test_store_settles_the_convergence_state_by_itself():
    # a caller that never asks for the iteration count must still read the converged buffer, and a
    # second run on the same engine must continue from it
    from b200conv import Engine
    img = rnd(128, 128, 1, seed=3)
    want = reference_oracle(img, 1000, "torus")          # fixed point, reached long before 1000
    eng = Engine(128, 128, 1, boundary="torus", check_every=20, tb=4)
    src = torch.from_numpy(img).cuda()
    out = torch.empty_like(src)
    eng.load(src)
    eng.exchange()                                        # torus: the rank is its own neighbour
    eng.run(999)                                          # odd pass count: buffer index matters
    eng.store(out)
    torch.cuda.synchronize()
    assert (out.cpu().numpy() == want).all()
    eng.run(7)
    eng.store(out)
    torch.cuda.synchronize()
    assert (out.cpu().numpy() == want).all()
    eng.close()
