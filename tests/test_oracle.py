"""T1: the C++ oracle against an independent numpy restatement, the sharded host emulation
against the 1-rank result, and the cuda-compat model's observable properties."""
import numpy as np
import pytest

from b200conv import numpy_oracle, reference_oracle, sharded_oracle
from b200conv.models import FILTERS


def rnd(h, w, ch, seed=0):
    return np.random.default_rng(seed).integers(0, 256, (h, w) if ch == 1 else (h, w, ch), dtype=np.uint8)


@pytest.mark.parametrize("boundary", ["zero", "torus"])
@pytest.mark.parametrize("ch", [1, 3])
@pytest.mark.parametrize("hw", [(1, 1), (2, 3), (3, 3), (17, 5), (96, 64)])
@pytest.mark.parametrize("reps", [0, 1, 2, 7])
def test_oracle_equals_numpy(boundary, ch, hw, reps):
    img = rnd(hw[0], hw[1], ch, seed=reps)
    assert (reference_oracle(img, reps, boundary) == numpy_oracle(img, reps, boundary)).all()


def test_reps_zero_is_identity():
    img = rnd(9, 7, 3)
    for b in ("zero", "torus", "cuda-compat"):
        assert (reference_oracle(img, 0, b) == img).all()


@pytest.mark.parametrize("boundary", ["zero", "torus"])
@pytest.mark.parametrize("grid", [(1, 2), (2, 1), (2, 2), (2, 4), (4, 2), (3, 3)])
@pytest.mark.parametrize("tb", [1, 2, 4])
def test_sharded_emulation_is_decomposition_independent(boundary, grid, tb):
    for ch in (1, 3):
        img = rnd(24, 36, ch, seed=grid[0] * 10 + grid[1])
        want = reference_oracle(img, 9, boundary)
        assert (sharded_oracle(img, 9, grid, boundary, tb=tb) == want).all()


def test_torus_constant_image_is_fixed_point():
    img = np.full((8, 8, 3), 95, dtype=np.uint8)
    assert (reference_oracle(img, 3, "torus") == 95).all()


def test_small_random_images_converge():
    # floor() makes the filter dissipative
    img = rnd(32, 32, 1, seed=3)
    z = reference_oracle(img, 400, "zero")
    assert (z == 0).all()
    t = reference_oracle(img, 300, "torus")
    assert len(np.unique(t)) == 1


def test_cuda_compat_grey_properties():
    img = rnd(20, 16, 1, seed=5)
    assert (reference_oracle(img, 1, "cuda-compat") == img).all()  # reps=1 returns the input (:93-97)
    r2, r3 = reference_oracle(img, 2, "cuda-compat"), reference_oracle(img, 3, "cuda-compat")
    assert (r2 == r3).all()  # 2*floor(reps/2) effective iterations
    assert (r2[0] == img[0]).all() and (r2[:, 0] == img[:, 0]).all()  # border frozen at the original
    # explicit model: iteration 1 reads A (orig border), iteration 2 reads B (zero border)
    x = img.astype(np.int32)
    def it(x, border):
        p = x.copy()
        p[0], p[-1], p[:, 0], p[:, -1] = border[0], border[-1], border[:, 0], border[:, -1]
        q = np.pad(p, 1)
        h = q[:, :-2] + 2 * q[:, 1:-1] + q[:, 2:]
        return (h[:-2] + 2 * h[1:-1] + h[2:]) >> 4
    a = it(x, x)
    b = it(a, np.zeros_like(x))
    b[0], b[-1], b[:, 0], b[:, -1] = x[0], x[-1], x[:, 0], x[:, -1]
    assert (r2 == b.astype(np.uint8)).all()


def test_cuda_compat_rgb_linear_wrap_model():
    img = rnd(10, 7, 3, seed=6)
    r2 = reference_oracle(img, 2, "cuda-compat")
    assert (r2[0] == img[0]).all() and (r2[1, 0] == img[1, 0]).all()  # row 0 and pixel (1,0) never written
    assert (r2[1:, 0][1:] != img[1:, 0][1:]).any()  # left border column IS rewritten (wraps to previous row)
    assert (reference_oracle(img, 1, "cuda-compat") == img).all()


@pytest.mark.parametrize("name", ["box", "edge"])
def test_float_filters_follow_reference_op_order(name):
    f = FILTERS[name]
    img = rnd(11, 13, 3, seed=8)
    got = reference_oracle(img, 2, "zero", filter=name)
    w = f.as_float32()
    x = img
    for _ in range(2):
        p = np.pad(x, ((1, 1), (1, 1), (0, 0))).astype(np.float32)
        acc = np.zeros(x.shape, dtype=np.float32)
        for i in range(3):
            for j in range(3):
                acc = (acc + (p[i:i + x.shape[0], j:j + x.shape[1]] * w[i, j]).astype(np.float32)).astype(np.float32)
        x = acc.astype(np.uint8)
    assert (got == x).all()


@pytest.mark.parametrize("boundary", ["zero", "torus", "cuda-compat"])
@pytest.mark.parametrize("threads", [1, 3, 0])
def test_threaded_host_blur_equals_oracle(boundary, threads):
    from b200conv import blur_cpu
    rng = np.random.default_rng(12)
    for shape in ((1, 1), (2, 5), (33, 47), (20, 31, 3), (3, 3, 3)):
        img = rng.integers(0, 256, shape, dtype=np.uint8)
        for reps in (0, 1, 5):
            assert (blur_cpu(img, reps, boundary, threads=threads) == reference_oracle(img, reps, boundary)).all()
    if boundary != "cuda-compat":
        img = rng.integers(0, 256, (19, 23), dtype=np.uint8)
        for filt in ("box", "edge"):
            assert (blur_cpu(img, 3, boundary, filter=filt, threads=threads) ==
                    reference_oracle(img, 3, boundary, filter=filt)).all()
