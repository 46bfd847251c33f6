"""kernel="resident": the register-resident persistent kernel for images that fit on chip
(csrc/cuda/stencil_resident.cu).  Bytes against the CPU oracle on small shapes (every exchange
interval, both boundaries, grey and rgb, partial last tiles), against the streaming kernel on the
1920x2520 configuration of BASELINE.json, and two runs in a row on one engine."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from b200conv import B2CError, Engine, blur, reference_oracle


def rnd(h, w, ch, seed=0):
    return np.random.default_rng(seed).integers(0, 256, (h, w) if ch == 1 else (h, w, ch), dtype=np.uint8)


@pytest.mark.parametrize("case", [
    (40, 64, 1, 9, "zero", 0), (40, 64, 1, 9, "torus", 4), (130, 488, 1, 13, "torus", 4), (130, 488, 1, 13, "zero", 8),
    (200, 96, 3, 11, "torus", 4), (77, 160, 3, 5, "zero", 0), (61, 8, 1, 7, "torus", 8), (4, 16, 1, 3, "zero", 0),
    (300, 520, 1, 1, "zero", 0), (300, 520, 1, 2, "torus", 0), (300, 520, 1, 3, "zero", 4),
    (130, 488, 1, 37, "torus", 16), (200, 96, 3, 35, "zero", 16), (90, 304, 1, 16, "zero", 16),
])
def test_resident_equals_oracle(case):
    h, w, ch, reps, boundary, tb = case
    img = rnd(h, w, ch, 1)
    got = blur(torch.from_numpy(img).cuda(), reps, boundary=boundary, kernel="resident", tb=tb).cpu().numpy()
    assert (got == reference_oracle(img, reps, boundary)).all()


@pytest.mark.parametrize("case", [(2520, 1920, 1, 1000, "zero", 0), (2520, 1920, 1, 999, "torus", 8),
                                  (1000, 1800, 3, 250, "zero", 0)])
def test_resident_equals_streaming_kernel_on_the_small_configs(case):
    h, w, ch, reps, boundary, tb = case
    img = torch.from_numpy(rnd(h, w, ch, 2)).cuda()
    got = blur(img, reps, boundary=boundary, kernel="resident", tb=tb)
    assert torch.equal(got, blur(img, reps, boundary=boundary, kernel="stream"))


def test_two_runs_on_one_engine_and_plan_report():
    eng = Engine(1920, 2520, 1, kernel="resident")
    assert "kernel=resident" in eng.describe()
    img = torch.from_numpy(rnd(2520, 1920, 1, 3)).cuda()
    out = torch.empty_like(img)
    eng.load(img)
    eng.run(10)
    eng.run(7)              # epochs of the exchange flags continue across launches
    eng.store(out)
    torch.cuda.synchronize()
    assert eng.iterations_done() == 7
    assert torch.equal(out, blur(img, 17, kernel="stream"))
    eng.close()


def test_images_that_do_not_fit_are_refused():
    with pytest.raises(B2CError, match="resident"):
        Engine(4096, 4096, 3, kernel="resident")        # 50 MB: far beyond the register files
    with pytest.raises(B2CError, match="resident"):
        Engine(100, 64, 1, kernel="resident")           # row bytes not a multiple of 8
