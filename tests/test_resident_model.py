"""The register-resident small-image kernel (csrc/cuda/stencil_resident.cu) without a GPU:
scripts/resident_model.py repeats the kernel statement by statement on [cta, warp, lane] arrays
(same tile plan, ring/halo classification, SHFL edge behaviour, scale alternation, in-place row
order, exchange images) and must reproduce the CPU oracle."""
import importlib.util
from pathlib import Path

import numpy as np
import pytest

from b200conv import reference_oracle

ROOT = Path(__file__).resolve().parent.parent
spec = importlib.util.spec_from_file_location("resident_model", ROOT / "scripts" / "resident_model.py")
model = importlib.util.module_from_spec(spec)
spec.loader.exec_module(model)


@pytest.mark.parametrize("case", [
    (40, 64, 1, 9, "zero", 148), (40, 64, 1, 9, "torus", 148), (130, 488, 1, 13, "torus", 148),
    (200, 96, 3, 11, "torus", 148), (77, 160, 3, 5, "zero", 148), (300, 520, 1, 6, "torus", 12),
    (61, 8, 1, 7, "torus", 148), (4, 16, 1, 3, "zero", 148), (90, 304, 1, 19, "torus", 148),
    (20, 248, 1, 9, "torus", 148), (60, 240, 3, 9, "torus", 148),   # last strip exactly one halo wide
])
@pytest.mark.parametrize("k", [4, 8, 16])
def test_lane_level_model_equals_the_oracle(case, k):
    h, w, ch, reps, boundary, sms = case
    model.K = k                                          # iterations between two exchanges
    if model.plan(w * ch, h, ch, sms) is None:
        pytest.skip("shape does not fit with this exchange interval")
    img = np.random.default_rng(5).integers(0, 256, (h, w) if ch == 1 else (h, w, ch), dtype=np.uint8)
    model.run.ch = ch
    got, _ = model.run(img.reshape(h, w * ch), reps, boundary, sms)
    assert (got == reference_oracle(img, reps, boundary).reshape(h, w * ch)).all()


def test_plan_limits():
    model.K = 4
    assert model.plan(1920, 2520, 1) == dict(rw=10, strips=8, ctarows=17, vr=152, vc=240)
    assert model.plan(1924, 2520, 1) is None            # row bytes must be a multiple of 8
    assert model.plan(16384, 16384, 1) is None          # far too large for the register files
    assert model.plan(5760, 2520, 3) is None            # 1920x2520 rgb does not fit either
    assert model.plan(5400, 1000, 3)["rw"] == 14
    assert model.plan(232, 64, 3) is None               # rgb: last strip (8 B) narrower than the 16 B halo


def test_cpp_planner_agrees_with_the_model():
    # the C++ planner (stencil_resident.cu, host code) and the model's plan() must be the same function
    import ctypes as C
    from b200conv._lib import lib
    rng = np.random.default_rng(8)
    for _ in range(300):
        bw, h = int(rng.integers(1, 900)) * 8, int(rng.integers(1, 4000))
        ch, k, sms = int(rng.choice([1, 3])), int(rng.choice([4, 8, 16])), int(rng.choice([12, 64, 148]))
        out = (C.c_int * 5)()
        ok = lib.b2c_resident_plan(bw, h, ch, k, sms, out)
        model.K = k
        want = model.plan(bw, h, ch, sms)
        assert bool(ok) == (want is not None), (bw, h, ch, k, sms)
        if want:
            assert list(out) == [want["rw"], want["strips"], want["ctarows"], want["vr"], want["vc"]]
    model.K = 4
