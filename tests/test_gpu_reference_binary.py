"""T5: the reference's actual CUDA binary (baseline/_ref/cuda, built unmodified with its own
Makefile) against this engine on the same GPU.

* grey: `cuda_convolution` == our `cuda-compat` mode byte for byte (frozen border, and the
  odd-repetition bug: odd counts return the result of reps-1 iterations, reps=1 the input).
* rgb: the reference kernel's guard (`y < 3W-3`, cuda_convolution.cu:34) reads and writes past
  its allocations, so its last rows depend on whatever follows the buffers; everything except the
  last two rows must still match the linear-wrap model, and the whole image usually does.
"""
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from b200conv import blur, reference_oracle
from b200conv.utils import read_raw, write_raw

ROOT = Path(__file__).resolve().parent.parent
EXE = ROOT / "baseline" / "_ref" / "cuda" / "cuda_convolution"


@pytest.fixture(scope="module")
def ref_cuda():
    if not EXE.exists():
        subprocess.run([sys.executable, str(ROOT / "baseline" / "install_ref.py")], capture_output=True, check=False)
    if not EXE.exists():
        pytest.skip("reference CUDA program could not be built here")
    return EXE


def run_ref(exe, tmp_path, img, reps):
    ch = 1 if img.ndim == 2 else 3
    write_raw(tmp_path / "in.raw", img)
    out = tmp_path / "blur_in.raw"
    if out.exists():
        out.unlink()
    p = subprocess.run([str(exe), "in.raw", str(img.shape[1]), str(img.shape[0]), str(reps), "grey" if ch == 1 else "rgb"],
                       cwd=tmp_path, capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stderr
    assert p.stdout.startswith("Execution time: ")
    return read_raw(out, img.shape[1], img.shape[0], ch)


@pytest.mark.parametrize("reps", [1, 2, 3, 10])
def test_grey_binary_equals_cuda_compat_mode(ref_cuda, tmp_path, reps):
    img = np.random.default_rng(reps).integers(0, 256, (75, 133), dtype=np.uint8)
    ref = run_ref(ref_cuda, tmp_path, img, reps)
    ours = blur(torch.from_numpy(img).cuda(), reps, boundary="cuda-compat").cpu().numpy()
    assert (ours == ref).all()
    assert (reference_oracle(img, reps, "cuda-compat") == ref).all()
    if reps == 1:
        assert (ref == img).all()


def test_rgb_binary_matches_linear_wrap_model(ref_cuda, tmp_path):
    # Measured on B200: cudaMalloc places the reference's two 9216-byte buffers back to back, so
    # its out-of-bounds rows H, H+1 of one buffer ARE rows 0, 1 of the other (and it races on them
    # inside one launch).  Junk therefore creeps in from the bottom AND the top, one row per
    # iteration; rows farther than that from both ends must equal the linear-wrap model.
    img = np.random.default_rng(7).integers(0, 256, (64, 48, 3), dtype=np.uint8)
    for reps in (2, 6):
        ref = run_ref(ref_cuda, tmp_path, img, reps)
        ours = blur(torch.from_numpy(img).cuda(), reps, boundary="cuda-compat").cpu().numpy()
        lo, hi = reps + 2, img.shape[0] - 2 - reps
        assert (ours[lo:hi] == ref[lo:hi]).all()
        assert (ours == reference_oracle(img, reps, "cuda-compat")).all()


def test_interior_of_binary_equals_default_mode_far_from_border(ref_cuda, tmp_path):
    """Away from the border the reference binary and the default (zero) mode agree: same filter."""
    img = np.random.default_rng(9).integers(0, 256, (80, 90), dtype=np.uint8)
    reps = 4
    ref = run_ref(ref_cuda, tmp_path, img, reps)
    ours = blur(torch.from_numpy(img).cuda(), reps).cpu().numpy()
    m = reps + 1
    assert (ours[m:-m, m:-m] == ref[m:-m, m:-m]).all()
