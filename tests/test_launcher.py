"""The native launcher `lib/b200conv`: reference CLI parity (usage text, exit status, output
name) on CPU, whole-program runs on GPU."""
import json
import subprocess

import numpy as np
import pytest

from b200conv import reference_oracle
from b200conv._lib import LAUNCHER_PATH
from b200conv.utils import read_raw, write_raw


def run(args, cwd):
    return subprocess.run([str(LAUNCHER_PATH), *map(str, args)], cwd=cwd, capture_output=True, text=True, timeout=300)


@pytest.fixture(scope="module")
def launcher():
    if not LAUNCHER_PATH.exists():
        pytest.skip("launcher not built")
    return LAUNCHER_PATH


def test_usage_error_like_the_reference(launcher, tmp_path):
    p = run(["x.raw", "10", "10"], tmp_path)
    assert p.returncode == 1
    assert p.stderr == f"\nError Input!\n{launcher} Image_name width height repetitions [rgb/grey].\n\n"
    p = run(["x.raw", "10", "10", "2", "gray"], tmp_path)
    assert p.returncode == 1 and "Error Input!" in p.stderr


def test_cannot_divide_and_missing_file(launcher, tmp_path):
    p = run(["x.raw", "10", "10", "2", "grey", "--gpus", "3"], tmp_path)
    assert p.returncode == 1 and "Cannot divide to processes" in p.stderr
    p = run(["nope.raw", "10", "10", "2", "grey", "--gpus", "1"], tmp_path)
    assert p.returncode == 1 and "cannot open nope.raw" in p.stderr


@pytest.mark.parametrize("color,boundary,cpu", [("grey", "zero", "1"), ("rgb", "zero", "3"), ("rgb", "torus", "all"),
                                                ("grey", "cuda-compat", "2")])
def test_whole_program_on_host_threads(launcher, tmp_path, color, boundary, cpu):
    # --cpu N|all: no CUDA call at all; same files, same two output lines as the GPU path
    ch = 1 if color == "grey" else 3
    img = np.random.default_rng(5).integers(0, 256, (37, 52) if ch == 1 else (37, 52, ch), dtype=np.uint8)
    write_raw(tmp_path / "in.raw", img)
    rep = tmp_path / "report.json"
    p = run(["in.raw", 52, 37, 6, color, "--cpu", cpu, "--boundary", boundary, "--json", rep], tmp_path)
    assert p.returncode == 0, p.stderr
    lines = p.stdout.splitlines()
    assert len(lines) == 2 and float(lines[0]) >= 0 and lines[1].startswith("Execution time: ")
    out = read_raw(tmp_path / "blur_in.raw", 52, 37, ch)
    assert (out == reference_oracle(img, 6, boundary)).all()
    d = json.loads(rep.read_text())
    assert d["device"] == "host" and d["repetitions"] == 6
    assert run(["in.raw", 52, 37, 6, color, "--cpu", "0"], tmp_path).returncode == 1      # 0 threads is not a count
    assert run(["nope.raw", 52, 37, 6, color, "--cpu", "1"], tmp_path).returncode == 1


@pytest.mark.gpu
@pytest.mark.parametrize("color,boundary", [("grey", "zero"), ("rgb", "zero"), ("rgb", "torus"), ("grey", "cuda-compat")])
def test_single_gpu_program(launcher, tmp_path, color, boundary):
    ch = 1 if color == "grey" else 3
    rng = np.random.default_rng(5)
    img = rng.integers(0, 256, (120, 96) if ch == 1 else (120, 96, 3), dtype=np.uint8)
    write_raw(tmp_path / "w.raw", img)
    p = run(["w.raw", 96, 120, 10, color, "--gpus", 1, "--boundary", boundary, "--json", "r.json", "--profile"], tmp_path)
    assert p.returncode == 0, p.stderr
    lines = p.stdout.strip().splitlines()
    float(lines[0])                                   # "%f\n": loop seconds, like mpi_convolution
    assert lines[1].startswith("Execution time: ") and lines[1].endswith(" sec")   # like cuda_convolution
    got = read_raw(tmp_path / "blur_w.raw", 96, 120, ch)
    assert (got == reference_oracle(img, 10, boundary)).all()
    rep = json.loads((tmp_path / "r.json").read_text())
    assert rep["gpus"] == 1 and rep["iterations_done"] == 10 and rep["gpix_it_per_s"] > 0
    assert "rank 0:" in p.stderr and "loop=" in p.stderr


@pytest.mark.gpu
def test_checkpoints_and_convergence(launcher, tmp_path):
    img = np.full((64, 64), 77, dtype=np.uint8)
    write_raw(tmp_path / "c.raw", img)
    p = run(["c.raw", 64, 64, 1000, "grey", "--gpus", 1, "--boundary", "torus", "--check-every", 20,
             "--checkpoint-every", 100, "--json", "r.json"], tmp_path)
    assert p.returncode == 0, p.stderr
    assert json.loads((tmp_path / "r.json").read_text())["iterations_done"] == 20   # constant image: fixed point
    assert (read_raw(tmp_path / "blur_c.raw", 64, 64, 1) == 77).all()


@pytest.mark.gpu
def test_multi_gpu_program(launcher, tmp_path):
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    n = 8 if n >= 8 else (4 if n >= 4 else 2)
    rng = np.random.default_rng(6)
    img = rng.integers(0, 256, (256, 512, 3), dtype=np.uint8)
    write_raw(tmp_path / "m.raw", img)
    for boundary in ("zero", "torus"):
        p = run(["m.raw", 512, 256, 23, "rgb", "--gpus", n, "--boundary", boundary], tmp_path)
        assert p.returncode == 0, p.stderr
        float(p.stdout.strip().splitlines()[0])
        assert (read_raw(tmp_path / "blur_m.raw", 512, 256, 3) == reference_oracle(img, 23, boundary)).all()


@pytest.mark.gpu
def test_worker_failure_takes_the_job_down(launcher, tmp_path):
    """Failure detection: a worker that cannot read its block makes the launcher exit 1
    (the reference would loop forever on a short file, cuda/functions.c:39-45)."""
    (tmp_path / "short.raw").write_bytes(b"\x07" * 1000)
    p = run(["short.raw", 64, 64, 3, "grey", "--gpus", 1], tmp_path)
    assert p.returncode == 1
    assert "short read" in p.stderr and "a worker failed" in p.stderr
