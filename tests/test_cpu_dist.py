"""Multi-process tier on CPUs: torch.distributed with the gloo backend, world_size 2 and 4.

The host-only arm of the package (GlooHaloBlur: 8-neighbour send/recv halo exchange + threaded
host stencil — ranks x threads, the shape of the reference's MPI and MPI+OpenMP programs,
mpi_convolution.c:156-240, omp_convolution.c:292,297) must give the 1-rank oracle's bytes on every
grid, including the 2-wide torus where two messages travel between the same pair of ranks."""
import socket
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest
import torch

from b200conv import reference_oracle
from b200conv.ops import stencil_step

ROOT = Path(__file__).resolve().parent.parent


def free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("world", [2, 4])
def test_gloo_halo_exchange_equals_one_rank_oracle(world):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(free_port()), str(ROOT / "tests" / "cpu_dist_worker.py")]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    lines = [l for l in p.stdout.splitlines() if l.startswith("case")]
    assert p.returncode == 0, (p.stdout[-2000:], p.stderr[-3000:])
    assert len(lines) >= 3 and all(l.endswith("OK") for l in lines), lines


@pytest.mark.parametrize("filt", ["gaussian", "box", "edge"])
@pytest.mark.parametrize("threads", [1, 3, 0])
def test_host_stencil_step_equals_oracle(filt, threads):
    rng = np.random.default_rng(2)
    for ch, (h, w) in ((1, (17, 33)), (3, (9, 20))):
        img = rng.integers(0, 256, (h, w) if ch == 1 else (h, w, ch), dtype=np.uint8)
        bw = w * ch
        pitch = bw + 2 * ch + 5
        src = torch.zeros((h + 2, pitch), dtype=torch.uint8)
        dst = torch.full((h + 2, pitch), 7, dtype=torch.uint8)
        src[1:h + 1, ch:ch + bw] = torch.from_numpy(img.reshape(h, bw))
        stencil_step(src, dst, h, bw, ch, filter=filt, threads=threads)
        got = dst[1:h + 1, ch:ch + bw].numpy().reshape(img.shape)
        assert (got == reference_oracle(img, 1, "zero", filter=filt)).all()
        assert (dst[0] == 7).all() and (dst[:, :ch] == 7).all()      # frame of dst untouched
