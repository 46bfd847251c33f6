"""CUDA-event timing helpers (layer L8: the device-side counterpart of MPI_Wtime /
micro_time(), mpi_convolution.c:154,242 and cuda/functions.c:47-51)."""
from __future__ import annotations

import torch


def time_cuda(fn, warmup=3, iters=5):
    """-> list of per-call milliseconds, CUDA events on the current stream, sync on both sides."""
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    out = []
    for _ in range(iters):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        e1.synchronize()
        out.append(e0.elapsed_time(e1))
    return out


def gpix_it_per_s(width, height, repetitions, ms) -> float:
    """BASELINE.json metric: W*H*reps / seconds / 1e9."""
    return width * height * repetitions / (ms * 1e-3) / 1e9
