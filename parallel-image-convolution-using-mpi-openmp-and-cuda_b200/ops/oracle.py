"""CPU oracles.

``reference_oracle``  — the C++ oracle in ``csrc/core/oracle.cpp`` (integer for the Gaussian,
                        fp32 reference op order for box/edge, zero/torus/cuda-compat boundaries).
``numpy_oracle``      — an independent numpy restatement of ``out = (sum w*p) >> 4`` used to
                        cross-check the C++ oracle (mpi/mpi_convolution.c:301-322 semantics at
                        one rank; torus per README.md:63-66).
``blur_cpu``          — the threaded, vectorised host path (what the launcher's ``--cpu N`` runs).
``sharded_oracle``    — host emulation of the multi-rank frame/epoch halo protocol.
``threaded_resident_model`` — one thread per CTA of the resident kernel (tile plan, ring/halo exchange images, epochs).
``threaded_halo_model`` — the same protocol with one thread per rank, atomics for the epoch flags
                        and the convergence slots, and injected delays (csrc/core/halo_emul.cpp).
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from .. import _lib
from .._lib import BOUNDARY, FILTER, check, lib


def _as_image(img):
    """-> (contiguous uint8 ndarray [H, W*CH], W, H, CH, restore) for ndarray or torch input."""
    torch_like = hasattr(img, "detach")
    a = img.detach().cpu().numpy() if torch_like else np.asarray(img)
    if a.dtype != np.uint8 or a.ndim not in (2, 3) or (a.ndim == 3 and a.shape[2] not in (1, 3)):
        raise ValueError("image must be uint8 [H,W] or [H,W,3]")
    ch = 1 if a.ndim == 2 else a.shape[2]
    h, w = a.shape[0], a.shape[1]
    flat = np.ascontiguousarray(a).reshape(h, w * ch)

    def restore(out):
        out = out.reshape(a.shape)
        if torch_like:
            import torch

            return torch.from_numpy(out)
        return out

    return flat, w, h, ch, restore


def reference_oracle(img, reps: int, boundary: str = "zero", filter: str = "gaussian"):
    flat, w, h, ch, restore = _as_image(img)
    out = np.empty_like(flat)
    check(lib.b2c_oracle_blur(flat.ctypes.data, out.ctypes.data, w, h, ch, int(reps), BOUNDARY[boundary],
                              FILTER[filter]), "oracle_blur")
    return restore(out)


def blur_cpu(img, reps: int, boundary: str = "zero", filter: str = "gaussian", threads: int = 0):
    """Whole blur on host threads (``b2c_blur_host``): the no-GPU path, exact, ndarray or CPU tensor in,
    same kind out.  ``threads`` = 0 uses every core."""
    flat, w, h, ch, restore = _as_image(img)
    out = np.empty_like(flat)
    check(lib.b2c_blur_host(flat.ctypes.data, out.ctypes.data, w, h, ch, int(reps), BOUNDARY[boundary],
                            FILTER[filter], int(threads)), "blur_host")
    return restore(out)


def sharded_oracle(img, reps: int, grid, boundary: str = "zero", tb: int = 1):
    flat, w, h, ch, restore = _as_image(img)
    out = np.empty_like(flat)
    check(lib.b2c_oracle_blur_sharded(flat.ctypes.data, out.ctypes.data, w, h, ch, int(reps), BOUNDARY[boundary],
                                      int(grid[0]), int(grid[1]), int(tb)), "oracle_blur_sharded")
    return restore(out)


def threaded_halo_model(img, reps: int, grid, boundary: str = "zero", tb: int = 1, check_every: int = 0,
                        max_delay_us: int = 0, seed: int = 0, timeout_ms: int = 0, single_slot_set: bool = False,
                        stall_rank: int = -1, stall_ms: int = 0, images: int = 1, halo_depth: int = 0,
                        courier: bool = False):
    """-> (image, iterations run); with ``images`` > 1 a list of results (image j = input XOR 37*j, pushed
    through the same buffers back to back with no barrier in between, like ``ShardedPipeline``).
    Raises B2CError (code B2C_ETIMEOUT) when a rank deadlocks."""
    flat, w, h, ch, restore = _as_image(img)
    n = max(1, int(images))
    out = np.empty((n,) + flat.shape, dtype=np.uint8)
    o = _lib.EmulOpts(int(tb), int(check_every), int(max_delay_us), int(seed), int(timeout_ms),
                      int(bool(single_slot_set)), int(stall_rank), int(stall_ms), int(images), int(halo_depth),
                      int(bool(courier)))
    iters = C.c_int(0)
    check(lib.b2c_emulate_halo_threads(flat.ctypes.data, out.ctypes.data, w, h, ch, int(reps), BOUNDARY[boundary],
                                       int(grid[0]), int(grid[1]), C.byref(o), C.byref(iters)),
          "emulate_halo_threads")
    if n == 1:
        return restore(out[0]), iters.value
    return [restore(out[j]) for j in range(n)], iters.value


def threaded_resident_model(img, reps: int, boundary: str = "zero", k: int = 4, sms: int = 148,
                            max_delay_us: int = 0, seed: int = 0):
    """Host model of the resident kernel's CTA protocol (csrc/core/resident_emul.cpp): one thread per CTA."""
    flat, w, h, ch, restore = _as_image(img)
    out = np.empty_like(flat)
    check(lib.b2c_emulate_resident_threads(flat.ctypes.data, out.ctypes.data, w, h, ch, int(reps), BOUNDARY[boundary],
                                           int(k), int(sms), int(max_delay_us), int(seed)), "emulate_resident_threads")
    return restore(out)


def threaded_persist_model(img, reps: int, k: int = 4, vw: int = 480, chunk_rows: int = 32, workers: int = 8,
                           max_delay_us: int = 0, seed: int = 0):
    """Host model of the persistent kernel's work queue and tile dependencies (csrc/core/persist_emul.cpp)."""
    flat, w, h, ch, restore = _as_image(img)
    out = np.empty_like(flat)
    check(lib.b2c_emulate_persist_threads(flat.ctypes.data, out.ctypes.data, w, h, ch, int(reps), int(k), int(vw),
                                          int(chunk_rows), int(workers), int(max_delay_us), int(seed)),
          "emulate_persist_threads")
    return restore(out)


def numpy_oracle(img, reps: int, boundary: str = "zero"):
    """Gaussian only.  Vectorised shifts on a padded int32 array; exact."""
    flat, w, h, ch, restore = _as_image(img)
    x = flat.reshape(h, w, ch).astype(np.int32)
    mode = "wrap" if boundary == "torus" else "constant"
    if boundary not in ("zero", "torus"):
        raise ValueError("numpy_oracle supports zero/torus")
    for _ in range(int(reps)):
        p = np.pad(x, ((1, 1), (1, 1), (0, 0)), mode=mode)
        hsum = p[:, :-2] + 2 * p[:, 1:-1] + p[:, 2:]
        x = (hsum[:-2] + 2 * hsum[1:-1] + hsum[2:]) >> 4
    return restore(x.astype(np.uint8).reshape(h, w * ch))
