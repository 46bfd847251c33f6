"""Headerless interleaved-uint8 ``.raw`` files (layer L3): grey = H rows x W bytes, rgb =
H rows x 3W bytes R,G,B interleaved, no header, no row padding
(mpi/mpi_convolution.c:126-140,244-262; cuda/main.c:23-30,35-45).  The byte moving is done
by the C++ core (64-bit ``pread``/``pwrite``)."""
from __future__ import annotations

import ctypes as C

import numpy as np

from .._lib import check, lib


def _shape(h, w, ch):
    return (h, w) if ch == 1 else (h, w, ch)


def read_block(path, width, height, channels, row0, col0, rows, cols, out=None) -> np.ndarray:
    """One rank's R x C block, read with per-row offsets (the MPI-IO pattern of :126-140)."""
    if out is None:
        out = np.empty(_shape(rows, cols, channels), dtype=np.uint8)
    check(lib.b2c_read_block(str(path).encode(), out.ctypes.data, cols * channels, int(width), int(height),
                             int(channels), int(row0), int(col0), int(rows), int(cols)), "read_block")
    return out


def write_block(path, block: np.ndarray, width, height, channels, row0, col0, create=True):
    block = np.ascontiguousarray(block)
    rows, cols = block.shape[0], block.shape[1]
    check(lib.b2c_write_block(str(path).encode(), block.ctypes.data, cols * channels, int(width), int(height),
                              int(channels), int(row0), int(col0), rows, cols, 1 if create else 0), "write_block")


def read_raw(path, width, height, channels) -> np.ndarray:
    return read_block(path, width, height, channels, 0, 0, height, width)


def write_raw(path, img: np.ndarray):
    img = np.ascontiguousarray(img)
    ch = 1 if img.ndim == 2 else img.shape[2]
    write_block(path, img, img.shape[1], img.shape[0], ch, 0, 0)


def output_name(image: str) -> str:
    """``"blur_" + argv[1]`` — the whole argument string, like :244-246 / cuda/main.c:36-38."""
    buf = C.create_string_buffer(len(image) + 16)
    check(lib.b2c_output_name(image.encode(), buf, len(image) + 16))
    return buf.value.decode()
