"""The allocation-free entry point `b2c_blur_device_noalloc`, called straight through the C ABI.

It exists for callers that are not allowed to allocate device memory (a plugin inside a host
application that owns the GPU's memory, a real-time loop that cannot take cudaMalloc's
device-wide stall, a job on a fixed memory budget); it works out of a fixed scratch array inside
the library instead.  Here its output is compared byte for byte — the tolerance every other test
of this integer filter uses — with the ordinary allocating entry point `b2c_blur_device`, on two
calls in a row that share (and must each re-initialise) that scratch.

This file sorts last on purpose: a `-x` run has executed every other test before it gets here.
"""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from b200conv._lib import LIB_PATH, Opts

ZERO, TORUS = 0, 1


@pytest.fixture(scope="module")
def abi():
    lib = C.CDLL(str(LIB_PATH))
    vp, i64, i = C.c_void_p, C.c_int64, C.c_int
    lib.b2c_blur_device_noalloc.restype = i
    lib.b2c_blur_device_noalloc.argtypes = [vp, i64, vp, i64, i, i, i, i, i, vp]
    lib.b2c_blur_device.restype = i
    lib.b2c_blur_device.argtypes = [vp, i64, vp, i64, i, i, i, i, C.POINTER(Opts), vp]
    lib.b2c_default_opts.restype = None
    lib.b2c_default_opts.argtypes = [C.POINTER(Opts)]
    lib.b2c_last_error.restype = C.c_char_p
    return lib


def both(abi, img, reps, boundary):
    """-> (allocating entry point's output, allocation-free entry point's output)."""
    h, w = img.shape[0], img.shape[1]
    ch = 1 if img.ndim == 2 else 3
    src = torch.from_numpy(img).cuda().contiguous()
    want = torch.empty_like(src)
    got = torch.zeros_like(src)
    stream = torch.cuda.current_stream().cuda_stream
    o = Opts()
    abi.b2c_default_opts(C.byref(o))
    o.boundary = boundary
    rc = abi.b2c_blur_device(src.data_ptr(), 0, want.data_ptr(), 0, w, h, ch, reps, C.byref(o), stream)
    assert rc == 0, abi.b2c_last_error()
    rc = abi.b2c_blur_device_noalloc(src.data_ptr(), 0, got.data_ptr(), 0, w, h, ch, reps, boundary, stream)
    assert rc == 0, abi.b2c_last_error()
    torch.cuda.synchronize()
    return want.cpu().numpy(), got.cpu().numpy()


def test_two_calls_in_a_row_match_the_allocating_entry_point(abi):
    rng = np.random.default_rng(42)
    # First call: a large rgb image fills most of the scratch with bright data and torus epochs.
    big = rng.integers(128, 256, (600, 1000, 3), dtype=np.uint8)
    want, got = both(abi, big, 13, TORUS)
    assert (got == want).all()
    # Second call right after it: a smaller grey image with the ZERO boundary — anything the
    # first call left in the halo frames or flag words would show up along the edges.
    small = rng.integers(0, 256, (97, 203), dtype=np.uint8)
    want, got = both(abi, small, 9, ZERO)
    assert (got == want).all()


def test_device_memory_is_left_alone(abi):
    """The point of the entry point: no device memory is taken or given back.  After one warm-up
    call (the first call on a device may load the module that holds the scratch array), the free
    memory the driver reports must be the same to the byte before and after two more calls —
    while the allocating entry point, run in between for contrast, is seen to move it."""
    rng = np.random.default_rng(3)
    img = rng.integers(0, 256, (1200, 1600, 3), dtype=np.uint8)
    src = torch.from_numpy(img).cuda().contiguous()
    out = torch.empty_like(src)
    stream = torch.cuda.current_stream().cuda_stream
    h, w = img.shape[0], img.shape[1]

    def call(boundary, reps):
        rc = abi.b2c_blur_device_noalloc(src.data_ptr(), 0, out.data_ptr(), 0, w, h, 3, reps, boundary, stream)
        assert rc == 0, abi.b2c_last_error()

    call(ZERO, 3)
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    call(TORUS, 12)
    free_mid, _ = torch.cuda.mem_get_info()      # the work is still in flight: nothing may be held meanwhile either
    call(ZERO, 7)
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free0 == free_mid == free1, (free0, free_mid, free1)


def test_matches_the_oracle_too(abi):
    from b200conv import reference_oracle

    img = np.random.default_rng(1).integers(0, 256, (120, 160, 3), dtype=np.uint8)
    for reps, boundary, name in ((0, ZERO, "zero"), (5, ZERO, "zero"), (8, TORUS, "torus")):
        _, got = both(abi, img, reps, boundary)
        assert (got == reference_oracle(img, reps, name)).all()


def test_rejects_images_beyond_the_scratch(abi):
    src = torch.zeros((4096, 4096, 3), dtype=torch.uint8, device="cuda")
    dst = torch.empty_like(src)
    rc = abi.b2c_blur_device_noalloc(src.data_ptr(), 0, dst.data_ptr(), 0, 4096, 4096, 3, 1, ZERO,
                                     torch.cuda.current_stream().cuda_stream)
    assert rc == -5  # B2C_ETOOBIG
    msg = abi.b2c_last_error().decode()
    assert "scratch" in msg and "33554432" in msg
    # the README-sized image (1920x2520 rgb) is the documented upper end and must fit
    src = torch.zeros((2520, 1920, 3), dtype=torch.uint8, device="cuda")
    dst = torch.empty_like(src)
    rc = abi.b2c_blur_device_noalloc(src.data_ptr(), 0, dst.data_ptr(), 0, 1920, 2520, 3, 2, ZERO,
                                     torch.cuda.current_stream().cuda_stream)
    assert rc == 0, abi.b2c_last_error()
    torch.cuda.synchronize()
    assert int(dst.max()) == 0
