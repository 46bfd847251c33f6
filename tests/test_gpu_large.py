"""64-bit paths on one GPU: an image larger than 2^31 bytes (where every reference program
overflows its `int` sizes and offsets, cuda/main.c:27, cuda_convolution.cu:21,38) must blur
correctly across the 2^31-byte line.  Checked against an independent blur of a crop that
straddles the line (rows far enough from the crop's own borders are unaffected by them)."""
import pytest
import torch

pytestmark = pytest.mark.gpu

import b200conv


def test_image_beyond_2_gib_is_consistent_across_the_int_overflow_line():
    free, _ = torch.cuda.mem_get_info()
    W, H, ch, reps = 32768, 32768, 3, 4
    nbytes = W * H * ch
    assert nbytes > 2 ** 31
    if free < 5 * nbytes:
        pytest.skip("not enough free device memory")
    g = torch.Generator(device="cuda").manual_seed(5)
    img = torch.randint(0, 256, (H, W, ch), dtype=torch.uint8, device="cuda", generator=g)
    out = b200conv.blur(img, reps, boundary="zero", tb=4)
    torch.cuda.synchronize()
    row = (2 ** 31) // (W * ch)           # the row that contains byte offset 2^31
    lo, hi = row - 40, row + 40
    crop = img[lo:hi].contiguous()
    ref = b200conv.blur(crop, reps, boundary="zero", tb=1, kernel="naive")   # different kernel on purpose
    torch.cuda.synchronize()
    m = reps + 1
    assert torch.equal(out[lo + m:hi - m], ref[m:-m])
    # and the far end of the buffer (last rows) against the zero boundary
    tail = img[H - 64:].contiguous()
    ref2 = b200conv.blur(tail, reps, boundary="zero", tb=1, kernel="naive")
    assert torch.equal(out[H - 64 + m:], ref2[m:])
    b200conv.clear_engine_cache()
