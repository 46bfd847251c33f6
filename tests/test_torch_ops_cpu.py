"""The PyTorch extension (csrc/torch/binding.cpp) on CPU tensors: `torch.ops.b200conv.blur`
dispatches to the exact CPU oracle, `stencil_step` to the threaded host step, `rows_division` to
the C++ RowsDivision (mpi_convolution.c:350-364).  The CUDA dispatch is covered by test_gpu_stencil."""
import numpy as np
import pytest
import torch

from b200conv import reference_oracle
from b200conv.ops import load_torch_ops


@pytest.fixture(scope="module")
def ops():
    return load_torch_ops()


@pytest.mark.parametrize("boundary", ["zero", "torus", "cuda-compat"])
@pytest.mark.parametrize("ch", [1, 3])
def test_blur_op_on_cpu_tensors_is_the_oracle(ops, boundary, ch):
    img = np.random.default_rng(1).integers(0, 256, (19, 27) if ch == 1 else (19, 27, ch), dtype=np.uint8)
    got = ops.blur(torch.from_numpy(img), 6, boundary)
    assert got.dtype == torch.uint8 and tuple(got.shape) == img.shape
    assert (got.numpy() == reference_oracle(img, 6, boundary)).all()


def test_blur_op_checks_its_arguments(ops):
    img = torch.zeros((4, 4), dtype=torch.uint8)
    with pytest.raises(RuntimeError, match="boundary"):
        ops.blur(img, 1, "mirror")
    with pytest.raises(RuntimeError, match="kernel"):
        ops.blur(img, 1, "zero", "gaussian", 0, "warp-speed")
    with pytest.raises(RuntimeError, match="uint8"):
        ops.blur(img.float(), 1)
    with pytest.raises(RuntimeError):
        ops.blur(torch.zeros((4, 4, 2), dtype=torch.uint8), 1)
    # non-contiguous input is accepted (made contiguous inside)
    big = torch.arange(64, dtype=torch.uint8).reshape(8, 8)
    view = big[:, ::2]
    assert (ops.blur(view, 2).numpy() == reference_oracle(view.contiguous().numpy(), 2, "zero")).all()


def test_stencil_step_op_on_cpu(ops):
    h, w, ch = 9, 14, 3
    img = np.random.default_rng(2).integers(0, 256, (h, w, ch), dtype=np.uint8)
    bw, pitch = w * ch, w * ch + 2 * ch + 2
    src = torch.zeros((h + 2, pitch), dtype=torch.uint8)
    dst = torch.zeros_like(src)
    src[1:h + 1, ch:ch + bw] = torch.from_numpy(img.reshape(h, bw))
    ops.stencil_step(src, dst, h, bw, ch, "gaussian", 2)
    assert (dst[1:h + 1, ch:ch + bw].numpy().reshape(img.shape) == reference_oracle(img, 1, "zero")).all()
    with pytest.raises(RuntimeError, match="frame"):
        ops.stencil_step(src, dst, h + 1, bw, ch)


def test_rows_division_op(ops):
    assert ops.rows_division(4, 2520, 1920) == 2 and ops.rows_division(7, 10, 10) == 0
