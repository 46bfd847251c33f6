"""b200conv — Blackwell-native iterative 3x3 image convolution (B200, sm_100a).

Capabilities of theopaid/Parallel-Image-Convolution-using-MPI-OPENMP-and-CUDA (its MPI,
MPI+OpenMP and CUDA programs) on one C++/CUDA core:

  ``blur(img, reps, boundary=...)``        single GPU, torch tensors
  ``BlurPipeline``                         single GPU, stream of host images, copies overlapped
  ``parallel.ShardedBlur``                 one process per GPU, peer-memory halo exchange
  ``parallel.ShardedPipeline``             stream of sharded images, PCIe copies overlapped with the loop
  ``parallel.LocalGrid``                   every rank in one process (tests, multi-GPU threads)
  ``reference_oracle`` / ``numpy_oracle``  CPU oracles
  ``rows_division`` / ``decompose`` / ``neighbours``   the reference's decomposition rule
  ``utils.read_raw`` / ``write_raw``       headerless uint8 .raw files
  ``lib/b200conv``                         the native launcher with the reference's CLI
"""
from . import _lib
from ._lib import B2CError
from .ops import (BlurPipeline, Engine, blur, blur_cpu, blur_file, clear_engine_cache, numpy_oracle, reference_oracle, sharded_oracle,
                  stencil_step)
from .parallel import LocalGrid, ShardedBlur, ShardedPipeline, block_of, decompose, neighbours, rows_division
from . import models, ops, parallel, utils

__version__ = "0.1.0"


def total_launches() -> int:
    """Kernel launches issued by the native library in this process."""
    return int(_lib.lib.b2c_total_launches())
