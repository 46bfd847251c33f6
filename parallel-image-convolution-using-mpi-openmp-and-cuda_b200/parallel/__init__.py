"""Multi-GPU layer: decomposition, peer-memory halo engine, NCCL baseline."""
from .decomposition import block_of, decompose, neighbours, rows_division
from .local import LocalGrid
from .sendrecv import GlooHaloBlur, NcclHaloBlur, SendRecvHaloBlur
from .sharded import ShardedBlur, ShardedPipeline

__all__ = ["rows_division", "decompose", "neighbours", "block_of", "LocalGrid", "ShardedBlur", "ShardedPipeline",
           "SendRecvHaloBlur", "NcclHaloBlur", "GlooHaloBlur"]
