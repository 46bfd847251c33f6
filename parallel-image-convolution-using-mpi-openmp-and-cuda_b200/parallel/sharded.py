"""One process per GPU under ``torch.distributed`` (backend ``nccl``): the product path.

``torch.distributed`` is only the plumbing — rendezvous, shipping the 256-byte peer
handles, barriers and the max-over-ranks reduction of the timer (the jobs of
MPI_Init/Bcast/Barrier/Send/Recv in mpi/mpi_convolution.c:22-25,65-70,151,264-275).
Inside the iteration loop there is no NCCL, MPI or cudaMemcpyPeer call: halos are stored
straight into the neighbours' frames by the stencil kernel over NVLink peer mappings
(cudaIpc), ordered by release/acquire epoch flags.
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from .._lib import HANDLE_BYTES
from ..ops.stencil import Engine
from .decomposition import decompose


class ShardedBlur:
    def __init__(self, width, height, channels, group=None, grid=None, device=None, **opts):
        if not dist.is_initialized():
            raise RuntimeError("torch.distributed is not initialised")
        self.group = group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.width, self.height, self.channels = int(width), int(height), int(channels)
        self.grid = tuple(grid) if grid is not None else decompose(self.world, height, width)
        if self.grid[0] * self.grid[1] != self.world:
            raise ValueError(f"grid {self.grid} does not match world size {self.world}")
        self.device = torch.cuda.current_device() if device is None else int(device)
        self.engine = Engine(width, height, channels, grid=self.grid, rank=self.rank, device=self.device, **opts)
        needs_peers = self.world > 1
        if needs_peers:
            mine = torch.frombuffer(bytearray(self.engine.export_handle()), dtype=torch.uint8).to(f"cuda:{self.device}")
            allh = [torch.empty(HANDLE_BYTES, dtype=torch.uint8, device=mine.device) for _ in range(self.world)]
            dist.all_gather(allh, mine, group=group)
            self.engine.import_handles([bytes(t.cpu().numpy().tobytes()) for t in allh])
        self._ev0 = torch.cuda.Event(enable_timing=True)
        self._ev1 = torch.cuda.Event(enable_timing=True)

    # geometry of this rank's block
    @property
    def block(self):
        e = self.engine
        return e.row0, e.col0, e.rows, e.cols

    @property
    def block_shape(self):
        return self.engine.block_shape

    def close(self):
        # cudaIpc tear-down order: (1) nobody may still be writing into a peer -> barrier,
        # (2) every rank unmaps its imports, (3) barrier, (4) only then free the exported memory
        # (freeing memory a peer still has mapped is undefined behaviour and did corrupt later
        # engines on 8 GPUs).
        torch.cuda.synchronize(self.device)
        if self.world > 1:
            dist.barrier(group=self.group)
            self.engine.disconnect()
            dist.barrier(group=self.group)
        self.engine.close()

    def load(self, block: torch.Tensor):
        """Copy this rank's block in and run the initial halo exchange (collective)."""
        self.engine.load(block)
        torch.cuda.current_stream().synchronize()
        if self.world > 1:
            dist.barrier(group=self.group)  # every rank's block is in place before frames are written
        self.engine.exchange()

    def run(self, repetitions: int):
        self.engine.run(repetitions)

    def store(self, out: torch.Tensor):
        return self.engine.store(out)

    def timed_run(self, repetitions: int) -> float:
        """Barrier, run, device-timed loop in milliseconds, max over ranks
        (mpi_convolution.c:151-154,242,264-275)."""
        torch.cuda.synchronize(self.device)
        if self.world > 1:
            dist.barrier(group=self.group)
        self._ev0.record()
        self.engine.run(repetitions)
        self._ev1.record()
        self._ev1.synchronize()
        ms = torch.tensor([self._ev0.elapsed_time(self._ev1)], dtype=torch.float64, device=f"cuda:{self.device}")
        if self.world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX, group=self.group)
        return float(ms.item())

    def iterations_done(self) -> int:
        return self.engine.iterations_done()
