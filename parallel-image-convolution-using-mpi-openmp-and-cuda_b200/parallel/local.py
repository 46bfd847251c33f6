"""All ranks of a grid inside ONE process — on one GPU (sequential emulation used by the
tests) or on several GPUs with peer access.

On a single GPU the ranks' kernels share the device, so a rank may never wait for a kernel
that has not been launched yet: ``run`` therefore steps pass by pass over the ranks
(every wait is already satisfied when its kernel starts).  With one GPU per rank the whole
run of every rank is enqueued at once and the devices synchronise through the epoch flags.
"""
from __future__ import annotations

import torch

from ..ops.stencil import Engine
from .decomposition import block_of


class LocalGrid:
    def __init__(self, width, height, channels, grid, devices=None, **opts):
        self.width, self.height, self.channels = int(width), int(height), int(channels)
        self.grid = (int(grid[0]), int(grid[1]))
        n = self.grid[0] * self.grid[1]
        self.devices = [torch.cuda.current_device()] * n if devices is None else list(devices)
        self.multi_device = len(set(self.devices)) == n and n > 1
        self.engines = []
        for k in range(n):
            with torch.cuda.device(self.devices[k]):
                self.engines.append(Engine(width, height, channels, grid=self.grid, rank=k,
                                           device=self.devices[k], **opts))
        if n > 1 or opts.get("boundary") == "torus":
            Engine.connect_local(self.engines)
        # fused iterations per pass as the engine planned them ("tb=K" in its description)
        import re

        m = re.search(r"tb=(\d+)", self.engines[0].describe())
        self.tb = int(m.group(1)) if m else 1

    def close(self):
        for e in self.engines:
            e.close()

    def _sync(self):
        for d in set(self.devices):
            torch.cuda.synchronize(d)

    def load(self, img: torch.Tensor):
        """Scatter a full image ([H,W] / [H,W,3] uint8, CPU or CUDA) to the ranks and exchange halos."""
        for k, e in enumerate(self.engines):
            r0, c0, r, c = block_of(k, self.grid, self.height, self.width)
            with torch.cuda.device(self.devices[k]):
                blk = img[r0:r0 + r, c0:c0 + c].contiguous()
                if blk.is_cuda and blk.device.index != self.devices[k]:
                    blk = blk.to(f"cuda:{self.devices[k]}")
                e.load(blk)
        self._sync()
        for k, e in enumerate(self.engines):
            with torch.cuda.device(self.devices[k]):
                e.exchange()
        self._sync()

    def run(self, repetitions: int):
        if self.multi_device:
            for k, e in enumerate(self.engines):
                with torch.cuda.device(self.devices[k]):
                    e.run(repetitions)
            return
        if int(self.engines[0].opts.check_every) and len(self.engines) > 1:
            # the cross-rank OR waits on every rank's kernel: impossible to emulate on one GPU
            raise ValueError("check_every needs one GPU per rank")
        if int(self.engines[0].opts.check_every) or self.engines[0].describe().startswith("kernel=resident"):
            for r, e in enumerate(self.engines):
                with torch.cuda.device(self.devices[r]):
                    e.run(repetitions)
            return
        # exactly ONE pass per rank per round, in the engines' own pass order (groups of passes
        # between two halo exchanges included): every flag a kernel waits for is set before it starts
        counts = [e.plan_run(repetitions) for e in self.engines]
        assert len(set(counts)) == 1, counts
        for i in range(counts[0]):
            for r, e in enumerate(self.engines):
                with torch.cuda.device(self.devices[r]):
                    e.run_pass(i)

    def gather(self) -> torch.Tensor:
        shape = (self.height, self.width) if self.channels == 1 else (self.height, self.width, self.channels)
        out = torch.empty(shape, dtype=torch.uint8)
        self._sync()
        for k, e in enumerate(self.engines):
            r0, c0, r, c = block_of(k, self.grid, self.height, self.width)
            with torch.cuda.device(self.devices[k]):
                blk = torch.empty(e.block_shape, dtype=torch.uint8, device=f"cuda:{self.devices[k]}")
                e.store(blk)
                out[r0:r0 + r, c0:c0 + c] = blk.cpu()
        return out
