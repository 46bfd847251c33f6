"""Halo exchange by point-to-point messages over ``torch.distributed`` — the *baseline*, not the product.

``SendRecvHaloBlur`` is the transliteration of the reference's Isend/Irecv/Wait loop
(mpi/mpi_convolution.c:156-240) onto ``torch.distributed``, with the missing corner exchange
added so that the result is exact on every grid.  Per iteration: pack the four edge rows/columns
and four corners (torch slicing), one grouped ``batch_isend_irecv``, unpack into the 1-pixel halo of
a padded tensor, then one plain stencil pass over the block.

* backend ``nccl`` (``NcclHaloBlur``): CUDA tensors, ncclGroupStart + ncclSend/ncclRecv, byte-per-thread
  kernel ``b2c_stencil_step_device`` — the NCCL baseline that the fused peer-memory engine is
  measured against.
* backend ``gloo`` (``GlooHaloBlur``): CPU tensors and the threaded ``b2c_stencil_step_host`` — ranks
  x threads on the host cores, i.e. what the reference's MPI and MPI+OpenMP programs are
  (omp_convolution.c:292,297), usable where no GPU exists and run by the CPU test tier with
  ``world_size`` 2 and 4.
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from ..ops.stencil import stencil_step
from .decomposition import block_of, decompose, neighbours

_DIRS = {"N": (-1, 0), "S": (1, 0), "W": (0, -1), "E": (0, 1), "NW": (-1, -1), "NE": (-1, 1), "SW": (1, -1), "SE": (1, 1)}
_OPP = {"N": "S", "S": "N", "W": "E", "E": "W", "NW": "SE", "SE": "NW", "NE": "SW", "SW": "NE"}


class SendRecvHaloBlur:
    def __init__(self, width, height, channels, boundary="zero", group=None, device=None, threads=0):
        self.group = group
        self.threads = int(threads)   # CPU only: threads per rank for the stencil pass (0 = all cores)
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        self.W, self.H, self.ch = int(width), int(height), int(channels)
        self.grid = decompose(self.world, height, width)
        self.nb = neighbours(self.rank, self.grid, boundary)
        self.r0, self.c0, self.R, self.C = block_of(self.rank, self.grid, height, width)
        self.bw = self.C * self.ch
        pitch = (self.bw + 2 * self.ch + 127) // 128 * 128
        if device is None:
            device = "cuda" if dist.get_backend(group) == "nccl" else "cpu"
        dev = torch.device("cuda", torch.cuda.current_device()) if str(device).startswith("cuda") else torch.device("cpu")
        self.buf = [torch.zeros((self.R + 2, pitch), dtype=torch.uint8, device=dev) for _ in range(2)]
        self.cur = 0

    def _region(self, t, d, halo):
        """Slice of the padded buffer: the edge we send towards d (halo=False) or the halo we
        receive from d (halo=True)."""
        ch, R, bw = self.ch, self.R, self.bw
        dr, dc = _DIRS[d]
        if halo:
            rows = slice(0, 1) if dr < 0 else (slice(R + 1, R + 2) if dr > 0 else slice(1, R + 1))
            cols = slice(0, ch) if dc < 0 else (slice(ch + bw, 2 * ch + bw) if dc > 0 else slice(ch, ch + bw))
        else:
            rows = slice(1, 2) if dr < 0 else (slice(R, R + 1) if dr > 0 else slice(1, R + 1))
            cols = slice(ch, 2 * ch) if dc < 0 else (slice(bw, bw + ch) if dc > 0 else slice(ch, ch + bw))
        return t[rows, cols]

    def load(self, block: torch.Tensor):
        self.buf[self.cur][1:self.R + 1, self.ch:self.ch + self.bw] = block.reshape(self.R, self.bw).to(self.buf[0].device)

    def store(self) -> torch.Tensor:
        out = self.buf[self.cur][1:self.R + 1, self.ch:self.ch + self.bw].contiguous()
        return out if self.ch == 1 else out.reshape(self.R, self.C, self.ch)

    def exchange(self):
        src = self.buf[self.cur]
        ops, recvs = [], []
        # One "flow direction" d at a time: I send my d-side edge towards nb[d] and receive, from
        # nb[opposite(d)], the data flowing the same way into my opposite-side halo.  Sender and
        # receiver enumerate flows in the same order, so two messages between one pair of ranks
        # (2-wide torus: W and E neighbour are the same rank) cannot be matched crosswise — the
        # single-tag hazard noted in SURVEY §2.6.
        for d in _DIRS:
            peer, o = self.nb[d], _OPP[d]
            if peer == self.rank:  # torus with a 1-wide grid dimension: local wrap copy
                self._region(src, o, True).copy_(self._region(src, d, False))
                continue
            if peer >= 0:
                send = self._region(src, d, False).contiguous()                   # pack
                ops.append(dist.P2POp(dist.isend, send, peer, self.group))
            if self.nb[o] >= 0:
                recv = torch.empty_like(self._region(src, o, True).contiguous())
                ops.append(dist.P2POp(dist.irecv, recv, self.nb[o], self.group))
                recvs.append((o, recv))
        if ops:
            for w in dist.batch_isend_irecv(ops):
                w.wait()
        for d, recv in recvs:
            self._region(src, d, True).copy_(recv)                                # unpack

    def run(self, repetitions: int):
        for _ in range(int(repetitions)):
            self.exchange()
            stencil_step(self.buf[self.cur], self.buf[self.cur ^ 1], self.R, self.bw, self.ch, threads=self.threads)
            self.cur ^= 1


class NcclHaloBlur(SendRecvHaloBlur):
    """CUDA tensors over NCCL send/recv: the baseline arm of the multi-GPU benches."""

    def __init__(self, width, height, channels, boundary="zero", group=None):
        super().__init__(width, height, channels, boundary, group, device="cuda")


class GlooHaloBlur(SendRecvHaloBlur):
    """CPU tensors over gloo, ``threads`` stencil threads per rank: the host-only arm."""

    def __init__(self, width, height, channels, boundary="zero", group=None, threads=0):
        super().__init__(width, height, channels, boundary, group, device="cpu", threads=threads)
