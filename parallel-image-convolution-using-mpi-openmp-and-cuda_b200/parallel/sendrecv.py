"""Halo exchange by point-to-point messages over ``torch.distributed`` — the *baseline*, not the product.

``SendRecvHaloBlur`` is the transliteration of the reference's Isend/Irecv/Wait loop
(mpi/mpi_convolution.c:156-240) onto ``torch.distributed``, with the missing corner exchange
added so that the result is exact on every grid.  Per iteration: pack the four edge rows/columns
and four corners (torch slicing), one grouped ``batch_isend_irecv``, unpack into the 1-pixel halo of
a padded tensor, then one plain stencil pass over the block.  With ``halo_depth`` F > 1 the halo is
F pixels deep and exchanged once per F iterations; in between, a rank recomputes the part of the
halo that is still valid itself (its block extended by F-1, F-2, ... pixels towards the sides that
have a neighbour) — the communication-avoiding schedule of the fused GPU engine, with messages.

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
    def __init__(self, width, height, channels, boundary="zero", group=None, device=None, threads=0, halo_depth=1):
        self.group = group
        self.threads = int(threads)   # CPU only: threads per rank for the stencil pass (0 = all cores)
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        self.W, self.H, self.ch = int(width), int(height), int(channels)
        self.grid = decompose(self.world, height, width)
        self.nb = neighbours(self.rank, self.grid, boundary)
        self.r0, self.c0, self.R, self.C = block_of(self.rank, self.grid, height, width)
        self.bw = self.C * self.ch
        # halo depth F: a halo pixel must come from the adjacent rank only
        self.F = max(1, min(int(halo_depth), self.R, self.C))
        self.fb = self.F * self.ch
        pitch = (self.bw + 2 * self.fb + 127) // 128 * 128
        if device is None:
            device = "cuda" if dist.get_backend(group) == "nccl" else "cpu"
        dev = torch.device("cuda", torch.cuda.current_device()) if str(device).startswith("cuda") else torch.device("cpu")
        self.buf = [torch.zeros((self.R + 2 * self.F, pitch), dtype=torch.uint8, device=dev) for _ in range(2)]
        self.cur = 0
        self.exchanges = 0

    def _region(self, t, d, halo):
        """Slice of the padded buffer: the F-deep edge we send towards d (halo=False) or the F-deep halo
        we receive from d (halo=True)."""
        R, bw, F, fb = self.R, self.bw, self.F, self.fb
        dr, dc = _DIRS[d]
        if halo:
            rows = slice(0, F) if dr < 0 else (slice(R + F, R + 2 * F) if dr > 0 else slice(F, R + F))
            cols = slice(0, fb) if dc < 0 else (slice(fb + bw, 2 * fb + bw) if dc > 0 else slice(fb, fb + bw))
        else:
            rows = slice(F, 2 * F) if dr < 0 else (slice(R, R + F) if dr > 0 else slice(F, R + F))
            cols = slice(fb, 2 * fb) if dc < 0 else (slice(bw, bw + fb) if dc > 0 else slice(fb, fb + bw))
        return t[rows, cols]

    def load(self, block: torch.Tensor):
        self.buf[self.cur][self.F:self.R + self.F, self.fb:self.fb + self.bw] = \
            block.reshape(self.R, self.bw).to(self.buf[0].device)

    def store(self) -> torch.Tensor:
        out = self.buf[self.cur][self.F:self.R + self.F, self.fb:self.fb + self.bw].contiguous()
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
        self.exchanges += 1

    def run(self, repetitions: int):
        left = int(repetitions)
        F, fb, ch, R, bw = self.F, self.fb, self.ch, self.R, self.bw
        has = {d: self.nb[d] >= 0 for d in ("N", "S", "W", "E")}
        while left > 0:
            g = min(F, left)
            self.exchange()                      # one F-deep exchange per group of up to F iterations
            for j in range(g):
                ext = g - 1 - j                  # iterations left in the group: how far the block is extended
                xn, xs = (ext if has["N"] else 0), (ext if has["S"] else 0)
                xw, xe = (ext * ch if has["W"] else 0), (ext * ch if has["E"] else 0)
                rows, row_bytes = R + xn + xs, bw + xw + xe
                r0, c0 = F - xn - 1, fb - xw - ch   # view whose [1, ch] element is the first pixel computed
                src = self.buf[self.cur][r0:r0 + rows + 2, c0:]
                dst = self.buf[self.cur ^ 1][r0:r0 + rows + 2, c0:]
                stencil_step(src, dst, rows, row_bytes, ch, threads=self.threads)
                self.cur ^= 1
            left -= g


class NcclHaloBlur(SendRecvHaloBlur):
    """CUDA tensors over NCCL send/recv: the baseline arm of the multi-GPU benches."""

    def __init__(self, width, height, channels, boundary="zero", group=None, halo_depth=1):
        super().__init__(width, height, channels, boundary, group, device="cuda", halo_depth=halo_depth)


class GlooHaloBlur(SendRecvHaloBlur):
    """CPU tensors over gloo, ``threads`` stencil threads per rank: the host-only arm."""

    def __init__(self, width, height, channels, boundary="zero", group=None, threads=0, halo_depth=1):
        super().__init__(width, height, channels, boundary, group, device="cpu", threads=threads, halo_depth=halo_depth)
