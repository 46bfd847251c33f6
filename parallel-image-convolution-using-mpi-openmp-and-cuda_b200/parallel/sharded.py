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
        self.tuned = None
        self.nvls = False
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
            self.nvls = self._setup_nvls()
            self.tuned = self._tune_coupled()
        self._ev0 = torch.cuda.Event(enable_timing=True)
        self._ev1 = torch.cuda.Event(enable_timing=True)

    def _setup_nvls(self) -> bool:
        """NVLS multicast for the two collectives, when the box exposes it: a 256-word symmetric
        buffer from torch's symmetric memory (which does the cuMulticast plumbing); the engine then
        issues one ``multimem.red`` per convergence check / start barrier instead of one peer store
        and one poll per rank.  ``B200CONV_NVLS=0`` keeps the peer stores."""
        import os

        ok = 0
        if os.environ.get("B200CONV_NVLS", "1") != "0":
            try:
                import torch.distributed._symmetric_memory as symm_mem

                buf = symm_mem.empty(256, dtype=torch.uint32, device=f"cuda:{self.device}")
                buf.zero_()
                hdl = symm_mem.rendezvous(buf, group=self.group if self.group is not None else dist.group.WORLD)
                mc = int(getattr(hdl, "multicast_ptr", 0) or 0)
                if mc:
                    self._nvls_buf, self._nvls_hdl, self._nvls_mc = buf, hdl, mc
                    ok = 1
            except Exception:  # noqa: BLE001  (no symmetric memory / no multicast on this box)
                ok = 0
        # all ranks or none
        flag = torch.tensor([ok], dtype=torch.int32, device=f"cuda:{self.device}")
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=self.group)
        torch.cuda.synchronize(self.device)
        if int(flag.item()) == 1:
            self.engine.set_nvls(self._nvls_buf.data_ptr(), self._nvls_mc, 256)
            return True
        return False

    def _tune_coupled(self):
        """Rows per warp are chosen by timing a few candidates with the halo exchange live, all
        ranks together (the optimum without neighbours is wrong once ring warps gate them)."""
        n = self.engine.tune_count()
        if n <= 0:
            return None
        dist.barrier(group=self.group)   # every rank has imported: nobody writes into an unmapped peer
        ms = torch.tensor([self.engine.tune_time(i) for i in range(n)], dtype=torch.float64,
                          device=f"cuda:{self.device}")
        dist.all_reduce(ms, op=dist.ReduceOp.MAX, group=self.group)
        best = int(torch.argmin(ms).item())
        self.engine.tune_set(best)
        torch.cuda.synchronize(self.device)
        dist.barrier(group=self.group)
        return {"candidate": best, "ms": [round(float(x), 4) for x in ms.tolist()]}

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

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()   # collective: every rank leaves the block together
        return False

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
            self.engine.barrier()   # device-side: the timed region starts within microseconds on all ranks
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


class ShardedPipeline:
    """A stream of same-shaped sharded images through one :class:`ShardedBlur`.

    Per image the work is unchanged — host->device copy of this rank's block from pinned memory,
    initial halo exchange, ``repetitions`` iterations, device->host copy of the result — but the
    PCIe copies run on two copy streams through device staging buffers, so the upload of image
    i+1 and the download of image i-1 overlap the iterations of image i (the copy engines work
    beside the SMs; PCIe is full duplex).  Only one stream ever carries the halo-exchanging
    kernels, so no two kernels that wait on flags share a GPU.

    No host barrier between images: a rank that runs ahead only writes frames that its
    neighbours have finished reading (the last pass a rank completed required the neighbours'
    flags of the pass before, exactly the slot-reuse rule of the iteration loop), and it cannot
    start iterating before the neighbours' exchange of the same image has published its epoch.
    The threaded host model runs exactly this (``images`` > 1, csrc/core/halo_emul.cpp) with skewed
    ranks, under ThreadSanitizer too (tests/test_halo_model.py).

        pipe = ShardedPipeline(sharded, repetitions=1000)
        for src, dst in zip(pinned_blocks_in, pinned_blocks_out):
            pipe.submit(src, dst)          # every rank submits the same number of images
        pipe.wait()
    """

    def __init__(self, sharded: ShardedBlur, repetitions: int, depth: int = 2, trace: bool = False):
        self.trace = bool(trace)
        self._marks = []   # trace: per image {name: timing event}
        if int(sharded.engine.opts.check_every):
            raise ValueError("ShardedPipeline does not support check_every (the host must read the latch per image)")
        self.sharded = sharded
        self.repetitions = int(repetitions)
        self.depth = int(depth)
        dev = torch.device(f"cuda:{sharded.device}")
        shape = sharded.block_shape
        with torch.cuda.device(dev):
            self._in = [torch.empty(shape, dtype=torch.uint8, device=dev) for _ in range(self.depth)]
            self._out = [torch.empty(shape, dtype=torch.uint8, device=dev) for _ in range(self.depth)]
            self.s_compute = torch.cuda.Stream(device=dev)
            self.s_h2d = torch.cuda.Stream(device=dev)
            self.s_d2h = torch.cuda.Stream(device=dev)
            mk = lambda: [torch.cuda.Event() for _ in range(self.depth)]  # noqa: E731
            self._in_ready, self._in_free, self._out_ready, self._out_free = mk(), mk(), mk(), mk()
            torch.cuda.synchronize(dev)
        if sharded.world > 1:
            dist.barrier(group=sharded.group)   # once: every rank's engine is idle and mapped
        self._n = 0

    def submit(self, src: torch.Tensor, dst: torch.Tensor):
        """Enqueue one image: ``src``/``dst`` are this rank's block (pinned CPU tensors). Returns at once."""
        eng = self.sharded.engine
        i = self._n % self.depth
        first_use = self._n < self.depth
        self._n += 1
        marks = {}

        def mark(name, stream):
            if self.trace:
                ev = torch.cuda.Event(enable_timing=True)
                ev.record(stream)
                marks[name] = ev

        with torch.cuda.device(self.sharded.device):
            if not first_use:
                self.s_h2d.wait_event(self._in_free[i])
            with torch.cuda.stream(self.s_h2d):
                mark("h2d_start", self.s_h2d)
                self._in[i].copy_(src, non_blocking=True)
                self._in_ready[i].record(self.s_h2d)
                mark("h2d_end", self.s_h2d)
            mark("compute_free", self.s_compute)   # the previous image's store has been issued before this
            self.s_compute.wait_event(self._in_ready[i])
            eng.load(self._in[i], stream=self.s_compute)
            self._in_free[i].record(self.s_compute)
            eng.exchange(stream=self.s_compute)
            mark("run_start", self.s_compute)
            eng.run(self.repetitions, stream=self.s_compute)
            mark("run_end", self.s_compute)
            if not first_use:
                self.s_compute.wait_event(self._out_free[i])
            eng.store(self._out[i], stream=self.s_compute)
            self._out_ready[i].record(self.s_compute)
            mark("store_end", self.s_compute)
            self.s_d2h.wait_event(self._out_ready[i])
            with torch.cuda.stream(self.s_d2h):
                dst.copy_(self._out[i], non_blocking=True)
                self._out_free[i].record(self.s_d2h)
                mark("d2h_end", self.s_d2h)
        if self.trace:
            self._marks.append(marks)

    def wait(self):
        for s in (self.s_h2d, self.s_compute, self.s_d2h):
            s.synchronize()

    def timeline(self):
        """With ``trace=True``: per submitted image, milliseconds (device clock, relative to the first
        image's upload start) at which each phase boundary passed on its stream."""
        self.wait()
        if not self._marks:
            return []
        t0 = self._marks[0]["h2d_start"]
        return [{k: round(t0.elapsed_time(ev), 3) for k, ev in m.items()} for m in self._marks]
