"""PyTorch-facing stencil operators over ``libb200conv_core``.

``blur`` is the single-GPU public call (the equivalent of running the reference's
``cuda_convolution`` program, cuda/main.c:10-53 + cuda/cuda_convolution.cu:49-102, on an
in-memory image).  ``Engine`` exposes one rank's persistent plan (buffers, peers, passes).
All GPU work runs on the current torch CUDA stream.
"""
from __future__ import annotations

import ctypes as C
from collections import OrderedDict

import torch

from .._lib import HANDLE_BYTES, B2CError, check, last_error, lib, make_opts


def _stream_ptr(stream=None) -> int:
    if stream is None:
        stream = torch.cuda.current_stream()
    return int(stream.cuda_stream)


def _require_cuda():
    if not torch.cuda.is_available():
        raise B2CError("b200conv GPU operators need a CUDA device (B200, sm_100a); none is visible")


def _img_meta(t: torch.Tensor):
    if t.dtype != torch.uint8 or t.dim() not in (2, 3) or (t.dim() == 3 and t.shape[2] not in (1, 3)):
        raise ValueError("image must be a uint8 tensor of shape [H,W] or [H,W,3]")
    h, w = int(t.shape[0]), int(t.shape[1])
    ch = 1 if t.dim() == 2 else int(t.shape[2])
    return w, h, ch


class Engine:
    """One rank's block of a ``grid[0] x grid[1]`` decomposition on one device."""

    def __init__(self, width, height, channels, grid=(1, 1), rank=0, device=None, **opts):
        _require_cuda()
        self.device = torch.cuda.current_device() if device is None else int(device)
        self.opts = make_opts(**opts)
        self.width, self.height, self.channels = int(width), int(height), int(channels)
        self.grid, self.rank = (int(grid[0]), int(grid[1])), int(rank)
        self._h = lib.b2c_engine_create(self.device, self.width, self.height, self.channels, self.grid[0],
                                        self.grid[1], self.rank, C.byref(self.opts))
        if not self._h:
            raise B2CError(f"engine_create: {last_error()}")
        r0, c0, r, c = C.c_int(), C.c_int(), C.c_int(), C.c_int()
        lib.b2c_engine_block(self._h, C.byref(r0), C.byref(c0), C.byref(r), C.byref(c))
        self.row0, self.col0, self.rows, self.cols = r0.value, c0.value, r.value, c.value

    # -- lifetime -----------------------------------------------------------
    def close(self):
        if getattr(self, "_h", None):
            lib.b2c_engine_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()
        return False

    # -- geometry -----------------------------------------------------------
    @property
    def block_shape(self):
        return (self.rows, self.cols) if self.channels == 1 else (self.rows, self.cols, self.channels)

    def describe(self) -> str:
        buf = C.create_string_buffer(512)
        lib.b2c_engine_describe(self._h, buf, 512)
        return buf.value.decode()

    @property
    def launches(self) -> int:
        return int(lib.b2c_engine_launches(self._h))

    # -- peers --------------------------------------------------------------
    def export_handle(self) -> bytes:
        buf = C.create_string_buffer(HANDLE_BYTES)
        check(lib.b2c_engine_export(self._h, buf), "engine_export")
        return buf.raw

    def import_handles(self, blobs):
        data = b"".join(blobs)
        check(lib.b2c_engine_import(self._h, data, len(blobs)), "engine_import")

    @staticmethod
    def connect_local(engines):
        arr = (C.c_void_p * len(engines))(*[e._h for e in engines])
        check(lib.b2c_engine_connect_local(arr, len(engines)), "engine_connect_local")

    def disconnect(self):
        """Unmap imported peer buffers (every rank, then a barrier, then close())."""
        check(lib.b2c_engine_disconnect(self._h), "engine_disconnect")

    def set_delay(self, cycles: int):
        check(lib.b2c_engine_set_delay(self._h, int(cycles)))

    # -- data ---------------------------------------------------------------
    def _check_block(self, t: torch.Tensor):
        if t.dtype != torch.uint8 or tuple(t.shape) != tuple(self.block_shape) or not t.is_contiguous():
            raise ValueError(f"expected contiguous uint8 block of shape {self.block_shape}, got {tuple(t.shape)}")

    def load(self, block: torch.Tensor, stream=None):
        """Copy this rank's block (CPU — ideally pinned — or CUDA tensor) into the engine."""
        self._check_block(block)
        fn = lib.b2c_engine_load_device if block.is_cuda else lib.b2c_engine_upload
        check(fn(self._h, block.data_ptr(), 0, _stream_ptr(stream)), "engine_load")

    def store(self, out: torch.Tensor, stream=None):
        self._check_block(out)
        fn = lib.b2c_engine_store_device if out.is_cuda else lib.b2c_engine_download
        check(fn(self._h, out.data_ptr(), 0, _stream_ptr(stream)), "engine_store")
        return out

    def exchange(self, stream=None):
        check(lib.b2c_engine_exchange(self._h, _stream_ptr(stream)), "engine_exchange")

    def run(self, repetitions: int, stream=None):
        check(lib.b2c_engine_run(self._h, int(repetitions), _stream_ptr(stream)), "engine_run")

    def iterations_done(self) -> int:
        return check(lib.b2c_engine_iterations_done(self._h), "engine_iterations_done")

    def plan_run(self, repetitions: int) -> int:
        """Number of passes of a ``repetitions`` run; enqueue them one by one with :meth:`run_pass`."""
        return check(lib.b2c_engine_plan_run(self._h, int(repetitions)), "engine_plan_run")

    def run_pass(self, i: int, stream=None):
        check(lib.b2c_engine_run_pass(self._h, int(i), _stream_ptr(stream)), "engine_run_pass")

    # -- coupled tuning, device barrier, protocol counters (multi-rank) --------
    def tune_count(self) -> int:
        return check(lib.b2c_engine_tune_count(self._h), "engine_tune_count")

    def tune_time(self, i: int, stream=None) -> float:
        ms = float(lib.b2c_engine_tune_time(self._h, int(i), _stream_ptr(stream)))
        if ms < 0:
            raise B2CError(f"engine_tune_time: {last_error()}")
        return ms

    def tune_set(self, i: int):
        check(lib.b2c_engine_tune_set(self._h, int(i)), "engine_tune_set")

    def barrier(self, stream=None):
        """Device-side barrier over peer memory (collective)."""
        check(lib.b2c_engine_barrier(self._h, _stream_ptr(stream)), "engine_barrier")

    def set_nvls(self, local_ptr: int, multicast_ptr: int, words: int):
        """Use an NVLS multicast mapping of a small symmetric buffer for the convergence OR and the
        start barrier (collective; the caller keeps the buffer alive)."""
        check(lib.b2c_engine_set_nvls(self._h, local_ptr or None, multicast_ptr or None, int(words)), "engine_set_nvls")

    def barrier_wait_cycles(self) -> int:
        return check(lib.b2c_engine_barrier_wait_cycles(self._h), "engine_barrier_wait_cycles")

    def set_profile(self, on: bool = True):
        check(lib.b2c_engine_set_profile(self._h, int(bool(on))), "engine_set_profile")

    def persist_profile(self) -> dict:
        """B200CONV_PERSIST_PROF=1: where the persistent kernel's warps spent their cycles."""
        buf = (C.c_uint64 * 5)()
        check(lib.b2c_engine_persist_profile(self._h, C.byref(buf)), "engine_persist_profile")
        return dict(zip(("wait_cycles", "tile_cycles", "total_cycles", "items", "waited_items"), list(buf)))

    def halo_profile(self, reset: bool = True) -> dict:
        """Wait-time counters of the halo protocol (SM cycles, summed over ring warps)."""
        from .._lib import DIRS

        buf = (C.c_uint64 * 40)()
        check(lib.b2c_engine_halo_profile(self._h, C.byref(buf), int(bool(reset))), "engine_halo_profile")
        v = list(buf)
        return {"data_wait_cycles": dict(zip(DIRS, v[0:8])), "ack_wait_cycles": dict(zip(DIRS, v[8:16])),
                "waits": v[16], "publish_cycles": v[17], "publishes": v[18],
                "data_wait_max": dict(zip(DIRS, v[24:32])), "ack_wait_max": dict(zip(DIRS, v[32:40]))}


_CACHE: "OrderedDict[tuple, Engine]" = OrderedDict()
_CACHE_MAX = 4


def clear_engine_cache():
    while _CACHE:
        _CACHE.popitem()[1].close()


def _cached_engine(w, h, ch, device, opts: dict) -> Engine:
    key = (w, h, ch, device, tuple(sorted(opts.items())))
    eng = _CACHE.get(key)
    if eng is None:
        eng = Engine(w, h, ch, device=device, **opts)
        _CACHE[key] = eng
        while len(_CACHE) > _CACHE_MAX:
            _CACHE.popitem(last=False)[1].close()
    else:
        _CACHE.move_to_end(key)
    return eng


def blur(img: torch.Tensor, repetitions: int, boundary: str = "zero", filter: str = "gaussian", tb: int = 0,
         kernel: str = "auto", out: torch.Tensor | None = None, return_iterations: bool = False, **opts):
    """Apply the 3x3 filter ``repetitions`` times on one GPU.

    ``img``: uint8 ``[H,W]`` (grey) or ``[H,W,3]`` (rgb, interleaved).  A CUDA tensor is
    processed in place on its device; a CPU tensor is copied host->device->host (pinned
    memory makes those copies asynchronous).  Returns a tensor like ``img``.
    """
    _require_cuda()
    w, h, ch = _img_meta(img)
    img = img.contiguous()
    device = img.device.index if img.is_cuda else torch.cuda.current_device()
    all_opts = dict(boundary=boundary, filter=filter, tb=tb, kernel=kernel, **opts)
    with torch.cuda.device(device):
        eng = _cached_engine(w, h, ch, device, all_opts)
        if out is None:
            out = torch.empty_like(img, pin_memory=(not img.is_cuda))
        eng.load(img)
        eng.exchange()
        eng.run(repetitions)
        iters = eng.iterations_done() if (return_iterations or all_opts.get("check_every", 0) or
                                          boundary.startswith("cuda")) else repetitions
        eng.store(out)
        if not out.is_cuda:
            torch.cuda.current_stream().synchronize()
    return (out, iters) if return_iterations else out


class BlurPipeline:
    """Throughput-oriented single-GPU API for a stream of same-shaped host images.

    ``depth`` engines, each with its own CUDA stream, work round-robin: while one image is
    being blurred, the next one is copied host->device and the previous result device->host
    (the copy engines run beside the SMs; PCIe is full duplex).  Per image the work is still
    H2D -> ``repetitions`` iterations -> D2H; only the waiting is overlapped.

        pipe = BlurPipeline(W, H, 3, repetitions=1000)
        for src, dst in zip(pinned_inputs, pinned_outputs):
            pipe.submit(src, dst)
        pipe.wait()
    """

    def __init__(self, width, height, channels, repetitions, depth=2, device=None, **opts):
        _require_cuda()
        self.repetitions = int(repetitions)
        self.device = torch.cuda.current_device() if device is None else int(device)
        with torch.cuda.device(self.device):
            self.engines = [Engine(width, height, channels, device=self.device, **opts) for _ in range(depth)]
            self.streams = [torch.cuda.Stream(device=self.device) for _ in range(depth)]
        self._next = 0

    def submit(self, src: torch.Tensor, dst: torch.Tensor):
        """Enqueue one image (pinned CPU or CUDA tensors); returns immediately."""
        i = self._next
        self._next = (i + 1) % len(self.engines)
        eng, st = self.engines[i], self.streams[i]
        with torch.cuda.device(self.device):
            eng.load(src, stream=st)
            eng.exchange(stream=st)
            eng.run(self.repetitions, stream=st)
            if int(eng.opts.check_every):
                eng.iterations_done()
            eng.store(dst, stream=st)

    def wait(self):
        for st in self.streams:
            st.synchronize()

    def close(self):
        self.wait()
        for e in self.engines:
            e.close()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()
        return False


def stencil_step(src: torch.Tensor, dst: torch.Tensor, rows: int, row_bytes: int, channels: int,
                 filter: str = "gaussian", threads: int = 0):
    """One plain iteration on caller-framed buffers, CUDA (``b2c_stencil_step_device``, current
    stream) or CPU (``b2c_stencil_step_host``, ``threads`` threads, 0 = all cores).

    ``src``/``dst``: uint8 ``[rows+2, pitch]`` with the block at ``[1:rows+1, channels:channels+row_bytes]``.
    """
    from .._lib import FILTER

    pitch = src.stride(0)
    off = pitch + channels
    if src.is_cuda:
        check(lib.b2c_stencil_step_device(src.data_ptr() + off, dst.data_ptr() + off, pitch, int(rows), int(row_bytes),
                                          int(channels), FILTER[filter], _stream_ptr()), "stencil_step")
    else:
        check(lib.b2c_stencil_step_host(src.data_ptr() + off, dst.data_ptr() + off, pitch, int(rows), int(row_bytes),
                                        int(channels), FILTER[filter], int(threads)), "stencil_step")


def load_torch_ops():
    """Load ``lib/_b200conv_torch.so`` (the thin C++ PyTorch binding) and return
    ``torch.ops.b200conv`` — ``torch.ops.b200conv.blur(img, reps, boundary, filter, tb, kernel)``."""
    from .._lib import PKG_DIR

    so = PKG_DIR / "lib" / "_b200conv_torch.so"
    if not so.exists():
        raise B2CError(f"{so} is not built; run `python build.py`")
    torch.ops.load_library(str(so))
    return torch.ops.b200conv


def blur_file(path: str, width: int, height: int, color: str, repetitions: int, **opts):
    """Whole-program run on one GPU: read ``path``, blur, write ``blur_<path>``.
    Returns ``(loop_ms, total_ms)``."""
    _require_cuda()
    ch = {"grey": 1, "rgb": 3}[color]
    o = make_opts(**opts)
    loop, total = C.c_double(), C.c_double()
    check(lib.b2c_blur_file(str(path).encode(), int(width), int(height), ch, int(repetitions), C.byref(o),
                            C.byref(loop), C.byref(total)), "blur_file")
    return loop.value, total.value
