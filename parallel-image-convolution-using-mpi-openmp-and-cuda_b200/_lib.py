"""ctypes binding of ``lib/libb200conv_core.so`` (declared in ``include/b200conv.h``).

The library is built in-tree by ``build.py``.  If it is missing and ``nvcc`` is on the
machine it is built on first import; otherwise importing fails loudly — there is no
Python fallback for the GPU path.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import sys
from pathlib import Path

PKG_DIR = Path(__file__).resolve().parent
ROOT = PKG_DIR.parent
# B200CONV_DEBUG_LIB=1: the -DB2C_DEBUG build (python build.py --debug) with address / epoch checks in the kernels
LIB_PATH = PKG_DIR / "lib" / ("libb200conv_core_debug.so" if os.environ.get("B200CONV_DEBUG_LIB", "0") not in ("", "0")
                              else "libb200conv_core.so")
LAUNCHER_PATH = PKG_DIR / "lib" / "b200conv"
HANDLE_BYTES = 256

ZERO, TORUS, CUDA_COMPAT = 0, 1, 2
GAUSSIAN, BOX, EDGE = 0, 1, 2
KERNEL_AUTO, KERNEL_NAIVE, KERNEL_STREAM, KERNEL_TMA, KERNEL_RESIDENT = 0, 1, 2, 3, 4
BOUNDARY = {"zero": ZERO, "torus": TORUS, "cuda-compat": CUDA_COMPAT, "cuda_compat": CUDA_COMPAT}
FILTER = {"gaussian": GAUSSIAN, "box": BOX, "edge": EDGE}
KERNEL = {"auto": KERNEL_AUTO, "naive": KERNEL_NAIVE, "stream": KERNEL_STREAM, "tma": KERNEL_TMA,
          "resident": KERNEL_RESIDENT}
DIRS = ("N", "S", "W", "E", "NW", "NE", "SW", "SE")


class Opts(C.Structure):
    _fields_ = [(n, C.c_int) for n in
                ("boundary", "filter", "kernel", "tb", "strip_bytes", "chunk_rows", "use_graph", "check_every",
                 "halo_depth", "persistent")]


class EmulOpts(C.Structure):
    _fields_ = [("tb", C.c_int), ("check_every", C.c_int), ("max_delay_us", C.c_int), ("seed", C.c_uint),
                ("timeout_ms", C.c_int), ("single_slot_set", C.c_int), ("stall_rank", C.c_int), ("stall_ms", C.c_int), ("images", C.c_int),
                ("halo_depth", C.c_int), ("courier", C.c_int)]


class Cli(C.Structure):
    _fields_ = [("image", C.c_char_p), ("width", C.c_int), ("height", C.c_int), ("repetitions", C.c_int),
                ("channels", C.c_int), ("gpus", C.c_int), ("boundary", C.c_int), ("filter", C.c_int),
                ("tb", C.c_int), ("kernel", C.c_int), ("check_every", C.c_int), ("checkpoint_every", C.c_int),
                ("profile", C.c_int), ("json", C.c_char_p), ("cpu", C.c_int), ("halo_depth", C.c_int)]


class B2CError(RuntimeError):
    pass


def _build():
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    if not Path(nvcc).exists():
        raise ImportError(f"{LIB_PATH} is missing and nvcc was not found; run `python build.py`")
    subprocess.run([sys.executable, str(ROOT / "build.py"), "--no-torch"], check=True)


def _load() -> C.CDLL:
    if not LIB_PATH.exists():
        _build()
    lib = C.CDLL(str(LIB_PATH), mode=C.RTLD_GLOBAL)
    P, I, L, V = C.c_void_p, C.c_int, C.c_int64, C.c_void_p
    u8p = C.c_void_p
    sig = {
        "b2c_last_error": (C.c_char_p, []),
        "b2c_version": (C.c_char_p, []),
        "b2c_rows_division": (I, [I, I, I]),
        "b2c_decompose": (I, [I, I, I, C.POINTER(I), C.POINTER(I)]),
        "b2c_neighbours": (I, [I, I, I, I, C.POINTER(I * 8)]),
        "b2c_parse_cli": (I, [I, C.POINTER(C.c_char_p), C.POINTER(Cli)]),
        "b2c_usage_text": (I, [C.c_char_p, I, C.c_char_p, C.c_size_t]),
        "b2c_output_name": (I, [C.c_char_p, C.c_char_p, C.c_size_t]),
        "b2c_read_block": (L, [C.c_char_p, u8p, L, I, I, I, I, I, I, I]),
        "b2c_write_block": (L, [C.c_char_p, u8p, L, I, I, I, I, I, I, I, I]),
        "b2c_oracle_blur": (I, [u8p, u8p, I, I, I, I, I, I]),
        "b2c_oracle_blur_sharded": (I, [u8p, u8p, I, I, I, I, I, I, I, I]),
        "b2c_emulate_halo_threads": (I, [u8p, u8p, I, I, I, I, I, I, I, C.POINTER(EmulOpts), C.POINTER(I)]),
        "b2c_stencil_step_host": (I, [u8p, u8p, L, I, I, I, I, I]),
        "b2c_blur_host": (I, [u8p, u8p, I, I, I, I, I, I, I]),
        "b2c_default_opts": (None, [C.POINTER(Opts)]),
        "b2c_engine_create": (P, [I, I, I, I, I, I, I, C.POINTER(Opts)]),
        "b2c_engine_destroy": (None, [P]),
        "b2c_engine_block": (I, [P, C.POINTER(I), C.POINTER(I), C.POINTER(I), C.POINTER(I)]),
        "b2c_engine_export": (I, [P, V]),
        "b2c_engine_import": (I, [P, V, I]),
        "b2c_engine_connect_local": (I, [C.POINTER(P), I]),
        "b2c_engine_disconnect": (I, [P]),
        "b2c_engine_upload": (I, [P, u8p, L, V]),
        "b2c_engine_download": (I, [P, u8p, L, V]),
        "b2c_engine_load_device": (I, [P, u8p, L, V]),
        "b2c_engine_store_device": (I, [P, u8p, L, V]),
        "b2c_engine_exchange": (I, [P, V]),
        "b2c_engine_run": (I, [P, I, V]),
        "b2c_engine_iterations_done": (I, [P]),
        "b2c_engine_plan_run": (I, [P, I]),
        "b2c_engine_run_pass": (I, [P, I, V]),
        "b2c_engine_tune_count": (I, [P]),
        "b2c_engine_tune_time": (C.c_double, [P, I, V]),
        "b2c_engine_tune_set": (I, [P, I]),
        "b2c_engine_barrier": (I, [P, V]),
        "b2c_engine_set_nvls": (I, [P, V, V, I]),
        "b2c_engine_barrier_wait_cycles": (L, [P]),
        "b2c_engine_set_profile": (I, [P, I]),
        "b2c_engine_halo_profile": (I, [P, C.POINTER(C.c_uint64 * 40), I]),
        "b2c_engine_persist_profile": (I, [P, C.POINTER(C.c_uint64 * 5)]),
        "b2c_debug_pass_plans": (I, [I, I, I, I, I, I, I, I, I, I, C.POINTER(I), I]),
        "b2c_engine_launches": (L, [P]),
        "b2c_resident_plan": (I, [I, I, I, I, I, C.POINTER(I)]),
        "b2c_debug_map_work": (I, [I, I, I, I, I, C.POINTER(I), C.POINTER(I)]),
        "b2c_debug_band_plan": (I, [I, I, I, I, I, I, I, I, I, C.POINTER(I)]),
        "b2c_emulate_persist_threads": (I, [u8p, u8p, I, I, I, I, I, I, I, I, I, C.c_uint]),
        "b2c_emulate_resident_threads": (I, [u8p, u8p, I, I, I, I, I, I, I, I, C.c_uint]),
        "b2c_engine_describe": (I, [P, C.c_char_p, C.c_size_t]),
        "b2c_engine_set_delay": (I, [P, L]),
        "b2c_blur_device": (I, [u8p, L, u8p, L, I, I, I, I, C.POINTER(Opts), V]),
        "b2c_stencil_step_device": (I, [u8p, u8p, L, I, I, I, I, V]),
        "b2c_blur_file": (I, [C.c_char_p, I, I, I, I, C.POINTER(Opts), C.POINTER(C.c_double), C.POINTER(C.c_double)]),
        "b2c_total_launches": (L, []),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)
        fn.restype, fn.argtypes = res, args
    return lib


lib = _load()


def last_error() -> str:
    return (lib.b2c_last_error() or b"").decode(errors="replace")


def check(rc: int, what: str = "b200conv") -> int:
    if rc is None or rc < 0:
        raise B2CError(f"{what}: {last_error()} (code {rc})")
    return rc


def make_opts(boundary="zero", filter="gaussian", kernel="auto", tb=0, strip_bytes=0, chunk_rows=0,
              use_graph=-1, check_every=0, halo_depth=0, persistent=-1) -> Opts:
    o = Opts()
    lib.b2c_default_opts(C.byref(o))
    o.boundary = BOUNDARY[boundary] if isinstance(boundary, str) else int(boundary)
    o.filter = FILTER[filter] if isinstance(filter, str) else int(filter)
    o.kernel = KERNEL[kernel] if isinstance(kernel, str) else int(kernel)
    o.tb, o.strip_bytes, o.chunk_rows = int(tb), int(strip_bytes), int(chunk_rows)
    o.use_graph, o.check_every, o.halo_depth = int(use_graph), int(check_every), int(halo_depth)
    o.persistent = int(persistent)
    return o
