#!/usr/bin/env python3
"""In-tree build of every native artefact (the reference's Makefile, cuda/Makefile:1-14,
and README one-liners README.md:18,27,32 are the counterpart).

  lib/libb200conv_core.so   C++/CUDA core, sm_100a only, no torch dependency
  lib/b200conv              C launcher (fork one worker per GPU)
  lib/_b200conv_torch.so    thin PyTorch binding over the core

Objects are cached under build/ by content hash; `python build.py --force` rebuilds all.
Everything lands inside the package directory so it travels to the GPU box with the
repo snapshot (built files are git-ignored).
"""
from __future__ import annotations

import argparse
import concurrent.futures as cf
import hashlib
import os
import shlex
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent
PKG = ROOT / "parallel-image-convolution-using-mpi-openmp-and-cuda_b200"
LIB = PKG / "lib"
BUILD = ROOT / "build"
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
INCLUDES = ["-I", str(ROOT / "include"), "-I", str(ROOT / "csrc/core"), "-I", str(ROOT / "csrc/cuda")]
NVCC_FLAGS = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "-Xcompiler", "-ffp-contract=off",
              "-Xptxas", "-v", "--expt-relaxed-constexpr"] + ARCH + INCLUDES

CORE_SOURCES = [
    "csrc/core/decomp.cpp", "csrc/core/cli.cpp", "csrc/core/rawio.cpp", "csrc/core/oracle.cpp", "csrc/core/halo_emul.cpp", "csrc/core/resident_emul.cpp", "csrc/core/persist_emul.cpp",
    "csrc/cuda/kernels.cu", "csrc/cuda/stencil_tma.cu", "csrc/cuda/stencil_resident.cu", "csrc/cuda/engine.cu", "csrc/cuda/program.cu",
]


def run(cmd, log=None):
    p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if log is not None:
        log.write_text(p.stdout)
    if p.returncode != 0:
        sys.stderr.write(" ".join(shlex.quote(str(c)) for c in cmd) + "\n" + p.stdout)
        raise SystemExit(f"build failed: {cmd[0]}")
    return p.stdout


def digest(path: Path, extra: str) -> str:
    h = hashlib.sha256(extra.encode())
    h.update(path.read_bytes())
    for hdr in sorted(list((ROOT / "include").glob("*.h")) + list((ROOT / "csrc").rglob("*.h")) +
                      list((ROOT / "csrc").rglob("*.cuh"))):
        h.update(hdr.read_bytes())
    return h.hexdigest()[:16]


def compile_one(src: str, force: bool, extra=(), bdir: Path = None) -> Path:
    bdir = bdir or BUILD
    sp = ROOT / src
    flags = [*NVCC_FLAGS, *extra]
    tag = digest(sp, " ".join(flags))
    obj = bdir / (sp.stem + "." + tag + ".o")
    if obj.exists() and not force:
        return obj
    for old in bdir.glob(sp.stem + ".*.o"):
        old.unlink()
    run([NVCC, *flags, "-c", str(sp), "-o", str(obj)], log=bdir / (sp.stem + ".ptxas.log"))
    return obj


def build_core(force=False, debug=False) -> Path:
    """debug=True: lib/libb200conv_core_debug.so with -DB2C_DEBUG — every global / peer address the
    streaming kernels form is checked against its allocation and epochs are checked for running
    ahead (stencil_stream.cuh); load it with B200CONV_DEBUG_LIB=1.  Stands in for compute-sanitizer."""
    bdir = BUILD / "debug" if debug else BUILD
    bdir.mkdir(parents=True, exist_ok=True)
    LIB.mkdir(parents=True, exist_ok=True)
    srcs = [s for s in CORE_SOURCES if (ROOT / s).exists()]
    extra = ("-DB2C_DEBUG",) if debug else ()
    with cf.ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        objs = list(ex.map(lambda s: compile_one(s, force, extra, bdir), srcs))
    out = LIB / ("libb200conv_core_debug.so" if debug else "libb200conv_core.so")
    stamp = bdir / "core.link"
    key = " ".join(o.name for o in objs)
    if force or not out.exists() or not stamp.exists() or stamp.read_text() != key:
        run([NVCC, "-shared", *ARCH, "-o", str(out), *map(str, objs), "-lpthread", "-ldl", "-lrt"])
        stamp.write_text(key)
    return out


def build_launcher(force=False) -> Path:
    src = ROOT / "csrc/launcher/main.cpp"
    out = LIB / "b200conv"
    if not src.exists():
        return out
    tag = digest(src, "launcher")
    stamp = BUILD / "launcher.stamp"
    core = LIB / "libb200conv_core.so"
    key = tag + str(core.stat().st_mtime_ns)
    if force or not out.exists() or not stamp.exists() or stamp.read_text() != key:
        run([NVCC, "-O2", "-std=c++17", *ARCH, *INCLUDES, str(src), "-o", str(out),
             "-L", str(LIB), "-lb200conv_core", "-Xlinker", "-rpath", "-Xlinker", "$ORIGIN", "-lpthread"])
        stamp.write_text(key)
    return out


def build_torch_ext(force=False) -> Path:
    src = ROOT / "csrc/torch/binding.cpp"
    out = LIB / "_b200conv_torch.so"
    if not src.exists():
        return out
    import sysconfig

    import torch
    from torch.utils import cpp_extension as ce

    tag = digest(src, "torch" + torch.__version__)
    stamp = BUILD / "torch.stamp"
    core = LIB / "libb200conv_core.so"
    key = tag + str(core.stat().st_mtime_ns)
    if not (force or not out.exists() or not stamp.exists() or stamp.read_text() != key):
        return out
    inc = []
    for p in ce.include_paths() + [sysconfig.get_paths()["include"], "/usr/local/cuda/include"]:
        inc += ["-I", p]
    tlib = str(Path(torch.__file__).parent / "lib")
    abi = int(torch._C._GLIBCXX_USE_CXX11_ABI)
    run(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", f"-D_GLIBCXX_USE_CXX11_ABI={abi}",
         "-DTORCH_EXTENSION_NAME=_b200conv_torch", "-DTORCH_API_INCLUDE_EXTENSION_H",
         *INCLUDES, *inc, str(src), "-o", str(out),
         "-L", str(LIB), "-lb200conv_core", "-L", tlib, "-ltorch", "-ltorch_cpu", "-ltorch_python", "-lc10",
         "-lc10_cuda", "-ltorch_cuda", "-L", "/usr/local/cuda/lib64", "-lcudart",
         f"-Wl,-rpath,$ORIGIN", f"-Wl,-rpath,{tlib}"])
    stamp.write_text(key)
    return out


def build_all(force=False, torch_ext=True):
    core = build_core(force)
    launcher = build_launcher(force)
    ext = build_torch_ext(force) if torch_ext else None
    return core, launcher, ext


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--no-torch", action="store_true")
    ap.add_argument("--debug", action="store_true", help="also build lib/libb200conv_core_debug.so (-DB2C_DEBUG)")
    a = ap.parse_args()
    if a.debug:
        print(build_core(a.force, debug=True), "ok")
    for p in build_all(a.force, not a.no_torch):
        if p is not None:
            print(p, "ok" if p.exists() else "skipped")
