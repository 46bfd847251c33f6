#!/usr/bin/env python3
"""Install and build the UNMODIFIED reference under baseline/_ref (git-ignored).

1. The prescribed ``pip install --target baseline/_ref /root/reference`` cannot work: the
   reference is three C/CUDA programs with no setup.py/pyproject.toml (outcome recorded in
   DESIGN.md).  So the tree is copied verbatim instead.
2. ``cuda/``: built with the reference's own Makefile (``make``), no flags changed.
3. ``mpi/`` and ``open-mp/``: no MPI exists on this image, so the unmodified sources are
   compiled against baseline/mpi_shim (threads as ranks) with ``-Dmain=ref_main``.
   The OpenMP file is built as documented (``-openmp`` is not a gcc OpenMP flag, so the
   pragmas are inert: plain build) and, separately and labelled as PATCHED, with
   ``collapse(3)`` -> ``collapse(2)`` and ``-fopenmp`` (the shipped file does not compile
   with OpenMP enabled).
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from pathlib import Path

HERE = Path(__file__).resolve().parent
REF = HERE / "_ref"
SRC = Path(os.environ.get("B200CONV_REFERENCE", "/root/reference"))
SHIM = HERE / "mpi_shim"


def sh(cmd, cwd=None):
    p = subprocess.run(cmd, cwd=cwd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    return p.returncode, p.stdout


def copy_tree():
    if (REF / "cuda" / "cuda_convolution.cu").exists():
        return "present"
    if not SRC.exists():
        return f"missing: {SRC}"
    REF.mkdir(parents=True, exist_ok=True)
    for item in SRC.iterdir():
        dst = REF / item.name
        if item.is_dir():
            shutil.copytree(item, dst, dirs_exist_ok=True)
        else:
            shutil.copy2(item, dst)
    return "copied"


def build_cuda():
    d = REF / "cuda"
    if (d / "cuda_convolution").exists():
        return 0, "present"
    env_path = "/usr/local/cuda/bin:" + os.environ.get("PATH", "")
    p = subprocess.run(["make"], cwd=d, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True,
                       env={**os.environ, "PATH": env_path})
    return p.returncode, p.stdout


def build_mpi():
    out = {}
    bdir = REF / "build"
    bdir.mkdir(exist_ok=True)
    shim = ["-I", str(SHIM), str(SHIM / "mpi_shim.c"), "-lpthread", "-Dmain=ref_main", "-w"]
    out["mpi"] = sh(["gcc", "-O2", "-o", str(bdir / "mpi_convolution"), str(REF / "mpi" / "mpi_convolution.c"), *shim])
    out["omp_as_documented"] = sh(["gcc", "-O2", "-o", str(bdir / "omp_convolution"),
                                   str(REF / "open-mp" / "omp_convolution.c"), *shim])
    patched = bdir / "omp_convolution_PATCHED_collapse2.c"
    patched.write_text((REF / "open-mp" / "omp_convolution.c").read_text().replace("collapse(3)", "collapse(2)"))
    out["omp_PATCHED_collapse2"] = sh(["gcc", "-O2", "-fopenmp", "-o", str(bdir / "omp_convolution_PATCHED"),
                                       str(patched), *shim])
    return out


def main():
    rc, log = sh([sys.executable, "-m", "pip", "install", "--no-index", "--no-build-isolation", "--find-links",
                  "/opt/wheelhouse", "--target", str(REF), str(SRC)])
    print(f"pip install: rc={rc}: {log.strip().splitlines()[-1] if log.strip() else ''}")
    print("copy:", copy_tree())
    rc, log = build_cuda()
    print(f"cuda make: rc={rc}")
    if rc:
        print(log)
    for name, (rc, log) in build_mpi().items():
        print(f"{name}: rc={rc}")
        if rc:
            print(log)


if __name__ == "__main__":
    main()
