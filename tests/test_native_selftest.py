"""The CPU half of the native library under AddressSanitizer + UndefinedBehaviorSanitizer.

tests/cpp/core_selftest.cpp drives every public CPU entry point (decomposition, CLI parser,
64-bit raw I/O incl. a sparse file beyond 2^31 bytes, the oracles, both host models of the halo
protocol) through its edge cases; the sanitizers turn silent out-of-bounds accesses, signed
overflow and leaks into failures ("race detection / sanitizers" — the reference has
none, and its int offsets and unchecked reads are exactly what such a build catches).
"""
import shutil
import subprocess
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
SOURCES = ["decomp.cpp", "cli.cpp", "rawio.cpp", "oracle.cpp", "halo_emul.cpp", "resident_emul.cpp"]


@pytest.mark.skipif(shutil.which("g++") is None, reason="needs g++")
@pytest.mark.parametrize("flags", [["-fsanitize=address,undefined", "-fno-sanitize-recover=undefined"], []],
                         ids=["asan-ubsan", "plain"])
def test_core_selftest(tmp_path, flags):
    exe = tmp_path / "core_selftest"
    cmd = ["g++", "-std=c++17", "-O1", "-g", "-Wall", "-Wextra", "-Werror", "-pthread", *flags,
           "-I", str(ROOT / "include"), "-I", str(ROOT / "csrc/core"), str(ROOT / "tests/cpp/core_selftest.cpp"),
           *[str(ROOT / "csrc/core" / s) for s in SOURCES], "-o", str(exe)]
    b = subprocess.run(cmd, capture_output=True, text=True)
    if b.returncode != 0 and flags and ("asan" in b.stderr.lower() or "ubsan" in b.stderr.lower()):
        pytest.skip("sanitizer runtime is not installed")
    assert b.returncode == 0, b.stderr[-4000:]
    r = subprocess.run([str(exe), str(tmp_path)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "0 failure(s)" in r.stdout, (r.stdout[-2000:], r.stderr[-4000:])
