"""Python entry to the native launcher ``lib/b200conv`` (same CLI as the reference:
``<prog> image.raw width height repetitions grey|rgb``, README.md:16-34)."""
from __future__ import annotations

import ctypes as C
import subprocess
import sys

from .._lib import LAUNCHER_PATH, Cli, lib


def parse_cli(argv):
    """-> dict of parsed fields, or None when the reference would print its usage error."""
    arr = (C.c_char_p * len(argv))(*[a.encode() for a in argv])
    out = Cli()
    if lib.b2c_parse_cli(len(argv), arr, C.byref(out)) != 0:
        return None
    d = {n: getattr(out, n) for n, _ in Cli._fields_}
    d["image"] = d["image"].decode() if d["image"] else None
    d["json"] = d["json"].decode() if d["json"] else None
    return d


def usage_text(argv0: str, style: int = 0) -> str:
    buf = C.create_string_buffer(1024)
    lib.b2c_usage_text(argv0.encode(), style, buf, 1024)
    return buf.value.decode()


def run_launcher(args, cwd=None, **kw):
    if not LAUNCHER_PATH.exists():
        raise FileNotFoundError(f"{LAUNCHER_PATH} is not built; run `python build.py`")
    return subprocess.run([str(LAUNCHER_PATH), *map(str, args)], cwd=cwd, capture_output=True, text=True, **kw)


def main(argv=None):
    argv = sys.argv[1:] if argv is None else argv
    p = subprocess.run([str(LAUNCHER_PATH), *argv])
    return p.returncode


if __name__ == "__main__":
    sys.exit(main())
