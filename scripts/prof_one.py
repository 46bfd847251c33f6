#!/usr/bin/env python3
"""Tiny driver for ncu: a few passes of one kernel variant on one shape."""
import argparse
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import b200conv  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--shape", default="16384x16384xrgb")
ap.add_argument("--tb", type=int, default=4)
ap.add_argument("--strip", type=int, default=16)
ap.add_argument("--chunk-rows", type=int, default=0)
ap.add_argument("--reps", type=int, default=16)
ap.add_argument("--kernel", default="stream")
a = ap.parse_args()
w, h, c = a.shape.split("x")
w, h, ch = int(w), int(h), 3 if c == "rgb" else 1
eng = b200conv.Engine(w, h, ch, tb=a.tb, strip_bytes=a.strip, chunk_rows=a.chunk_rows, kernel=a.kernel, use_graph=0)
eng.load(torch.randint(0, 256, eng.block_shape, dtype=torch.uint8, device="cuda"))
eng.run(a.reps)
torch.cuda.synchronize()
print(eng.describe())
eng.close()
