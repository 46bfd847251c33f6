#!/usr/bin/env python3
"""Kernel-variant sweep on one GPU: Gpix-it/s for (tb, strip, chunk_rows) on the BASELINE.json shapes."""
import argparse
import json
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import b200conv  # noqa: E402


def bench(w, h, ch, reps, n=3, **opts):
    eng = b200conv.Engine(w, h, ch, **opts)
    blk = torch.randint(0, 256, eng.block_shape, dtype=torch.uint8, device="cuda")
    eng.load(blk)
    eng.exchange()
    for _ in range(2):
        eng.run(reps)
    torch.cuda.synchronize()
    best = 1e30
    for _ in range(n):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        eng.run(reps)
        e1.record()
        e1.synchronize()
        best = min(best, e0.elapsed_time(e1))
    d = eng.describe()
    eng.close()
    return w * h * reps / (best * 1e-3) / 1e9, best, d


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--shapes", default="16384x16384xrgb,16384x16384xgrey,1920x2520xgrey,1920x2520xrgb")
    ap.add_argument("--reps", type=int, default=200)
    ap.add_argument("--configs", default="1:16,2:16,4:16,4:8,8:8,4:4,8:4")
    ap.add_argument("--chunks", default="0")
    ap.add_argument("--kernel", default="stream")
    ap.add_argument("--boundary", default="zero")
    a = ap.parse_args()
    for shp in a.shapes.split(","):
        w, h, c = shp.split("x")
        w, h, ch = int(w), int(h), 3 if c == "rgb" else 1
        for cfg in a.configs.split(","):
            tb, strip = map(int, cfg.split(":"))
            for cr in map(int, a.chunks.split(",")):
                try:
                    v, ms, d = bench(w, h, ch, a.reps, tb=tb, strip_bytes=strip, chunk_rows=cr, kernel=a.kernel,
                                     boundary=a.boundary)
                    print(json.dumps({"shape": shp, "tb": tb, "strip": strip, "chunk_rows": cr, "gpix_it_s": round(v, 1),
                                      "byte_it_T": round(v * ch / 1e3, 3), "ms": round(ms, 3), "plan": d}), flush=True)
                except Exception as e:  # noqa: BLE001
                    print(json.dumps({"shape": shp, "tb": tb, "strip": strip, "error": str(e)[:200]}), flush=True)
