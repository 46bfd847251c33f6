#!/usr/bin/env python3
"""Rows per warp x {persistent kernel, one launch per pass} on one GPU.
python scripts/persist_sweep.py --width 8192 --height 4096 --color rgb [--reps 1000]"""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200conv  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--width", type=int, default=8192)
ap.add_argument("--height", type=int, default=4096)
ap.add_argument("--color", default="rgb")
ap.add_argument("--reps", type=int, default=1000)
ap.add_argument("--strip", type=int, default=16)
ap.add_argument("--tb", type=int, default=4)
ap.add_argument("--rows", default="24,32,40,48,64,80,96,128,160,192,256")
a = ap.parse_args()
ch = 3 if a.color == "rgb" else 1
img = torch.randint(0, 256, (a.height, a.width) + ((3,) if ch == 3 else ()), dtype=torch.uint8, device="cuda")
print(f"{a.width}x{a.height} {a.color} reps={a.reps} strip={a.strip} tb={a.tb}")
print("rows   persistent   per-pass   (ms per run)")
for cr in [int(x) for x in a.rows.split(",")]:
    res = []
    for pm in (1, 0):
        eng = b200conv.Engine(a.width, a.height, ch, tb=a.tb, strip_bytes=a.strip, chunk_rows=cr, persistent=pm)
        eng.load(img)
        eng.exchange()
        st = torch.cuda.Stream()
        with torch.cuda.stream(st):
            eng.run(a.reps)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(3):
                eng.run(a.reps)
            e1.record()
        torch.cuda.synchronize()
        res.append(e0.elapsed_time(e1) / 3)
        if pm:
            pp = eng.persist_profile()
        eng.close()
    extra = ""
    if pp["items"]:
        tot = max(1, pp["total_cycles"])
        extra = (f"   wait {100 * pp['wait_cycles'] / tot:5.1f}%  tile {100 * pp['tile_cycles'] / tot:5.1f}%  "
                 f"cyc/item {pp['tile_cycles'] // pp['items']:7d}  waited {100 * pp['waited_items'] / pp['items']:5.1f}%")
    print(f"{cr:4d}   {res[0]:10.3f} {res[1]:10.3f}{extra}", flush=True)
