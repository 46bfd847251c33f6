#!/usr/bin/env python3
"""Short single-GPU workloads for `ncu` (one kernel family each).  python scripts/ncu_case.py CASE
  persist   stream_persist_kernel   8192x16384 rgb block, 4 passes in one launch
  perpass   stream_kernel           same block, one launch per pass
  tma       stream_tma_kernel       same block, TMA feed
  general   stream_kernel           160-pixel-wide torus image: every warp is a ring warp (masks + pushes)
  resident  resident_kernel         1920x2520 grey, 1000 iterations in registers
  launches  every kernel of a torus run with convergence checks (push, converge, passes)"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200conv  # noqa: E402

case = sys.argv[1]
cfg = {
    "persist": dict(w=16384, h=8192, ch=3, reps=16, opts=dict(kernel="stream", persistent=1, chunk_rows=160, tb=4)),
    "perpass": dict(w=16384, h=8192, ch=3, reps=16, opts=dict(kernel="stream", persistent=0, use_graph=0, chunk_rows=160, tb=4)),
    "tma": dict(w=16384, h=8192, ch=3, reps=16, opts=dict(kernel="tma", use_graph=0, chunk_rows=160, tb=4)),
    "general": dict(w=160, h=16384, ch=3, reps=16, opts=dict(kernel="stream", boundary="torus", persistent=0, use_graph=0, tb=4)),
    "resident": dict(w=1920, h=2520, ch=1, reps=1000, opts=dict(kernel="resident")),
    "launches": dict(w=2048, h=2048, ch=1, reps=200, opts=dict(kernel="stream", boundary="torus", persistent=0, use_graph=0,
                                                             tb=4, check_every=20)),
}[case]
shape = (cfg["h"], cfg["w"]) + ((3,) if cfg["ch"] == 3 else ())
img = torch.randint(0, 256, shape, dtype=torch.uint8, device="cuda")
eng = b200conv.Engine(cfg["w"], cfg["h"], cfg["ch"], **cfg["opts"])
print(eng.describe(), flush=True)
st = torch.cuda.Stream()
with torch.cuda.stream(st):
    eng.load(img)
    eng.exchange()
    eng.run(cfg["reps"])     # warm-up (tables, module load)
    if case == "launches":
        eng.iterations_done()
        eng.load(img)
        eng.exchange()
    eng.run(cfg["reps"])
torch.cuda.synchronize()
eng.close()
print("done", flush=True)
