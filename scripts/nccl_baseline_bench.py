#!/usr/bin/env python3
"""NCCL send/recv halo baseline vs the fused peer-memory engine, same shape, same GPUs.
torchrun --nproc-per-node N scripts/nccl_baseline_bench.py [--width W --height H --color rgb --reps 50]"""
import argparse
import json
import os
import sys
from pathlib import Path

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import b200conv  # noqa: E402
from b200conv.parallel.sendrecv import NcclHaloBlur  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--width", type=int, default=16384)
ap.add_argument("--height", type=int, default=16384)
ap.add_argument("--color", default="rgb")
ap.add_argument("--reps", type=int, default=40)
ap.add_argument("--check", action="store_true")
a = ap.parse_args()
local = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
rank, world = dist.get_rank(), dist.get_world_size()
ch = 3 if a.color == "rgb" else 1


def timed(fn):
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); fn(); e1.record(); e1.synchronize()
    t = torch.tensor([e0.elapsed_time(e1)], device="cuda", dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


base = NcclHaloBlur(a.width, a.height, ch)
blk = torch.randint(0, 256, (base.R, base.C) if ch == 1 else (base.R, base.C, ch), dtype=torch.uint8, device="cuda",
                    generator=torch.Generator(device="cuda").manual_seed(rank))
base.load(blk)
base.run(3)
ms_nccl = timed(lambda: base.run(a.reps))
out_nccl = base.store()

res = {}
for name, opts in (("fused_k1", dict(tb=1)), ("fused_k4", dict(tb=4)), ("fused_tma_k4", dict(tb=4, kernel="tma"))):
    eng = b200conv.ShardedBlur(a.width, a.height, ch, **opts)
    eng.load(blk)
    eng.run(3 + (1 if opts["tb"] == 4 else 0))
    ms = eng.timed_run(a.reps)
    if a.check and name == "fused_k1":
        out = torch.empty_like(blk)
        eng.store(out); torch.cuda.synchronize()
        same = bool((out.reshape(out_nccl.shape) == out_nccl).all())
        res["outputs_equal"] = same
    res[name] = a.width * a.height * a.reps / (ms * 1e-3) / 1e9
    eng.close()
if rank == 0:
    res["nccl_sendrecv_naive"] = a.width * a.height * a.reps / (ms_nccl * 1e-3) / 1e9
    res.update(gpus=world, shape=f"{a.width}x{a.height}x{a.color}", reps=a.reps, unit="Gpix-it/s")
    print(json.dumps(res), flush=True)
dist.destroy_process_group()
