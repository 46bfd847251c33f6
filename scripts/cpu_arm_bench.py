"""CPU arm (GlooHaloBlur: ranks over gloo x threads per rank) on the plumbing config of
BASELINE.json, next to the reference's own MPI program built against baseline/mpi_shim:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node=4 --master-addr 127.0.0.1 \
        scripts/cpu_arm_bench.py [--width 1920 --height 2520 --color rgb --reps 40 --threads 2]
Prints one JSON line on rank 0 (loop seconds = max over ranks, like mpi_convolution.c:264-275).
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200conv  # noqa: E402
from b200conv.parallel import GlooHaloBlur  # noqa: E402
from b200conv.utils import waterfall_image  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=2520)
    ap.add_argument("--color", default="rgb", choices=["grey", "rgb"])
    ap.add_argument("--reps", type=int, default=40)
    ap.add_argument("--threads", type=int, default=2)
    ap.add_argument("--halo-depth", type=int, default=1, help="iterations between two halo exchanges")
    ap.add_argument("--check", action="store_true")
    a = ap.parse_args()
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    ch = 3 if a.color == "rgb" else 1
    img = waterfall_image(a.height, a.width, ch, seed=0)
    eng = GlooHaloBlur(a.width, a.height, ch, threads=a.threads, halo_depth=a.halo_depth)
    blk = torch.from_numpy(np.ascontiguousarray(img[eng.r0:eng.r0 + eng.R, eng.c0:eng.c0 + eng.C]))
    eng.load(blk)
    eng.run(2)                       # warm-up
    eng.load(blk)
    dist.barrier()                   # MPI_Barrier, mpi_convolution.c:151
    t0 = time.perf_counter()
    eng.run(a.reps)
    sec = torch.tensor([time.perf_counter() - t0], dtype=torch.float64)
    dist.all_reduce(sec, op=dist.ReduceOp.MAX)
    ok = None
    if a.check:
        out = eng.store().contiguous()
        parts = [torch.empty_like(out) for _ in range(world)]
        dist.all_gather(parts, out)
        if rank == 0:
            full = np.empty_like(img)
            for k, part in enumerate(parts):
                a0, b0, r, c = b200conv.block_of(k, eng.grid, a.height, a.width)
                full[a0:a0 + r, b0:b0 + c] = part.numpy()
            ok = bool((full == b200conv.reference_oracle(img, a.reps, "zero")).all())
    if rank == 0:
        s = float(sec.item())
        print(json.dumps({"impl": "b200conv cpu arm (gloo ranks x threads)", "ranks": world, "threads_per_rank": a.threads, "halo_depth": eng.F, "exchanges": eng.exchanges,
                          "grid": list(eng.grid), "width": a.width, "height": a.height, "color": a.color,
                          "repetitions": a.reps, "loop_seconds": s,
                          "gpix_it_per_s": a.width * a.height * a.reps / s / 1e9, "equals_oracle": ok,
                          "host_cores": os.cpu_count()}), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
