"""Worker for test_cpu_dist.py: ranks are CPU processes under torch.distributed.run, backend gloo.
Each rank blurs its block with GlooHaloBlur (send/recv halo exchange + threaded host stencil);
rank 0 checks the gathered image against the 1-rank oracle."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200conv  # noqa: E402
from b200conv.parallel import GlooHaloBlur, block_of  # noqa: E402


def main():
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    ok = True
    # last field: halo depth F (one F-deep exchange per F iterations, the halo recomputed in between)
    cases = [(48, 64, 1, 7, "zero", 1, 1), (48, 64, 3, 5, "torus", 2, 1), (60, 36, 3, 9, "zero", 2, 1),
             (24, 24, 1, 12, "torus", 1, 1), (48, 64, 3, 23, "zero", 2, 4), (48, 64, 1, 17, "torus", 1, 8),
             (60, 36, 3, 10, "torus", 2, 3), (24, 24, 1, 9, "zero", 1, 100)]
    for (H, W, ch, reps, boundary, threads, depth) in cases:
        img = np.random.default_rng(3).integers(0, 256, (H, W) if ch == 1 else (H, W, ch), dtype=np.uint8)
        try:
            grid = b200conv.decompose(world, H, W)
        except b200conv.B2CError:
            continue
        eng = GlooHaloBlur(W, H, ch, boundary=boundary, threads=threads, halo_depth=depth)
        assert eng.grid == grid and eng.F == max(1, min(depth, eng.R, eng.C))
        eng.load(torch.from_numpy(np.ascontiguousarray(img[eng.r0:eng.r0 + eng.R, eng.c0:eng.c0 + eng.C])))
        eng.run(reps)
        out = eng.store().contiguous()
        parts = [torch.empty_like(out) for _ in range(world)]
        dist.all_gather(parts, out)
        if rank == 0:
            full = np.empty_like(img)
            for k, part in enumerate(parts):
                a0, b0, a, b = block_of(k, grid, H, W)
                full[a0:a0 + a, b0:b0 + b] = part.numpy()
            good = bool((full == b200conv.reference_oracle(img, reps, boundary)).all())
            good = good and eng.exchanges == -(-reps // eng.F)
            print(f"case {H}x{W}x{ch} reps={reps} {boundary} threads={threads} halo_depth={eng.F} exchanges={eng.exchanges} "
                  f"grid={grid} {'OK' if good else 'MISMATCH'}",
                  flush=True)
            ok = ok and good
    flag = torch.tensor([1 if ok else 0])
    dist.broadcast(flag, 0)
    dist.destroy_process_group()
    sys.exit(0 if int(flag.item()) else 1)


if __name__ == "__main__":
    main()
