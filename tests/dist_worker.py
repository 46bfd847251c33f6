"""Worker for test_gpu_dist.py: one rank per GPU under torch.distributed.run.
Each rank blurs its block with the peer-memory halo engine; rank 0 checks the gathered
image against the 1-rank CPU oracle."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200conv  # noqa: E402
from b200conv.parallel import block_of  # noqa: E402


def main():
    local = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
    rank, world = dist.get_rank(), dist.get_world_size()
    cases = [
        # (H, W, ch, reps, boundary, opts)
        (256, 512, 3, 11, "zero", dict(tb=4)),   # groups 4, 4, 2+1: the acknowledging pass is a k=2 kernel first used mid-run
        (256, 512, 3, 23, "torus", dict()),
        (256, 512, 1, 23, "torus", dict(tb=4)),
        (240, 200, 3, 9, "zero", dict(tb=2)),          # unaligned block width -> fallback push kernel
        (512, 1024, 3, 40, "torus", dict(tb=4, kernel="tma")),
        (128, 256, 1, 6, "zero", dict(kernel="naive")),
        (256, 256, 1, 400, "torus", dict(tb=4, check_every=20)),   # fused convergence across ranks
        # convergence on the fallback push path (unaligned width), and the same engine reloaded and
        # run again: the exchange after a converged run must not be skipped by the stale latch
        (120, 100, 1, 1000, "torus", dict(tb=2, check_every=20, twice=True)),
        (128, 512, 3, 1000, "torus", dict(tb=4, delay=20000)),      # 250 epochs with skewed ranks
        # one halo exchange per group of passes (deep halo, ack epochs), skewed ranks, graph replay
        (256, 512, 3, 203, "torus", dict(tb=4, halo_depth=8, delay=30000)),
        (256, 1024, 1, 330, "zero", dict(tb=4, halo_depth=16, delay=10000, twice=True)),
        (512, 1024, 3, 1000, "torus", dict(tb=4, halo_depth=32, use_graph=1, twice=True)),
        # the persistent kernel with neighbours: its own courier warp, tail passes (203 = 25 groups of 8 + 2 + 1)
        (512, 1024, 3, 300, "torus", dict(tb=4, halo_depth=16, persistent=1, twice=True)),
        (256, 512, 3, 203, "zero", dict(tb=4, halo_depth=8, persistent=1, delay=20000)),
        (256, 512, 1, 90, "torus", dict(tb=4, halo_depth=12, persistent=1)),    # groups 4+4+4, last group 4+2
        (240, 200, 3, 57, "torus", dict(tb=4, halo_depth=12)),      # unaligned width: push kernel after each group
        (512, 1024, 3, 90, "zero", dict(tb=4, halo_depth=16, kernel="tma")),
        (512, 512, 1, 800, "torus", dict(tb=4, halo_depth=16, check_every=20)),
        (128, 512, 3, 501, "zero", dict(tb=2, delay=50000)),
    ]
    ok = True
    case_no = 0
    for (H, W, ch, reps, boundary, opts) in cases:
        rng = np.random.default_rng(7)
        img = rng.integers(0, 256, (H, W) if ch == 1 else (H, W, ch), dtype=np.uint8)
        try:
            grid = b200conv.decompose(world, H, W)
        except b200conv.B2CError:
            continue
        delay = opts.pop("delay", 0)
        twice = opts.pop("twice", False)
        eng = b200conv.ShardedBlur(W, H, ch, boundary=boundary, **opts)
        if delay:
            # T3 stress: skew the ranks against each other before every flag wait so a
            # slot-reuse race (a frame overwritten while its owner still reads it) would corrupt bytes
            eng.engine.set_delay(delay * (1 + (rank * 7) % 5))
        r0, c0, r, c = eng.block
        blk = torch.from_numpy(np.ascontiguousarray(img[r0:r0 + r, c0:c0 + c])).cuda()
        # every other case runs on a stream of its own: only there can a run be replayed as a CUDA
        # graph and have a courier beside it (the legacy default stream allows neither)
        own = torch.cuda.Stream() if case_no % 2 == 0 else torch.cuda.current_stream()
        case_no += 1
        torch.cuda.synchronize()
        with torch.cuda.stream(own):
            for _ in range(2 if twice else 1):
                eng.load(blk)
                ms = eng.timed_run(reps)
                iters = eng.iterations_done()
            out = torch.empty_like(blk)
            eng.store(out)
        torch.cuda.synchronize()
        parts = [torch.empty_like(out) for _ in range(world)]
        dist.all_gather(parts, out)
        if rank == 0:
            full = np.empty_like(img)
            for k, part in enumerate(parts):
                a0, b0, a, b = block_of(k, grid, H, W)
                full[a0:a0 + a, b0:b0 + b] = part.cpu().numpy()
            want = b200conv.reference_oracle(img, reps, boundary)
            good = bool((full == want).all())
            if opts.get("check_every"):
                good = good and iters < reps and iters % opts["check_every"] == 0
            print(f"case {H}x{W}x{ch} reps={reps} {boundary} {opts} grid={grid} iters={iters} ms={ms:.3f} "
                  f"{'OK' if good else 'MISMATCH'}", flush=True)
            ok = ok and good
        eng.close()
    # streaming API: different images back to back through one engine, copies overlapped, no host
    # barrier between images; an odd pass count (4+4+1) flips the buffer parity from image to image
    big = os.environ.get("B200CONV_DIST_BIG", "0") != "0"   # the bench shape's 8-GPU block, byte-checked
    pipe_cases = [("torus", (192, 384, 3, 9), {}), ("zero", (256, 256, 1, 40), {}),
                  ("torus", (512, 1024, 3, 100), dict(halo_depth=16))]
    if big:
        pipe_cases.append(("zero", (16384, 16384, 3, 100), dict(big=True)))   # 8 GPUs: 8192 x 4096 blocks
    for boundary, (H, W, ch, reps), extra in pipe_cases:
        try:
            grid = b200conv.decompose(world, H, W)
        except b200conv.B2CError:
            continue
        is_big = extra.pop("big", False)
        sh = b200conv.ShardedBlur(W, H, ch, boundary=boundary, tb=4, **extra)
        if not is_big:
            sh.engine.set_delay(3000 * (1 + (rank * 5) % 3))
        pipe = b200conv.ShardedPipeline(sh, reps)
        r0, c0, r, c = sh.block
        imgs = [np.random.default_rng(100 + j).integers(0, 256, (H, W) if ch == 1 else (H, W, ch), dtype=np.uint8)
                for j in range(2 if is_big else 5)]
        srcs = [torch.from_numpy(np.ascontiguousarray(im[r0:r0 + r, c0:c0 + c])).pin_memory() for im in imgs]
        dsts = [torch.empty_like(t).pin_memory() for t in srcs]
        for a, b in zip(srcs, dsts):
            pipe.submit(a, b)
        pipe.wait()
        good = True
        for im, d in zip(imgs, dsts):
            out = d.cuda()
            parts = [torch.empty_like(out) for _ in range(world)]
            dist.all_gather(parts, out)
            if rank == 0:
                full = np.empty_like(im)
                for k, part in enumerate(parts):
                    a0, b0, a, b = block_of(k, grid, H, W)
                    full[a0:a0 + a, b0:b0 + b] = part.cpu().numpy()
                if is_big:
                    # 805 MB x 100 iterations is out of reach of the scalar CPU oracle: the whole image is
                    # compared with the single-GPU engine (itself oracle-checked in test_gpu_large.py), and a
                    # full-width slab across the block boundary with the CPU oracle (rows further than
                    # `reps` from the slab's cut edges depend on nothing outside it)
                    one = b200conv.blur(torch.from_numpy(im).cuda(), reps, boundary=boundary).cpu().numpy()
                    good = good and bool((full == one).all())
                    mid, half = H // 2, 64
                    slab = im[mid - half - reps:mid + half + reps]
                    want = b200conv.reference_oracle(slab, reps, boundary)[reps:-reps]
                    good = good and bool((full[mid - half:mid + half] == want).all())
                    del one
                else:
                    good = good and bool((full == b200conv.reference_oracle(im, reps, boundary)).all())
        if rank == 0:
            print(f"case pipeline {H}x{W}x{ch} reps={reps} {boundary} grid={grid} {'OK' if good else 'MISMATCH'}",
                  flush=True)
            ok = ok and good
        sh.close()
    # the NCCL send/recv baseline must be exact as well (it is the comparison arm of the benches)
    from b200conv.parallel.sendrecv import NcclHaloBlur
    for boundary in ("zero", "torus"):
        H, W, ch, reps = 96, 192, 3, 7
        img = np.random.default_rng(9).integers(0, 256, (H, W, ch), dtype=np.uint8)
        nb = NcclHaloBlur(W, H, ch, boundary=boundary)
        nb.load(torch.from_numpy(np.ascontiguousarray(img[nb.r0:nb.r0 + nb.R, nb.c0:nb.c0 + nb.C])).cuda())
        nb.run(reps)
        out = nb.store()
        parts = [torch.empty_like(out) for _ in range(world)]
        dist.all_gather(parts, out)
        if rank == 0:
            full = np.empty_like(img)
            for k, part in enumerate(parts):
                a0, b0, a, b = block_of(k, nb.grid, H, W)
                full[a0:a0 + a, b0:b0 + b] = part.cpu().numpy()
            good = bool((full == b200conv.reference_oracle(img, reps, boundary)).all())
            print(f"case nccl-baseline {boundary} {'OK' if good else 'MISMATCH'}", flush=True)
            ok = ok and good
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.broadcast(flag, 0)
    dist.destroy_process_group()
    sys.exit(0 if int(flag.item()) else 1)


if __name__ == "__main__":
    main()
