#!/usr/bin/env python3
"""Where does the end-to-end pipeline lose time against the device-timed loop?
torchrun --nproc-per-node N scripts/e2e_probe.py [--width W --height H --reps R --images K]

Runs ShardedPipeline with the PCIe copies switched on and off (host<->device, device<->device
stand-ins) and prints, per variant, the steady-state milliseconds per image and the duration of
the iteration loop inside the pipeline, max over ranks."""
import argparse
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200conv  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--width", type=int, default=16384)
    ap.add_argument("--height", type=int, default=16384)
    ap.add_argument("--reps", type=int, default=1000)
    ap.add_argument("--images", type=int, default=8)
    ap.add_argument("--color", default="rgb")
    a = ap.parse_args()
    local = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
    rank, world = dist.get_rank(), dist.get_world_size()
    ch = 3 if a.color == "rgb" else 1
    sh = b200conv.ShardedBlur(a.width, a.height, ch, device=local)
    shape = sh.block_shape
    host_in = torch.randint(0, 256, shape, dtype=torch.uint8).pin_memory()
    host_out = [torch.empty(shape, dtype=torch.uint8).pin_memory() for _ in range(2)]
    dev_in = host_in.cuda()
    dev_out = [torch.empty_like(dev_in) for _ in range(2)]

    def mx(x):
        t = torch.tensor([x], dtype=torch.float64, device=f"cuda:{local}")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # device-timed loop on resident data, for reference
    sh.load(dev_in)
    sh.engine.set_profile(True)
    sh.timed_run(a.reps)
    loop_ms = sh.timed_run(a.reps)
    hp = sh.engine.halo_profile(reset=True)
    if rank == 0:
        print("device-timed loop: avg publish cycles", hp["publish_cycles"] // max(1, hp["publishes"]), "avg wait",
              (sum(hp["data_wait_cycles"].values()) + sum(hp["ack_wait_cycles"].values())) // max(1, hp["waits"]))
    rows = []
    for name, src, dsts in (("h2d+d2h", host_in, host_out), ("no copies", dev_in, dev_out),
                            ("h2d only", host_in, dev_out), ("d2h only", dev_in, host_out)):
        pipe = b200conv.ShardedPipeline(sh, a.reps, depth=2, trace=True)
        for i in range(a.images):
            pipe.submit(src, dsts[i % 2])
        tl = pipe.timeline()
        torch.cuda.synchronize()
        dist.barrier()
        steady = (tl[-1]["run_start"] - tl[2]["run_start"]) / (len(tl) - 3)
        run = sum(t["run_end"] - t["run_start"] for t in tl[2:]) / (len(tl) - 2)
        h2d = sum(t["h2d_end"] - t["h2d_start"] for t in tl[2:]) / (len(tl) - 2)
        prep = sum(t["run_start"] - t["compute_free"] for t in tl[2:]) / (len(tl) - 2)
        hp = sh.engine.halo_profile(reset=True)
        dw = sum(hp["data_wait_cycles"].values()) + sum(hp["ack_wait_cycles"].values())
        rows.append((name, mx(steady), mx(run), mx(h2d), mx(prep), mx(dw / max(1, hp["waits"])),
                     mx(hp["publish_cycles"] / max(1, hp["publishes"])), mx(max(hp["data_wait_max"].values()))))
    if rank == 0:
        print(f"{world} GPUs, {a.width}x{a.height} {a.color}, {a.reps} reps; device-timed loop {loop_ms:.3f} ms; plan {sh.engine.describe()}")
        print(f"{'variant':12s} {'ms/image':>9s} {'loop':>8s} {'input copy':>11s} {'load+exch':>10s} {'avg wait cyc':>13s} "
              f"{'avg publish cyc':>16s} {'max data wait':>14s}")
        for r in rows:
            print(f"{r[0]:12s} {r[1]:9.3f} {r[2]:8.3f} {r[3]:11.3f} {r[4]:10.3f} {r[5]:13.0f} {r[6]:16.0f} {r[7]:14.0f}", flush=True)
    sh.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
