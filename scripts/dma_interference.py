#!/usr/bin/env python3
"""Does a PCIe copy running beside the iteration loop slow the loop down?  One GPU, no peers.
Times `reps` iterations alone, then with a continuous H2D, D2H and both on other streams;
samples SM clock / power / throttle reasons meanwhile."""
import argparse
import os
import subprocess
import sys
import threading
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200conv  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--width", type=int, default=8192)
ap.add_argument("--height", type=int, default=4096)
ap.add_argument("--reps", type=int, default=1000)
ap.add_argument("--loops", type=int, default=12)
a = ap.parse_args()
img = torch.randint(0, 256, (a.height, a.width, 3), dtype=torch.uint8, device="cuda")
eng = b200conv.Engine(a.width, a.height, 3)
eng.load(img)
eng.exchange()
host = torch.empty(256 << 20, dtype=torch.uint8).pin_memory()
dev = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
s_run, s_h2d, s_d2h = torch.cuda.Stream(), torch.cuda.Stream(), torch.cuda.Stream()


def smi():
    out = subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,power.draw,clocks_throttle_reasons.active", "--format=csv,noheader",
                          "-i", "0"], capture_output=True, text=True).stdout.strip()
    return out


def trial(h2d, d2h):
    stop = threading.Event()
    samples = []

    def sampler():
        while not stop.is_set():
            samples.append(smi())
            time.sleep(0.05)

    th = threading.Thread(target=sampler)
    with torch.cuda.stream(s_run):
        eng.run(a.reps)
    torch.cuda.synchronize()
    th.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n_copies = 0
    with torch.cuda.stream(s_run):
        e0.record()
        for _ in range(a.loops):
            eng.run(a.reps)
        e1.record()
    # keep the copy engines busy until the loop is done
    while not e1.query():
        if h2d:
            with torch.cuda.stream(s_h2d):
                dev.copy_(host, non_blocking=True)
        if d2h:
            with torch.cuda.stream(s_d2h):
                host.copy_(dev, non_blocking=True)
        n_copies += 1
        s_h2d.synchronize()
        s_d2h.synchronize()
        if not (h2d or d2h):
            time.sleep(0.001)
    torch.cuda.synchronize()
    stop.set()
    th.join()
    ms = e0.elapsed_time(e1) / a.loops
    return ms, n_copies, samples[len(samples) // 2] if samples else ""


print(eng.describe())
for name, h, d in (("alone", 0, 0), ("with H2D", 1, 0), ("with D2H", 0, 1), ("with both", 1, 1), ("alone", 0, 0)):
    ms, n, smp = trial(h, d)
    print(f"{name:10s} {ms:8.3f} ms per {a.reps} iterations   copies of 256 MiB issued: {n:3d}   smi: {smp}", flush=True)
eng.close()
