"""GPU check of the resident kernel (kernel="resident"): bytes against the CPU oracle on small
shapes, against the streaming kernel on the 1920x2520 configs at 1000 repetitions, then timing of
both kernels with CUDA events.  Writes gpurun_out/resident_validate.json.

    timeout 60 python scripts/validate_resident.py
"""
import json
import os
import sys
import time
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from b200conv import Engine, blur, reference_oracle  # noqa: E402


def rnd(h, w, ch, seed):
    return np.random.default_rng(seed).integers(0, 256, (h, w) if ch == 1 else (h, w, ch), dtype=np.uint8)


def main():
    res = {"cases": [], "timing": {}}
    bad = 0
    small = [(40, 64, 1, 9, "zero"), (40, 64, 1, 9, "torus"), (130, 488, 1, 13, "torus"), (130, 488, 1, 13, "zero"),
             (200, 96, 3, 11, "torus"), (77, 160, 3, 5, "zero"), (61, 8, 1, 7, "torus"), (4, 16, 1, 3, "zero"),
             (300, 520, 1, 1, "zero"), (300, 520, 1, 2, "torus"), (300, 520, 1, 3, "zero"), (300, 520, 1, 4, "torus")]
    for (h, w, ch, reps, boundary) in small:
        img = rnd(h, w, ch, 1)
        got = blur(torch.from_numpy(img).cuda(), reps, boundary=boundary, kernel="resident").cpu().numpy()
        ok = bool((got == reference_oracle(img, reps, boundary)).all())
        bad += not ok
        res["cases"].append([h, w, ch, reps, boundary, "oracle", ok])
        print(h, w, ch, reps, boundary, "OK" if ok else "MISMATCH", flush=True)
    for (h, w, ch, reps, boundary, tb) in [(130, 488, 1, 21, "torus", 8), (200, 96, 3, 19, "zero", 8), (90, 304, 1, 8, "zero", 8),
                                           (130, 488, 1, 37, "torus", 16), (200, 96, 3, 35, "zero", 16), (90, 304, 1, 16, "zero", 16),
                                           (40, 64, 1, 9, "torus", 0)]:
        img = rnd(h, w, ch, 1)
        got = blur(torch.from_numpy(img).cuda(), reps, boundary=boundary, kernel="resident", tb=tb).cpu().numpy()
        ok = bool((got == reference_oracle(img, reps, boundary)).all())
        bad += not ok
        res["cases"].append([h, w, ch, reps, boundary, f"oracle, tb={tb}", ok])
        print(h, w, ch, reps, boundary, f"tb={tb}", "OK" if ok else "MISMATCH", flush=True)
    big = [(2520, 1920, 1, 1000, "zero", 4), (2520, 1920, 1, 999, "torus", 8), (2520, 1920, 1, 1000, "zero", 16),
           (1000, 1800, 3, 250, "zero", 8), (1000, 1800, 3, 37, "torus", 4)]
    for (h, w, ch, reps, boundary, tb) in big:
        img = torch.from_numpy(rnd(h, w, ch, 2)).cuda()
        got = blur(img, reps, boundary=boundary, kernel="resident", tb=tb)
        want = blur(img, reps, boundary=boundary, kernel="stream")
        ok = bool(torch.equal(got, want))
        bad += not ok
        res["cases"].append([h, w, ch, reps, boundary, f"stream kernel, tb={tb}", ok])
        print(h, w, ch, reps, boundary, tb, "OK" if ok else "MISMATCH", flush=True)
    # same engine run twice in a row (epochs continue across launches)
    eng = Engine(1920, 2520, 1, kernel="resident")
    img = torch.from_numpy(rnd(2520, 1920, 1, 3)).cuda()
    out = torch.empty_like(img)
    eng.load(img); eng.run(10); eng.run(7); eng.store(out)
    torch.cuda.synchronize()
    ok = bool(torch.equal(out, blur(img, 17, kernel="stream")))
    bad += not ok
    print("two runs on one engine", "OK" if ok else "MISMATCH", "|", eng.describe(), flush=True)
    res["plan"] = eng.describe()
    for kern in ("resident", "stream"):
        for (h, w, ch) in ((2520, 1920, 1), (1000, 1800, 3)):
            e = eng if (kern == "resident" and ch == 1) else Engine(w, h, ch, kernel=kern)
            x = torch.from_numpy(rnd(h, w, ch, 4)).cuda()
            e.load(x)
            for _ in range(3):
                e.run(1000)
            torch.cuda.synchronize()
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            ev0.record()
            for _ in range(5):
                e.run(1000)
            ev1.record()
            torch.cuda.synchronize()
            ms = ev0.elapsed_time(ev1) / 5
            e.iterations_done()
            res["timing"][f"{kern} {w}x{h}x{ch} 1000 reps"] = {"ms": ms, "gpix_it_per_s": w * h * 1000 / ms / 1e6}
            print(kern, w, h, ch, f"{ms:.3f} ms  {w * h * 1000 / ms / 1e6:.0f} Gpix-it/s", flush=True)
    # where the time goes: exchange interval 4 / 8, and the profiling switches of the kernel
    # (1 = no exchange at all); bytes are not checked here
    x = torch.from_numpy(rnd(2520, 1920, 1, 4)).cuda()
    for tb in (4, 8, 16):
        e = Engine(1920, 2520, 1, kernel="resident", tb=tb)
        for dbg in (0, 1):
            os.environ["B200CONV_RES_DEBUG"] = str(dbg)
            e.load(x)
            for _ in range(2):
                e.run(1000)
            torch.cuda.synchronize()
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            ev0.record()
            for _ in range(5):
                e.run(1000)
            ev1.record()
            torch.cuda.synchronize()
            ms = ev0.elapsed_time(ev1) / 5
            res["timing"][f"resident grey 1920x2520 tb={tb} debug={dbg}"] = ms
            print(f"tb={tb} debug={dbg}: {ms:.3f} ms | {e.describe()}", flush=True)
        os.environ["B200CONV_RES_DEBUG"] = "0"
        e.iterations_done()
    res["bad"] = bad
    os.makedirs(ROOT / "gpurun_out", exist_ok=True)
    (ROOT / "gpurun_out" / "resident_validate.json").write_text(json.dumps(res, indent=1))
    print("bad =", bad)
    return bad


if __name__ == "__main__":
    t0 = time.time()
    rc = main()
    print(f"{time.time() - t0:.1f} s")
    sys.exit(1 if rc else 0)
