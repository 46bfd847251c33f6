"""Dry-run planner (no GPU needed): what a run of ``b200conv image W H reps grey|rgb --gpus N``
would look like — process grid (RowsDivision, mpi/mpi_convolution.c:350-364), block per rank,
neighbour counts, halo bytes per iteration, k=1 HBM-roofline time, whether a block fits the
register-resident kernel, and the streaming plan of a rank: halo depth, passes and exchanges per
run, which kernel AUTO would take and how many launches a run costs.

    python -m b200conv.utils.plan 16384 16384 rgb --gpus 8 --reps 1000 [--boundary torus]
"""
from __future__ import annotations

import argparse
import ctypes as C
import json

from .._lib import B2CError, lib
from ..parallel.decomposition import block_of, decompose, neighbours

HBM_GBS = 6555.5          # measured copy bandwidth of one B200 (MEASURED_PEAKS.json)
NVLINK_GBS = 900.0        # per direction per GPU


def plan(width: int, height: int, channels: int, gpus: int = 1, reps: int = 1000, boundary: str = "zero",
         sms: int = 148) -> dict:
    try:
        grid = decompose(gpus, height, width)
    except B2CError:
        return {"error": "Cannot divide to processes", "width": width, "height": height, "gpus": gpus}
    r0, c0, rows, cols = block_of(0, grid, height, width)
    bw = cols * channels
    nb_counts = []
    for rank in range(gpus):
        peers = {p for p in neighbours(rank, grid, boundary).values() if p >= 0 and p != rank}
        nb_counts.append(len(peers))
    multi = gpus > 1
    halo = (2 * bw + 2 * rows * channels + 4 * channels) if multi or boundary == "torus" else 0
    resident = None
    for k in (16, 8, 4):
        out = (C.c_int * 5)()
        if lib.b2c_resident_plan(bw, rows, channels, k, sms, out):
            resident = {"exchange_every": k, "rows_per_warp": out[0], "ctas": [out[2], out[1]],
                        "valid_rows": out[3], "valid_bytes": out[4]}
            break
    # the streaming plan of one rank (host-side planner hooks of the C core; model plan, no timing)
    has_nb = multi or boundary == "torus"
    streaming = None
    buf = (C.c_int * (16 * 4096))()
    n = lib.b2c_debug_pass_plans(rows, cols, channels, int(has_nb), 0, 0, 0, 0, int(reps), sms, buf, 4096)
    if n > 0:
        passes = [list(buf[16 * i:16 * i + 16]) for i in range(min(n, 4096))]
        groups = sum(p[3] for p in passes) if has_nb else 0
        bp = (C.c_int * 10)()
        lib.b2c_debug_band_plan(rows, cols, channels, int(has_nb), 0, 0, 0, 0, sms, bp)
        block_bytes = rows * bw
        fits = resident is not None and not multi
        if fits:
            kernel = "resident (whole image in registers, one cooperative launch per run)"
        elif block_bytes >= 300e6 and (not has_nb or bw % bp[8] == 0):
            kernel = "stream-persistent (one launch per run, per-tile dependencies)"
        else:
            kernel = "stream (one CUDA graph of PDL-chained passes per run)"
        streaming = {
            "fused_iterations_per_pass": bp[7], "strip_bytes": bp[8], "halo_depth_iterations": bp[9] if has_nb else 0,
            "bands": bp[2], "chunks": bp[3], "rows_per_warp": bp[4],
            "passes_per_run": n, "halo_exchanges_per_run": groups,
            "extra_rows_recomputed_per_side_max": max(p[1] for p in passes),
            "auto_kernel": kernel,
            "host_launches_per_run": 1 if fits else (2 if has_nb else 1),
        }
    hbm_bytes = 2.0 * rows * bw                      # one read + one write per iteration without blocking
    t_hbm_ms = hbm_bytes * reps / (HBM_GBS * 1e9) * 1e3
    return {
        "image": [height, width, channels], "gpus": gpus, "grid": list(grid), "block": [rows, cols],
        "block_bytes": rows * bw, "neighbour_gpus_per_rank": nb_counts,
        "halo_bytes_per_gpu_per_iteration": halo,
        "halo_us_per_iteration_at_nvlink": halo / (NVLINK_GBS * 1e9) * 1e6,
        "hbm_bytes_per_gpu_per_iteration_k1": hbm_bytes,
        "k1_hbm_roofline_ms": t_hbm_ms,
        "k1_hbm_roofline_gpix_it_per_s": width * height * reps / (t_hbm_ms * 1e-3) / 1e9 if t_hbm_ms else None,
        "fits_resident_kernel": (resident if not multi else None),
        "streaming_plan": streaming,
        "index_width": "64-bit" if width * height * channels >= 2 ** 31 else "fits 32-bit (64-bit used anyway)",
    }


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__.splitlines()[0])
    ap.add_argument("width", type=int)
    ap.add_argument("height", type=int)
    ap.add_argument("color", choices=["grey", "rgb"])
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--reps", type=int, default=1000)
    ap.add_argument("--boundary", default="zero", choices=["zero", "torus"])
    a = ap.parse_args(argv)
    print(json.dumps(plan(a.width, a.height, 3 if a.color == "rgb" else 1, a.gpus, a.reps, a.boundary), indent=1))
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
