"""utils/plan.py: the dry-run planner reproduces the hand-computed numbers of SURVEY §2.6 / BASELINE §5."""
import json
import subprocess
import sys
from pathlib import Path

from b200conv.utils.plan import plan

ROOT = Path(__file__).resolve().parent.parent


def test_32768_rgb_on_8_gpus_matches_the_survey_numbers():
    p = plan(32768, 32768, 3, gpus=8, reps=1000, boundary="torus")
    assert p["grid"] == [2, 4] and p["block"] == [16384, 8192]
    assert p["halo_bytes_per_gpu_per_iteration"] == 147468
    assert abs(p["k1_hbm_roofline_ms"] - 122.8) < 0.2                 # BASELINE §5
    assert p["neighbour_gpus_per_rank"] == [5] * 8                    # 2-high torus: N and S are the same GPU
    assert plan(32768, 32768, 3, gpus=8)["neighbour_gpus_per_rank"] == [3, 5, 5, 3, 3, 5, 5, 3]
    assert p["fits_resident_kernel"] is None


def test_small_grey_image_fits_the_resident_kernel_and_rgb_does_not():
    g = plan(1920, 2520, 1)
    assert g["fits_resident_kernel"] == {"exchange_every": 16, "rows_per_warp": 12, "ctas": [16, 9],
                                         "valid_rows": 160, "valid_bytes": 224}
    assert abs(g["k1_hbm_roofline_ms"] - 1.476) < 0.01
    assert plan(1920, 2520, 3)["fits_resident_kernel"] is None
    assert plan(10, 10, 1, gpus=7) == {"error": "Cannot divide to processes", "width": 10, "height": 10, "gpus": 7}


def test_command_line():
    r = subprocess.run([sys.executable, "-m", "b200conv.utils.plan", "16384", "16384", "rgb", "--gpus", "4"],
                       capture_output=True, text=True, cwd=ROOT, timeout=120)
    assert r.returncode == 0, r.stderr[-1500:]
    d = json.loads(r.stdout)
    assert d["grid"] == [2, 2] and d["block"] == [8192, 8192] and d["neighbour_gpus_per_rank"] == [3, 3, 3, 3]


def test_streaming_plan_of_the_bench_shape_on_8_gpus():
    s = plan(16384, 16384, 3, gpus=8, reps=1000)["streaming_plan"]
    assert s["fused_iterations_per_pass"] == 4 and s["strip_bytes"] == 16 and s["bands"] == 26
    assert s["halo_depth_iterations"] == 32 and s["passes_per_run"] == 250
    assert s["halo_exchanges_per_run"] == 32                 # 31 groups of 32 iterations + one of 8
    assert s["extra_rows_recomputed_per_side_max"] == 28     # first pass of a group: F - K rows into the frame
    assert s["auto_kernel"].startswith("stream (") and s["host_launches_per_run"] == 2
    one = plan(16384, 16384, 3, gpus=1, reps=1000)["streaming_plan"]
    assert one["halo_exchanges_per_run"] == 0 and one["auto_kernel"].startswith("stream-persistent")
    assert plan(1920, 2520, 1)["streaming_plan"]["auto_kernel"].startswith("resident")
