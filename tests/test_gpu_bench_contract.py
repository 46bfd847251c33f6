"""bench.py prints ONE JSON line with the keys the driver reads (tiny config to stay fast)."""
import json
import subprocess
import sys
from pathlib import Path

import pytest

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent


def test_bench_json_contract():
    p = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--width", "2048", "--height", "2048", "--reps", "16",
                        "--steps", "2", "--warmup", "3"], capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [l for l in p.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
              "vs_baseline", "dtype", "data", "config", "clocks", "e2e", "gpu_launches"):
        assert k in d, k
    assert d["n_gpus"] == 1 and d["steps"] == 2 and d["warmup"] == 3 and d["higher_is_better"] is True
    assert d["value"] > 0 and 2 <= d["gpu_launches"] <= 2 * 4     # per step: one persistent kernel, or 4 passes of K=4
    assert d["e2e"]["value"] > 0 and d["e2e"]["h2d_bytes_per_step"] == 2048 * 2048 * 3
    # (for images this small the pipelined end-to-end rate may exceed the single-stream device
    #  rate: two engines on two streams fill a GPU that one latency-bound image cannot)
    assert set(d["clocks"]) >= {"sm_mhz", "sm_max_mhz", "reasons"}
