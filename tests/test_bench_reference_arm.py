"""bench.py --impl reference: the arm around the unmodified reference binary must survive the
reference's own faults (its rgb kernel reads and writes out of bounds, cuda/cuda_convolution.cu:34;
on one box it died with "illegal memory access", reported at :90).  A stand-in executable plays the
binary here (B200CONV_REF_EXE): a failed process is retried, fresh, up to three times per run of the
arm, the failures are reported, and three failures give an `unavailable` line with exit status 0."""
import json
import os
import stat
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent


def fake_binary(tmp_path, fail_first):
    """A program with the reference's command line that fails its first `fail_first` runs."""
    counter = tmp_path / "runs"
    counter.write_text("0")
    exe = tmp_path / "cuda_convolution"
    exe.write_text(f"""#!{sys.executable}
import sys
p = {str(counter)!r}
n = int(open(p).read()) + 1
open(p, "w").write(str(n))
assert len(sys.argv) == 6 and sys.argv[5] in ("grey", "rgb")
if n <= {fail_first}:
    sys.stderr.write("CUDA error: an illegal memory access was encountered, line 90\\n")
    sys.exit(1)
print("Execution time: 0.250 sec")
""")
    exe.chmod(exe.stat().st_mode | stat.S_IEXEC)
    return exe, counter


def run_arm(tmp_path, exe, *extra):
    env = dict(os.environ, B200CONV_REF_EXE=str(exe), TMPDIR=str(tmp_path))
    p = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--impl", "reference", "--width", "64", "--height", "48",
                        "--reps", "3", "--steps", "2", "--warmup", "1", *extra],
                       capture_output=True, text=True, env=env, timeout=300, cwd=ROOT)
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [l for l in p.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1, p.stdout
    return json.loads(lines[0])


def test_two_faults_are_retried_and_reported(tmp_path):
    exe, counter = fake_binary(tmp_path, fail_first=2)
    d = run_arm(tmp_path, exe)
    assert d["impl"] == "reference" and "unavailable" not in d
    assert d["retries"] == 2 and len(d["failures"]) == 2 and "illegal memory access" in d["failures"][0]
    assert d["steps"] == 2 and d["warmup"] == 1
    assert int(counter.read_text()) == 5                      # 2 failed + 1 warm-up + 2 timed processes
    assert abs(d["ms_per_step"] - 250.0) < 1e-6
    assert abs(d["value"] - 64 * 48 * 3 / 0.25 / 1e9) < 1e-12 and d["e2e"]["value"] == d["value"]


def test_three_faults_give_an_unavailable_line(tmp_path):
    exe, counter = fake_binary(tmp_path, fail_first=99)
    d = run_arm(tmp_path, exe)
    assert d["impl"] == "reference" and "failed 3 time" in d["unavailable"] and len(d["failures"]) == 3
    assert int(counter.read_text()) == 3


def test_a_clean_binary_needs_no_retry(tmp_path):
    exe, _ = fake_binary(tmp_path, fail_first=0)
    d = run_arm(tmp_path, exe, "--color", "grey")
    assert d["retries"] == 0 and d["failures"] == [] and d["config"]["color"] == "grey"
