#!/bin/bash
# usage: gpu_multi.sh N  — multi-GPU correctness + scaling runs on N GPUs of one box
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node=$N --master-addr 127.0.0.1"
F='^{|^case'
timeout 600 $TR --master-port 29601 tests/dist_worker.py 2>&1 | grep -E "$F|Error|error" | head -20
echo "== strong 16384^2 rgb"
timeout 600 $TR --master-port 29602 bench.py --gpus $N --steps 2 --warmup 3 2>&1 | grep -E "$F" | tee gpurun_out/bench_strong_$N.json | cut -c1-330
echo "== 32768^2 rgb torus"
timeout 600 $TR --master-port 29603 bench.py --gpus $N --steps 2 --warmup 3 --width 32768 --height 32768 --boundary torus 2>&1 | grep -E "$F" | tee gpurun_out/bench_32k_torus_$N.json | cut -c1-330
echo "== weak 16384^2 grey per GPU, check every 20"
timeout 600 $TR --master-port 29604 bench.py --gpus $N --steps 2 --warmup 3 --weak --color grey --check-every 20 2>&1 | grep -E "$F" | tee gpurun_out/bench_weak_$N.json | cut -c1-330
echo "== nccl baseline"
timeout 600 $TR --master-port 29605 scripts/nccl_baseline_bench.py --check 2>&1 | grep -E "$F" | tee gpurun_out/nccl_baseline_$N.json
timeout 300 python -m pytest tests/test_launcher.py -m gpu -q -k multi 2>&1 | tail -2
echo "== launcher, plumbing config (1920x2520 rgb, 40 reps) on $N GPUs"
python - <<'PY'
import sys
sys.path.insert(0, ".")
from b200conv.utils import waterfall_image, write_raw
write_raw("/tmp/waterfall_1920_2520.raw", waterfall_image(2520, 1920, 3, seed=0))
PY
( cd /tmp && $GRAFT_REPO_ROOT/parallel-image-convolution-using-mpi-openmp-and-cuda_b200/lib/b200conv waterfall_1920_2520.raw 1920 2520 40 rgb --gpus $( [ $N -ge 4 ] && echo 4 || echo $N ) --profile --json $GRAFT_REPO_ROOT/gpurun_out/launcher_plumbing_$N.json 2>&1 | tail -8 )
