#!/bin/bash
# usage: gpu_scale_only.sh N — the three scaling benches on N GPUs (no correctness suite)
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node=$N --master-addr 127.0.0.1"
echo "== strong 16384^2 rgb"
timeout 600 $TR --master-port 29602 bench.py --gpus $N --steps 3 --warmup 3 2>&1 | grep -E "^\{" | tee gpurun_out/bench_strong_$N.json | cut -c1-200
echo "== 32768^2 rgb torus"
timeout 600 $TR --master-port 29603 bench.py --gpus $N --steps 2 --warmup 3 --width 32768 --height 32768 --boundary torus 2>&1 | grep -E "^\{" | tee gpurun_out/bench_32k_torus_$N.json | cut -c1-200
echo "== weak 16384^2 grey per GPU, check every 20"
timeout 600 $TR --master-port 29604 bench.py --gpus $N --steps 2 --warmup 3 --weak --color grey --check-every 20 2>&1 | grep -E "^\{" | tee gpurun_out/bench_weak_$N.json | cut -c1-200
echo "== weak 16384^2 rgb per GPU"
timeout 600 $TR --master-port 29605 bench.py --gpus $N --steps 2 --warmup 3 --weak 2>&1 | grep -E "^\{" | tee gpurun_out/bench_weak_rgb_$N.json | cut -c1-200
