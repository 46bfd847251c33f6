#!/bin/bash
# GPU check: smoke, GPU tests, variant sweep, short bench.  Everything is wrapped in timeouts.
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/gpus.txt 2>&1
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
tail -2 gpurun_out/smoke.log
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -25 gpurun_out/pytest_gpu.log
timeout 600 python scripts/tune.py > gpurun_out/tune.jsonl 2>&1; echo "tune rc=$?"
cut -c1-150 gpurun_out/tune.jsonl
timeout 600 python bench.py --steps 2 --warmup 3 > gpurun_out/bench1.log 2>&1; echo "bench rc=$?"
tail -1 gpurun_out/bench1.log | cut -c1-400
