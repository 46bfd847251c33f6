#!/bin/bash
# usage: gpu_ncu.sh <tag> <prof_one.py args...>   — plain run first, then one ncu --set full capture
tag=$1; shift
mkdir -p gpurun_out
python scripts/prof_one.py "$@" > gpurun_out/plain_$tag.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:stream_ -s 1 -c 2 -f -o gpurun_out/prof_$tag \
    python scripts/prof_one.py "$@" > gpurun_out/ncu_$tag.log 2>&1
echo "rc=$?"; tail -3 gpurun_out/ncu_$tag.log
