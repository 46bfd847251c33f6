#!/bin/bash
# ncu captures of every kernel family (one GPU).  Each case runs plainly first.
mkdir -p gpurun_out
cap() {  # case kernel-regex skip
  python scripts/ncu_case.py $1 > gpurun_out/plain_$1.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:$2 -s $3 -c 1 -f -o gpurun_out/ncu_$1 python scripts/ncu_case.py $1 > gpurun_out/ncu_$1.log 2>&1
  echo "$1 rc=$?"; tail -2 gpurun_out/ncu_$1.log
}
cap persist stream_persist_kernel 1
cap perpass 'stream_kernel' 5
cap tma stream_tma_kernel 5
cap general 'stream_kernel' 5
cap resident resident_kernel 1
python scripts/ncu_case.py launches > gpurun_out/plain_launches.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file gpurun_out/launches.csv python scripts/ncu_case.py launches > gpurun_out/ncu_launches.log 2>&1
echo "launches rc=$?"
ls -la gpurun_out/*.ncu-rep
