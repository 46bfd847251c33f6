#!/bin/bash
# CPU baselines on the GPU box: the unmodified reference MPI / OpenMP sources (threads-as-ranks
# MPI shim), BASELINE.json config 1 (1920x2520 rgb, 40 reps, 4 ranks) and a rank sweep.
mkdir -p gpurun_out
python baseline/install_ref.py > gpurun_out/ref_build.log 2>&1
B=$GRAFT_REPO_ROOT/baseline/_ref/build
cd /tmp && python - <<'PY'
import numpy as np
np.random.default_rng(0).integers(0,256,size=(2520,1920*3),dtype=np.uint8).tofile("w_rgb.raw")
np.random.default_rng(0).integers(0,256,size=(2520,1920),dtype=np.uint8).tofile("w_grey.raw")
PY
echo "cores: $(nproc)"
for n in 1 4 16; do
  for exe in mpi_convolution omp_convolution omp_convolution_PATCHED; do
    t=$($B/$exe -n $n w_rgb.raw 1920 2520 40 rgb 2>&1 | tail -1)
    echo "$exe n=$n rgb 1920x2520 40 reps: $t s"
  done
done
t=$($B/mpi_convolution -n 16 w_grey.raw 1920 2520 20 grey | tail -1); echo "mpi_convolution n=16 grey 20 reps: $t s"
OMP_NUM_THREADS=8 $B/omp_convolution_PATCHED -n 4 w_rgb.raw 1920 2520 40 rgb | sed 's/^/omp_PATCHED n=4 x 8 threads rgb 40 reps: /'
