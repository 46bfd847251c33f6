#!/bin/bash
# Reference arm probe: build the unmodified reference on the box, time a few shapes.
mkdir -p gpurun_out
python baseline/install_ref.py > gpurun_out/ref_build.log 2>&1; tail -5 gpurun_out/ref_build.log
cd /tmp && python - <<'PY'
import numpy as np
for name,w,h,ch in (("g.raw",1920,2520,1),("c.raw",1920,2520,3),("big.raw",16384,16384,3)):
    np.random.default_rng(0).integers(0,256,size=(h,w*ch),dtype=np.uint8).tofile(name)
PY
REF=$GRAFT_REPO_ROOT/baseline/_ref/cuda/cuda_convolution
for args in "g.raw 1920 2520 1000 grey" "c.raw 1920 2520 100 rgb" "big.raw 16384 16384 10 rgb" "big.raw 16384 16384 100 rgb"; do
  echo "== $args"; ( time timeout 600 $REF $args ) 2>&1 | grep -v "^$\|user\|sys"
done
cd $GRAFT_REPO_ROOT
