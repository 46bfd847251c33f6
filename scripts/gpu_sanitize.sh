#!/bin/bash
# usage: gpu_sanitize.sh memcheck|racecheck|initcheck|synccheck   (one tool per gpurun call)
tool=${1:-memcheck}
mkdir -p gpurun_out
cat > /tmp/san_case.py <<'PY'
import sys, numpy as np, torch
sys.path.insert(0, ".")
import b200conv
rng = np.random.default_rng(0)
for ch, boundary, kw in ((1, "zero", dict(tb=4)), (3, "torus", dict(tb=4)), (3, "zero", dict(tb=2, kernel="tma")),
                         (3, "zero", dict(kernel="naive")), (1, "cuda-compat", {})):
    img = rng.integers(0, 256, (70, 150) if ch == 1 else (70, 150, 3), dtype=np.uint8)
    out = b200conv.blur(torch.from_numpy(img).cuda(), 6, boundary=boundary, **kw)
    torch.cuda.synchronize()
    assert (out.cpu().numpy() == b200conv.reference_oracle(img, 6, boundary)).all()
g = b200conv.LocalGrid(256, 64, 3, (2, 2), tb=4)
img = rng.integers(0, 256, (64, 256, 3), dtype=np.uint8)
g.load(torch.from_numpy(img).cuda()); g.run(8)
assert (g.gather().numpy() == b200conv.reference_oracle(img, 8, "zero")).all()
g.close()
print("sanitize case ok")
PY
timeout 900 compute-sanitizer --tool $tool --error-exitcode 3 python /tmp/san_case.py > gpurun_out/sanitizer_$tool.log 2>&1
echo "rc=$?"; grep -E "ERROR SUMMARY|sanitize case ok|Error|error" gpurun_out/sanitizer_$tool.log | head -10
