"""Three runs of the resident kernel on 1920x2520 grey (1000 repetitions each) for ncu:
    ncu --set full --clock-control none --import-source on -k regex:resident_kernel --launch-skip 1 -c 1 \
        -o gpurun_out/resident_k8 python scripts/prof_resident.py [tb]
"""
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from b200conv import Engine  # noqa: E402

tb = int(sys.argv[1]) if len(sys.argv) > 1 else 8
e = Engine(1920, 2520, 1, kernel="resident", tb=tb)
x = torch.from_numpy(np.random.default_rng(0).integers(0, 256, (2520, 1920), dtype=np.uint8)).cuda()
e.load(x)
for _ in range(3):
    e.run(1000)
torch.cuda.synchronize()
print(e.describe(), "iterations_done", e.iterations_done())
