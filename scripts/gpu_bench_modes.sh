#!/bin/bash
# bench.py at N GPUs with the persistent kernel forced on and off.  Usage: scripts/gpu_bench_modes.sh N [steps]
N=${1:-8}; STEPS=${2:-10}
for pm in 1 0; do
B200CONV_PERSIST=$pm python -m torch.distributed.run --nnodes=1 --nproc-per-node=$N --master-addr 127.0.0.1 --master-port 2954$pm bench.py --gpus $N --steps $STEPS --warmup 3 --no-solo --trace-e2e > gpurun_out/bm${N}_p$pm.json 2> gpurun_out/bm.err
python - <<PY
import json
d=json.loads(open("gpurun_out/bm${N}_p$pm.json").read().strip().splitlines()[-1])
print("persist=$pm value", round(d["value"],1), "ms", round(d["ms_per_step"],3), "launches", d["gpu_launches"], "e2e", round(d["e2e"]["value"],1), round(d["e2e"]["ms_per_step"],3))
print(d["config"]["plan"]); print(d["config"].get("coupled_tuning"))
tl=d["e2e"].get("timeline_ms_per_rank") or []
for r in tl[:2]: print([round(x["run_end"]-x["run_start"],2) for x in r[-4:]], [round(x["h2d_end"]-x["h2d_start"],2) for x in r[-4:]])
PY
done
