#!/bin/bash
# bench.py at N GPUs with the three flag-fence flavours.  Usage: scripts/gpu_fence_modes.sh N [steps]
N=${1:-8}; STEPS=${2:-10}
for fm in 1 2 0; do
B200CONV_FLAG_FENCE=$fm python -m torch.distributed.run --nnodes=1 --nproc-per-node=$N --master-addr 127.0.0.1 --master-port 2955$fm bench.py --gpus $N --steps $STEPS --warmup 3 --no-solo --trace-e2e > gpurun_out/bf${N}_$fm.json 2> gpurun_out/bf.err
python - <<PY
import json
d=json.loads(open("gpurun_out/bf${N}_$fm.json").read().strip().splitlines()[-1])
print("flag_fence=$fm value", round(d["value"],1), "ms", round(d["ms_per_step"],3), "e2e", round(d["e2e"]["value"],1), round(d["e2e"]["ms_per_step"],3))
tl=d["e2e"].get("timeline_ms_per_rank") or []
for r in tl[:8:4]: print([round(x["run_end"]-x["run_start"],2) for x in r[-4:]], [round(x["run_start"]-x["compute_free"],2) for x in r[-4:]])
PY
done
