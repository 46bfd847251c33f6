#!/bin/bash
# One multi-GPU measurement round: correctness worker, then the strong-scaling bench with the
# protocol counters and the end-to-end timeline.  Usage: scripts/gpu_scale_probe.sh N [steps]
N=${1:-8}; STEPS=${2:-10}
mkdir -p gpurun_out
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node=$N --master-addr 127.0.0.1 --master-port $1 "${@:2}"; }
B200CONV_DIST_BIG=${BIG:-0} run 29521 tests/dist_worker.py > gpurun_out/dist$N.log 2>&1; echo "dist rc=$?"
grep -c " OK" gpurun_out/dist$N.log; grep -v " OK" gpurun_out/dist$N.log | grep -v "^\*\|OMP_NUM\|^$\|NCCL version" | tail -20
run 29522 bench.py --gpus $N --steps $STEPS --warmup 3 --halo-prof --trace-e2e > gpurun_out/b$N.json 2> gpurun_out/b$N.err; echo "bench rc=$?"
grep -v "^\*\|OMP_NUM\|^$" gpurun_out/b$N.err | tail -5
python - <<PY
import json
d=json.loads(open("gpurun_out/b$N.json").read().strip().splitlines()[-1])
print("value", round(d["value"],1), "ms", round(d["ms_per_step"],3), "launches", d["gpu_launches"])
print(d["config"]["plan"]); print(d["config"].get("coupled_tuning"))
print("e2e", round(d["e2e"]["value"],1), round(d["e2e"]["ms_per_step"],3))
print(d.get("per_rank")); print("start wait", d.get("start_barrier_wait_cycles"))
for i,r in enumerate(d.get("halo_profile_per_rank") or []):
    w=max(1,r["waits"]); print(i, "waits", r["waits"], "avg data wait", {k:v for k,v in r["data_wait_cycles"].items() if v}, "max", {k:v for k,v in r["data_wait_max"].items() if v}, "ack", {k:v for k,v in r["ack_wait_cycles"].items() if v}, "publish avg", r["publish_cycles"]//max(1,r["publishes"]))
for r in (d["e2e"].get("timeline_ms_per_rank") or []): print(r[-2:])
PY
