#!/usr/bin/env python3
"""Key numbers of .ncu-rep files as one table: scripts/ncu_summary.py gpurun_out/ncu_*.ncu-rep"""
import csv
import io
import subprocess
import sys

KEYS = [
    ("gpu__time_duration.sum", "time"),
    ("launch__registers_per_thread", "regs"),
    ("launch__grid_size", "grid"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "occupancy %"),
    ("smsp__inst_executed.sum", "warp instr"),
    ("sm__inst_executed.avg.per_cycle_elapsed", "IPC / SM"),
    ("smsp__issue_active.avg.pct", "issue active %"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM thr %"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM thr %"),
    ("dram__bytes_read.sum", "DRAM read"),
    ("dram__bytes_write.sum", "DRAM write"),
    ("lts__t_sector_hit_rate.pct", "L2 hit %"),
    ("smsp__average_warp_latency_issue_stalled_short_scoreboard.pct", "stall short sb %"),
    ("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "stall long sb"),
    ("smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "stall short sb"),
    ("smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "stall wait"),
    ("smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio", "stall not selected"),
    ("smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio", "stall math throttle"),
    ("smsp__average_warps_issue_stalled_membar_per_issue_active.ratio", "stall membar"),
    ("smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio", "stall barrier"),
    ("smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio", "stall lg throttle"),
    ("smsp__inst_executed_pipe_alu.sum", "pipe alu"),
    ("smsp__inst_executed_pipe_fma.sum", "pipe fma"),
    ("smsp__inst_executed_pipe_fmaheavy.sum", "pipe fmaheavy"),
    ("smsp__inst_executed_pipe_lsu.sum", "pipe lsu"),
    ("sm__inst_executed_pipe_uniform.sum", "pipe uniform"),
    ("smsp__inst_executed_op_local_ld.sum", "local loads"),
    ("smsp__inst_executed_op_local_st.sum", "local stores"),
]
rows = {}
for path in sys.argv[1:]:
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    r = list(csv.reader(io.StringIO(out)))
    hdr, units, vals = r[0], r[1], r[2]
    d = {h: (v, u) for h, u, v in zip(hdr, units, vals)}
    name = d.get("Kernel Name", ("?", ""))[0]
    rows[path] = (name, d)
for path, (name, d) in rows.items():
    print(f"## {path.split('/')[-1]}  —  {name[:100]}")
    for k, label in KEYS:
        if k in d:
            print(f"  {label:22s} {d[k][0]:>16s} {d[k][1]}")
    print()
