#!/usr/bin/env python3
"""Hottest SASS lines of an .ncu-rep by warp-stall samples, and the opcode mix weighted by
executed instructions: scripts/ncu_hot.py file.ncu-rep [top]"""
import collections
import csv
import io
import subprocess
import sys

out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
print(rows[0][1])
hdr = rows[1]
ia, isrc, ist, iex = hdr.index("Address"), hdr.index("Source"), hdr.index("Warp Stall Sampling (All Samples)"), hdr.index("Instructions Executed")
body = [r for r in rows[2:] if len(r) > iex]
tot_s = sum(int(r[ist] or 0) for r in body)
tot_i = sum(int(r[iex] or 0) for r in body)
print(f"total stall samples {tot_s}, warp instructions executed {tot_i}")
mix = collections.Counter()
smp = collections.Counter()
for r in body:
    op = r[isrc].split()
    op = [o for o in op if not o.startswith("@")]
    name = op[0].split(".")[0] if op else "?"
    mix[name] += int(r[iex] or 0)
    smp[name] += int(r[ist] or 0)
print("opcode mix (share of executed warp instructions | share of stall samples):")
for k, v in mix.most_common(14):
    print(f"  {k:10s} {100 * v / tot_i:5.1f} % | {100 * smp[k] / max(1, tot_s):5.1f} %")
top = int(sys.argv[2]) if len(sys.argv) > 2 else 12
print(f"top {top} lines by stall samples:")
for r in sorted(body, key=lambda r: -int(r[ist] or 0))[:top]:
    print(f"  {100 * int(r[ist]) / max(1, tot_s):4.1f} %  {r[isrc].strip()[:90]}")
