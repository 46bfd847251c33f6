#!/usr/bin/env python3
"""Loop bodies of a SASS listing (cuobjdump -sass -fun ...): address range, instruction count and
opcode histogram of every backward branch.  Usage: sass_loops.py file.sass [min_instructions]"""
import collections
import re
import sys

ins = []
for line in open(sys.argv[1]):
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
    if m:
        ins.append((int(m.group(1), 16), m.group(2).strip()))
minlen = int(sys.argv[2]) if len(sys.argv) > 2 else 50
addr_index = {a: i for i, (a, _) in enumerate(ins)}
for i, (a, t) in enumerate(ins):
    m = re.search(r"\bBRA\b.*?(0x[0-9a-f]+)", t)
    if not m:
        continue
    tgt = int(m.group(1), 16)
    if tgt < a and tgt in addr_index:
        body = ins[addr_index[tgt]:i + 1]
        if len(body) < minlen:
            continue
        hist = collections.Counter()
        for _, s in body:
            s = re.sub(r"^@!?U?P\d+\s+", "", s)
            hist[s.split()[0].split(".")[0]] += 1
        print(f"loop {tgt:#x}..{a:#x}: {len(body)} instructions")
        print("   " + "  ".join(f"{k}:{v}" for k, v in hist.most_common()))
