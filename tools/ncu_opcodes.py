#!/usr/bin/env python
"""Executed warp instructions of an ncu capture by SASS opcode, with the local-memory (spill) share:
    python tools/ncu_opcodes.py prof.ncu-rep [top]"""
import csv
import subprocess
import sys
from collections import Counter

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
col = {n: i for i, n in enumerate(rows[hi])}
ops, tot = Counter(), 0
for r in rows[hi + 1:]:
    if len(r) < len(rows[hi]):
        continue
    n = int(float(r[col["Instructions Executed"]] or 0))
    src = r[col["Source"]].split()
    if not src:
        continue
    op = src[1] if src[0].startswith("@") and len(src) > 1 else src[0]
    ops[op.split(".")[0]] += n
    tot += n
print(f"total warp-inst {tot:,}")
for op, n in ops.most_common(top):
    print(f"  {op:12s} {n:>16,} {100 * n / tot:6.2f} %")
loc = ops["STL"] + ops["LDL"]
print(f"local memory: STL {ops['STL']:,}  LDL {ops['LDL']:,}  = {100 * loc / tot:.2f} % of instructions")
