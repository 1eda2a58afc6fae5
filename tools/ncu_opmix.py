#!/usr/bin/env python
"""SASS opcode mix of the first kernel in an ncu report:  python tools/ncu_opmix.py rep.ncu-rep [divisor]"""
import collections, csv, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv"], capture_output=True, text=True).stdout
div = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
rows = list(csv.reader(out.splitlines()))
start = next(i for i, r in enumerate(rows) if "Source" in r and "Instructions Executed" in r)
hdr = rows[start]
col = {h: i for i, h in enumerate(hdr)}
ops, tot = collections.Counter(), 0.0
for r in rows[start + 1:]:
    if len(r) < len(hdr):
        if tot > 0:
            break
        continue
    try:
        n = float(r[col["Instructions Executed"]])
    except ValueError:
        continue
    sass = r[col["Source"]].strip().split()
    op = (sass[1] if sass[0].startswith("@") else sass[0]).split(".")[0]
    ops[op] += n
    tot += n
print("total warp instructions", tot)
for op, n in ops.most_common(24):
    print(f"{op:12s} {100 * n / tot:5.1f}%  {n / div:.1f}")
