"""Opcode mix of one kernel from an ncu report's source page (developer tool).
    ncu -i rep.ncu-rep --page source --csv --kernel-name regex:NAME --launch-count 1 > k.csv
    python tools/ncu_opcode_mix.py k.csv [units]        units = divide counts by this (e.g. pixels)"""
import csv
import sys
from collections import Counter

rows = list(csv.reader(open(sys.argv[1])))
units = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
hdr = rows[1]
i_src, i_ex, i_smp = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
ops, smp = Counter(), Counter()
tot = 0
for r in rows[2:]:
    if len(r) <= i_ex:
        continue
    if r[i_ex] == "Instructions Executed":   # the next view (PTX / source) starts: SASS only
        break
    toks = r[i_src].split()
    if not toks:
        continue
    op = toks[1] if toks[0].startswith("@") and len(toks) > 1 else toks[0]
    op = op.split(".")[0] + ("." + op.split(".")[1] if op.startswith(("LDG", "STG", "LDS", "STS", "SHFL", "CREDUX", "REDUX", "LDGSTS", "SYNCS", "BAR", "UBLKCP")) and "." in op else "")
    n = int(float(r[i_ex] or 0))
    ops[op] += n
    smp[op] += int(float(r[i_smp] or 0))
    tot += n
print(f"total warp instructions {tot}  ({tot / units:.2f} per unit)")
ts = sum(smp.values()) or 1
for op, n in ops.most_common(40):
    print(f"  {op:14s} {n:12d}  {100.0 * n / tot:5.1f} %   {n / units:8.2f}/unit   stall samples {100.0 * smp[op] / ts:5.1f} %")
