#!/usr/bin/env python
"""Executed instructions per live group by SASS region (64 instructions each) and class, from an `ncu --page source --csv`
export: where the integer / control overhead of the streaming kernel sits.  usage: python scripts/ncu_regions.py src.csv [groups]"""
import collections
import csv
import re
import sys

rows = list(csv.reader(open(sys.argv[1])))
G = float(sys.argv[2]) if len(sys.argv) > 2 else 22900.0
hdr, data = rows[1], []
for r in rows[2:]:
    if r and r[0] == "Kernel Name":
        break
    if len(r) >= len(hdr):
        data.append(r)
FP = {"FFMA2", "FADD2", "FMUL2", "FFMA", "FADD", "FMUL", "MUFU", "FMNMX"}
MEM = {"LDS", "STS", "LDG", "STG", "SHFL", "LDC", "LDCU"}
tot = collections.Counter()
for s in range(0, len(data), 64):
    c, ops = collections.Counter(), collections.Counter()
    for r in data[s:s + 64]:
        op = re.sub(r"^@!?U?P\d+\s+", "", r[1].strip()).split()[0].split(".")[0]
        n = int(float(r[5] or 0)) / G
        cls = "fp" if op in FP else ("mem" if op in MEM else "int")
        c[cls] += n
        tot[cls] += n
        if cls == "int":
            ops[op] += n
    if sum(c.values()) > 3:
        top = " ".join(f"{k}:{v:.0f}" for k, v in ops.most_common(6) if v >= 1)
        print(f"{s:5d} int {c['int']:6.1f} mem {c['mem']:6.1f} fp {c['fp']:6.1f} | {top}")
print("per group:", {k: round(v, 1) for k, v in tot.items()}, "all", round(sum(tot.values()), 1))
