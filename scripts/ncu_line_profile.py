"""Attribute an `ncu --page source --csv` export (SASS rows) to CUDA source lines using nvdisasm's line info.
usage: cuobjdump -xelf all libpsla_frontend.so; nvdisasm --print-line-info X.cubin > dis.txt
       python scripts/ncu_line_profile.py src.csv dis.txt <mangled-kernel-substring> [top-n]
Rows are matched by position (the i-th instruction of the kernel in both listings)."""
import collections
import csv
import os
import re
import sys

src_csv, dis_txt, want = sys.argv[1], sys.argv[2], sys.argv[3]
topn = int(sys.argv[4]) if len(sys.argv) > 4 else 60

# ---- nvdisasm: one (file, line) per instruction of the wanted kernel ----
lines_of = []
cur, inside = ("?", 0), False
for ln in open(dis_txt):
    if ln.startswith(".section") or ln.startswith("\t.section"):
        inside = (".text." in ln) and (want in ln)
        continue
    if not inside:
        continue
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)', ln)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2)))
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,6}\*/\s+\S", ln):
        lines_of.append(cur)

# ---- ncu source page: first block (first profiled launch) ----
rows = list(csv.reader(open(src_csv)))
blocks, b = [], None
for r in rows:
    if r and r[0] == "Kernel Name":
        b = {"hdr": None, "rows": []}
        blocks.append(b)
    elif b is not None and b["hdr"] is None:
        b["hdr"] = r
    elif b is not None and len(r) >= len(b["hdr"]):
        b["rows"].append(r)
blk = blocks[0]
ix = {h: i for i, h in enumerate(blk["hdr"])}
n = min(len(lines_of), len(blk["rows"]))
if len(lines_of) != len(blk["rows"]):
    print(f"warning: {len(lines_of)} disassembled vs {len(blk['rows'])} profiled instructions", file=sys.stderr)
inst, samp, stat = collections.Counter(), collections.Counter(), collections.Counter()
for i in range(n):
    r = blk["rows"][i]
    k = lines_of[i]
    inst[k] += int(r[ix["Instructions Executed"]])
    samp[k] += int(r[ix["# Samples"]])
    stat[k] += 1
ti, ts = sum(inst.values()), sum(samp.values())
print(f"executed warp-instructions {ti}, samples {ts}, static {n}")
print(f"{'file:line':40s} {'static':>6s} {'executed':>10s} {'%':>6s} {'samples':>8s} {'%':>6s}")
for k, v in sorted(inst.items(), key=lambda kv: -samp[kv[0]])[:topn]:
    print(f"{k[0] + ':' + str(k[1]):40s} {stat[k]:6d} {v:10d} {100 * v / ti:6.2f} {samp[k]:8d} {100 * samp[k] / max(ts, 1):6.2f}")

# ---- optional phase roll-up: PSLA_PHASES="name:file:lo-hi,..." ----
spec = os.environ.get("PSLA_PHASES")
if spec:
    ph_i, ph_s = collections.Counter(), collections.Counter()
    rules = []
    for item in spec.split(","):
        name, f, rng = item.split(":")
        lo, hi = rng.split("-")
        rules.append((name, f, int(lo), int(hi)))
    for k in inst:
        name = next((nm for nm, f, lo, hi in rules if k[0] == f and lo <= k[1] <= hi), "other:" + k[0])
        ph_i[name] += inst[k]
        ph_s[name] += samp[k]
    print("\nphase roll-up")
    for name, v in sorted(ph_i.items(), key=lambda kv: -kv[1]):
        print(f"{name:40s} {v:10d} {100 * v / ti:6.2f}% inst   {ph_s[name]:6d} {100 * ph_s[name] / max(ts, 1):6.2f}% samples")
