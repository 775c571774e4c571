"""Summarise an `ncu --page source --csv` export: executed-instruction histogram by opcode and stall reasons.
usage: ncu -i X.ncu-rep --page source --csv > src.csv ; python scripts/ncu_source_summary.py src.csv [kernel-substring]"""
import collections
import csv
import re
import sys

rows = list(csv.reader(open(sys.argv[1])))
want = sys.argv[2] if len(sys.argv) > 2 else None
# the export holds one block per profiled launch: "Kernel Name" row, header row, instruction rows
blocks, cur = [], None
for r in rows:
    if r and r[0] == "Kernel Name":
        cur = {"name": r[1], "hdr": None, "rows": []}
        blocks.append(cur)
    elif cur is not None and cur["hdr"] is None:
        cur["hdr"] = r
    elif cur is not None:
        cur["rows"].append(r)
blk = [b for b in blocks if want is None or want in b["name"]][0]
hdr = blk["hdr"]
ix = {h: i for i, h in enumerate(hdr)}
stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
ops, st, ninst, wf, wf_ideal = collections.Counter(), collections.Counter(), 0, 0, 0
for r in blk["rows"]:
    if len(r) < len(hdr):
        continue
    n = int(r[ix["Instructions Executed"]])
    ninst += n
    m = re.match(r"\s*(@!?U?P\d+\s+)?([A-Z0-9_.]+)", r[ix["Source"]])
    ops[(m.group(2) if m else r[ix["Source"]]).split(".")[0]] += n
    for c in stall_cols:
        st[c] += int(r[ix[c]])
    wf += int(r[ix["L1 Wavefronts Shared"]] or 0)
    wf_ideal += int(r[ix["L1 Wavefronts Shared Ideal"]] or 0)
print(f"kernel: {blk['name']}\nstatic SASS instructions: {len(blk['rows'])}  executed warp-instructions: {ninst}")
print(f"shared wavefronts: {wf} (ideal {wf_ideal})")
for k, v in ops.most_common(30):
    print(f"  {k:12s} {v:10d} {100 * v / ninst:5.1f}%")
s = sum(st.values())
print("warp stall samples (all):")
for k, v in st.most_common():
    if v:
        print(f"  {k:28s} {v:8d} {100 * v / s:5.1f}%")
