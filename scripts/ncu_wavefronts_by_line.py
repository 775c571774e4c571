"""Shared-memory wavefronts of the streaming kernel per source line and per live group, from an `ncu --page source --csv`
export and nvdisasm's line info (same inputs as scripts/ncu_line_profile.py):
    python scripts/ncu_wavefronts_by_line.py src.csv dis.txt <exact-mangled-kernel-substring> <live-groups-per-launch> [top-n]
One wavefront = one cycle of the SM's L1 data pipe (profiles/r02e/microbench_lsu.md)."""
import collections
import csv
import os
import re
import sys

src_csv, dis_txt, want, groups = sys.argv[1], sys.argv[2], sys.argv[3], float(sys.argv[4])
topn = int(sys.argv[5]) if len(sys.argv) > 5 else 16
lines_of, cur, inside, seen = [], ("?", 0), False, False
for ln in open(dis_txt):
    if ln.lstrip().startswith(".section"):
        was = inside
        inside = (".text." in ln) and (want in ln) and not seen
        if was and not inside:
            seen = True  # first matching section only
        continue
    if not inside:
        continue
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)', ln)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2)))
    elif re.match(r"\s+/\*[0-9a-f]{4,6}\*/\s+\S", ln):
        lines_of.append(cur)
rows = list(csv.reader(open(src_csv)))
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
body = rows[2:2 + len(lines_of)]
wf, ins, ideal = collections.Counter(), collections.Counter(), collections.Counter()
for k, r in zip(lines_of, body):
    wf[k] += float(r[ix["L1 Wavefronts Shared"]] or 0)
    ideal[k] += float(r[ix["L1 Wavefronts Shared Ideal"]] or 0)
    if float(r[ix["L1 Wavefronts Shared"]] or 0) > 0:
        ins[k] += float(r[ix["Instructions Executed"]] or 0)
print(f"shared-memory wavefronts per live group: {sum(wf.values()) / groups:.1f} (ideal {sum(ideal.values()) / groups:.1f})")
print(f"{'file:line':28s} {'wavefronts/group':>16s} {'LDS+STS instr/group':>20s}")
for k, v in wf.most_common(topn):
    print(f"{k[0] + ':' + str(k[1]):28s} {v / groups:16.1f} {ins[k] / groups:20.1f}")
