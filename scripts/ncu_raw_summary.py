"""Pick the roofline-relevant metrics out of `ncu -i X.ncu-rep --page raw --csv` and write profiles/ncu_summary.json.
usage: python scripts/ncu_raw_summary.py raw.csv kernel_key out.json [tag]"""
import csv
import json
import os
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr, units, launches = rows[0], rows[1], rows[2:]
key, out = sys.argv[2], sys.argv[3]
tag = sys.argv[4] if len(sys.argv) > 4 else ""
KEEP = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__registers_per_thread",
        "smsp__inst_executed.sum", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__cycles_elapsed.max", "launch__grid_size", "launch__block_size",
        "launch__shared_mem_per_block_dynamic", "sm__inst_executed.avg.per_cycle_elapsed"]
ix = {h: i for i, h in enumerate(hdr)}
per = []
for r in launches:
    if len(r) < len(hdr):
        continue
    d = {}
    for k in KEEP:
        if k in ix:
            try:
                d[k] = float(r[ix[k]].replace(",", ""))
            except ValueError:
                d[k] = r[ix[k]]
            d[k + " [unit]"] = units[ix[k]]
    per.append(d)
def mb(v, u):
    return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
traffic = [mb(d["dram__bytes_read.sum"], d["dram__bytes_read.sum [unit]"]) + mb(d["dram__bytes_write.sum"], d["dram__bytes_write.sum [unit]"]) for d in per]
summary = {}
if os.path.exists(out):
    summary = json.load(open(out))
summary[key] = {"tag": tag, "launches_captured": len(per), "dram_bytes_per_launch": sum(traffic) / len(traffic),
                "note": "ncu --set full --clock-control none; dram__bytes_read.sum + dram__bytes_write.sum per launch", "metrics": per}
json.dump(summary, open(out, "w"), indent=1)
print(json.dumps({k: v for k, v in summary[key].items() if k != "metrics"}))
