#!/usr/bin/env python
"""Condenses an `ncu --set full` report into profiles/*.json: one record per distinct kernel
(the LAST launch of each), with the metrics the roofline discussion in DESIGN.md uses.
Usage: python profiles/ncu_summary.py report.ncu-rep out.json "source description" """
import csv
import json
import subprocess
import sys

KEEP = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "l1tex__throughput.avg.pct_of_peak_sustained_active",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block_dynamic"]

rep, out, src = sys.argv[1], sys.argv[2], sys.argv[3]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True,
                     text=True, check=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]
col = {h: i for i, h in enumerate(hdr)}
last = {}
for r in rows[2:]:
    name = r[col["Kernel Name"]]
    short = name.split("(")[0].replace("void ", "").replace("<unnamed>::", "").replace("frodo::", "")
    rec = {"kernel": short}
    for k in KEEP:
        if k in col:
            rec[k] = f"{r[col[k]]} {units[col[k]]}".strip()
    last[short] = rec
json.dump({"source": src, "kernels": list(last.values())}, open(out, "w"), indent=1)
print(out, len(last), "kernels")
