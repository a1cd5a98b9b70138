#!/usr/bin/env python
"""Per-kernel stall summary from an .ncu-rep (source page): total samples per stall reason and
the hottest instructions.  Usage: python profiles/ncu_stalls.py report.ncu-rep [top_n]"""
import csv
import subprocess
import sys

rep = sys.argv[1]
topn = int(sys.argv[2]) if len(sys.argv) > 2 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True,
                     text=True).stdout
rows = list(csv.reader(out.splitlines()))
print(rows[0][1][:100])
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
data = rows[2:]
num = lambda r, k: int(r[ix[k]] or 0)  # noqa: E731
tot = sum(num(r, "# Samples") for r in data)
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
agg = {s: sum(num(r, s) for r in data) for s in stalls}
print("samples", tot, "instructions", len(data), "warp-instr executed",
      sum(num(r, "Instructions Executed") for r in data))
print({k: f"{100 * v / tot:.1f}%" for k, v in sorted(agg.items(), key=lambda kv: -kv[1]) if v})
for r in sorted(data, key=lambda r: -num(r, "# Samples"))[:topn]:
    st = {s[6:]: num(r, s) for s in stalls if num(r, s) > 0}
    st = dict(sorted(st.items(), key=lambda kv: -kv[1])[:3])
    print(f'{r[ix["Address"]][-5:]} {num(r, "# Samples"):6d} {r[ix["Source"]][:64]:64s} {st}')
