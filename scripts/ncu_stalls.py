#!/usr/bin/env python
"""Top stalled SASS instructions of one kernel from `ncu --page source --csv` output (all kernels in one file)."""
import csv
import sys

path, want, nth = sys.argv[1], sys.argv[2], int(sys.argv[3]) if len(sys.argv) > 3 else 0
topn = int(sys.argv[4]) if len(sys.argv) > 4 else 40
rows = list(csv.reader(open(path)))
blocks, cur = [], None
for r in rows:
    if r and r[0] == "Kernel Name":
        cur = {"name": r[1], "rows": []}
        blocks.append(cur)
    elif cur is not None:
        cur["rows"].append(r)
sel = [b for b in blocks if want in b["name"]]
b = sel[nth]
hdr, data = b["rows"][0], [r for r in b["rows"][1:] if len(r) == len(b["rows"][0])]
iS, iSrc, iEx = hdr.index("# Samples"), hdr.index("Source"), hdr.index("Instructions Executed")
tot = sum(int(r[iS]) for r in data)
print(b["name"][:90], "| samples", tot, "| SASS instrs", len(data), "| warp-instr executed", sum(int(r[iEx]) for r in data))
stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
agg = {}
for r in data:
    for j in stall_cols:
        if r[j] not in ("", "0"):
            agg[hdr[j]] = agg.get(hdr[j], 0) + int(r[j])
print("stall totals:", sorted(agg.items(), key=lambda kv: -kv[1])[:8])
order = sorted(range(len(data)), key=lambda i: -int(data[i][iS]))[:topn]
for i in sorted(order):
    r = data[i]
    st = {hdr[j]: int(r[j]) for j in stall_cols if r[j] not in ("", "0")}
    top = sorted(st.items(), key=lambda kv: -kv[1])[:2]
    print(f"{i:5d} {int(r[iS]):6d} {100.0*int(r[iS])/tot:5.1f}% ex={r[iEx]:>8s}  {r[iSrc][:64]:64s} {top}")
