#!/usr/bin/env python
"""Hot spots of an `ncu --page source --csv` export (SASS view): the instructions with the most warp-stall samples, with
the dominant stall reason of each, and totals per stall reason.   tools/ncu_hot.py <source.csv> [top N]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
# a report with several launches has one section per launch ("Kernel Name" row, header row, instructions): take the first
starts = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"]
rows = rows[starts[0]:(starts[1] if len(starts) > 1 else len(rows))]
hdr = rows[1]
col = {h: i for i, h in enumerate(hdr)}
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
data = []
tot = {s: 0 for s in stalls}
for k, r in enumerate(rows[2:]):
    if len(r) < len(hdr) or not r[col["# Samples"]].strip().isdigit():
        continue
    n = int(r[col["# Samples"]] or 0)
    per = {s: int(r[col[s]] or 0) for s in stalls}
    for s in stalls:
        tot[s] += per[s]
    data.append((n, k, r[col["Source"]].strip(), per, int(r[col["Instructions Executed"]] or 0)))
total = sum(d[0] for d in data)
print(rows[0][1][:100], "| samples", total)
print("by reason:", ", ".join(f"{s[6:]} {100*v/total:.1f}%" for s, v in sorted(tot.items(), key=lambda kv: -kv[1])[:10]))
for n, k, src, per, ex in sorted(data, key=lambda d: -d[0])[:top]:
    why = ", ".join(f"{s[6:]} {v}" for s, v in sorted(per.items(), key=lambda kv: -kv[1])[:3] if v)
    print(f"{100*n/total:5.1f}%  #{k:5d} x{ex:7d}  {src[:70]:70s} {why}")
