#!/usr/bin/env python
"""Summarise an `ncu --page source --csv` dump: top SASS instructions by stall samples.
    ncu -i prof.ncu-rep --page source --csv > src.csv ; python tools/ncu_src_top.py src.csv [N]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 40
hdr_i = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hdr_i]
col = {h: i for i, h in enumerate(hdr)}
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
data = []
for r in rows[hdr_i + 1:]:
    if len(r) < len(hdr):
        continue
    try:
        s = int(r[col["# Samples"]] or 0)
    except ValueError:
        continue
    data.append((s, r))
total = sum(s for s, _ in data)
print(f"total samples {total}, instructions {len(data)}")
agg = {h: 0 for h in stalls}
for s, r in data:
    for h in stalls:
        try:
            agg[h] += int(r[col[h]] or 0)
        except ValueError:
            pass
print("stall totals:", {k: v for k, v in sorted(agg.items(), key=lambda x: -x[1]) if v})
idx = {id(r): i for i, (_, r) in enumerate(data)}
for s, r in sorted(data, key=lambda x: -x[0])[:n]:
    top = sorted(((int(r[col[h]] or 0), h) for h in stalls), reverse=True)[:2]
    print(f"{s:7d} {100.0 * s / max(1, total):5.1f}%  #{idx[id(r)]:5d} exec={r[col['Instructions Executed']]:>10}  "
          f"{top[0][1]}={top[0][0]} {top[1][1]}={top[1][0]}  {r[col['Source']][:90]}")
