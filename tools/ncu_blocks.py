#!/usr/bin/env python
"""Basic-block view of an `ncu --page source --csv --print-source sass` dump: consecutive SASS instructions with the same
executed count are one block; prints blocks by their share of all executed warp instructions.
    ncu -i prof.ncu-rep --page source --csv --print-source sass > sass.csv ; python tools/ncu_blocks.py sass.csv [min_share_pct]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
min_share = float(sys.argv[2]) if len(sys.argv) > 2 else 0.5
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
col = {h: i for i, h in enumerate(rows[hi])}
data = []
for r in rows[hi + 1:]:
    if len(r) < len(rows[hi]):
        continue
    try:
        data.append((int(r[col["Instructions Executed"]] or 0), int(r[col["# Samples"]] or 0), r[col["Source"]]))
    except ValueError:
        pass
total = sum(e for e, _, _ in data)
print(f"{len(data)} SASS instructions, {total:.4g} executed warp instructions")
blocks, start = [], 0
for i in range(1, len(data) + 1):
    if i == len(data) or data[i][0] != data[start][0]:
        blocks.append((start, i))
        start = i
for a, b in blocks:
    e = data[a][0]
    share = 100.0 * e * (b - a) / max(1, total)
    if share >= min_share:
        samples = sum(s for _, s, _ in data[a:b])
        ops = " ".join(x[2].split()[0] if not x[2].startswith("@") else x[2].split()[1] for x in data[a:b])
        print(f"#{a:5d}-{b:5d} n={b - a:3d} exec={e:.3g} share={share:5.1f}% samples={samples:6d} | {ops[:230]}")
