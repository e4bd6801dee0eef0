#!/usr/bin/env python3
"""Summarise an `ncu --page source --csv` SASS dump: stall-reason totals and the hottest
instructions.  usage: ncu_stalls.py dump.csv [top_n]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
h = rows[hi]
data = []
for r in rows[hi + 1:]:
    if not r or r[0] in ("Kernel Name", "Address"):
        break  # first kernel instance only
    if len(r) >= len(h) - 2:
        data.append(r)
si = h.index("# Samples"); src = h.index("Source"); ex = h.index("Instructions Executed")
stall_cols = [(i, c) for i, c in enumerate(h) if c.startswith("stall_") and "Not Issued" not in c]
tot = sum(int(r[si] or 0) for r in data)
print("kernel:", rows[0][1][:80], " samples:", tot, " instr:", sum(int(r[ex] or 0) for r in data))
agg = {c: sum(int(r[i] or 0) for r in data) for i, c in stall_cols}
print("stalls:", ", ".join(f"{c[6:]} {100*v/max(tot,1):.1f}%" for c, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]))
order = sorted(range(len(data)), key=lambda k: -int(data[k][si] or 0))[:top]
for k in sorted(order):
    r = data[k]
    st = sorted(((int(r[i] or 0), c[6:]) for i, c in stall_cols), reverse=True)[:2]
    print(f"{k:5d} {100*int(r[si])/max(tot,1):5.1f}% {r[src].strip()[:70]:70s} {st[0][1]}:{st[0][0]} {st[1][1]}:{st[1][0]}")
