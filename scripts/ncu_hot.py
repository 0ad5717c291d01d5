#!/usr/bin/env python
"""Summarise an ncu --page source --csv export: top SASS instructions by stall samples.
usage: ncu -i X.ncu-rep --page source --csv > src.csv ; python scripts/ncu_hot.py src.csv [N]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 40
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
body = [r for r in rows[2:] if len(r) == len(hdr)]
tot = sum(int(r[ix["# Samples"]] or 0) for r in body)
stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
print("total samples", tot, "instructions", len(body))
agg = {c: sum(int(r[ix[c]] or 0) for r in body) for c in stall_cols}
print("by reason:", {k: v for k, v in sorted(agg.items(), key=lambda kv: -kv[1]) if v})
order = sorted(range(len(body)), key=lambda i: -int(body[i][ix["# Samples"]] or 0))[:n]
for i in sorted(order):
    r = body[i]
    s = int(r[ix["# Samples"]] or 0)
    top = sorted(((int(r[ix[c]] or 0), c) for c in stall_cols), reverse=True)[:2]
    print(f"{i:5d} {s:6d} {100*s/tot:5.1f}%  ex={r[ix['Instructions Executed']]:>10}  {r[ix['Source']].strip()[:90]:90s} {top}")
