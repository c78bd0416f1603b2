#!/usr/bin/env python
"""Per-instruction view of an `ncu --page source --csv --print-source sass` export, for finding WHERE a kernel waits:
prints every instruction with >= a share of the samples, with its dominant stall reasons and a window of context, and
totals by stall reason.   usage: sass_regions.py export.csv [min_pct]"""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
minpct = float(sys.argv[2]) if len(sys.argv) > 2 else 0.6
start = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[start]; ix = {h: i for i, h in enumerate(hdr)}
body = [r for r in rows[start + 1:] if r and len(r) > ix["# Samples"]]
stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
tot = sum(int(r[ix["# Samples"]] or 0) for r in body)
by = collections.Counter()
for r in body:
    for h in stall_cols:
        by[h] += int(r[ix[h]] or 0)
print("total samples", tot, "; by reason:", ", ".join(f"{k[6:]} {100*v/tot:.1f}%" for k, v in by.most_common(10)))
execd = sum(int(r[ix["Instructions Executed"]] or 0) for r in body)
print("warp instructions executed", execd)
for i, r in enumerate(body):
    n = int(r[ix["# Samples"]] or 0)
    if 100.0 * n / tot >= minpct:
        st = sorted(((int(r[ix[h]] or 0), h[6:]) for h in stall_cols), reverse=True)[:3]
        print(f"{100*n/tot:5.2f}%  #{i:4d} exec {int(r[ix['Instructions Executed']] or 0):9d}  {r[ix['Source']].strip():60s} " + " ".join(f"{k}={v}" for v, k in st if v))
