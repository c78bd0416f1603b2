#!/usr/bin/env python
"""Summarise an `ncu --page source --csv --print-source sass` export: executed
warp instructions by opcode, and stall samples by reason (first kernel in the file)."""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
# find header row
start = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[start]
body = []
for r in rows[start + 1:]:
    if not r or r[0] in ("Kernel Name", "Address"):
        break
    body.append(r)
ix = {h: i for i, h in enumerate(hdr)}
ops = collections.Counter(); samples = collections.Counter()
tot = 0
for r in body:
    src = r[ix["Source"]].strip()
    parts = src.split()
    op = parts[1] if parts[0].startswith("@") else parts[0]
    op = op.split(".")[0]
    n = int(r[ix["Instructions Executed"]] or 0)
    ops[op] += n; tot += n
    samples[op] += int(r[ix["# Samples"]] or 0)
print("total warp instructions", tot, "static instructions", len(body))
for op, n in ops.most_common(30):
    print(f"{op:10s} {n:12d} {100*n/tot:5.1f}%  samples {samples[op]}")
stalls = collections.Counter()
for h in hdr:
    if h.startswith("stall_") and "Not Issued" not in h:
        stalls[h] = sum(int(r[ix[h]] or 0) for r in body)
ts = sum(stalls.values())
print("stall samples:", ", ".join(f"{k[6:]}={100*v/ts:.1f}%" for k, v in stalls.most_common(10)))
