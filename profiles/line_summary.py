#!/usr/bin/env python
"""Per source line totals (instructions executed, stall samples) from an
`ncu --page source --csv --print-source cuda,sass` export."""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 40
inst = collections.Counter(); samp = collections.Counter(); text = {}
fname = None; ix = None
for r in rows:
    if not r: continue
    if r[0] == "File Path": fname = r[1].split("/")[-1]; continue
    if r[0] == "Line No": ix = {h: i for i, h in enumerate(r)}; src_cols = [i for i, h in enumerate(r) if h == "Source"]; continue
    if ix is None or len(r) < len(ix): continue
    if not r[0].isdigit(): continue
    key = (fname, int(r[0]))
    text.setdefault(key, r[src_cols[0]].strip())
    try:
        inst[key] += int(r[ix["Instructions Executed"]] or 0); samp[key] += int(r[ix["# Samples"]] or 0)
    except ValueError:
        pass
tot = sum(inst.values()); ts = sum(samp.values())
print("total inst", tot, "samples", ts)
for key, v in inst.most_common(n):
    print(f"{key[0]}:{key[1]:4d} inst {v:11d} {100*v/tot:5.1f}%  samp {100*samp[key]/max(ts,1):5.1f}%  {text[key][:90]}")
