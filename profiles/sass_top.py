#!/usr/bin/env python
"""Top sampled SASS instructions (with neighbours) of an ncu source-page CSV export."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 15
start = next(i for i, r in enumerate(rows) if r and r[0] == 'Address')
hdr = rows[start]; ix = {h: i for i, h in enumerate(hdr)}
body = []
for r in rows[start + 1:]:
    if not r or r[0] in ("Kernel Name", "Address"): break
    body.append(r)
top = sorted(range(len(body)), key=lambda i: -int(body[i][ix['# Samples']] or 0))[:n]
for i in top:
    r = body[i]
    print(i, r[ix['Instructions Executed']], r[ix['# Samples']], r[ix['Source']].strip(), '| long_sb', r[ix['stall_long_sb']], 'wait', r[ix['stall_wait']], '| prev:', body[i-1][ix['Source']].strip()[:70])
