#!/usr/bin/env python
"""Per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv` launch list.
usage: launch_summary.py <launches.csv> "<command that was profiled>" > profiles/r01_launches_<tag>.txt"""
import collections, csv, os, sys

path, cmd = sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else "?"
rows = [r for r in csv.reader(l for l in open(path) if l.startswith('"'))]
ix = {h: i for i, h in enumerate(rows[0])}
tot, cnt = collections.Counter(), collections.Counter()
for r in rows[1:]:
    if r[ix["Metric Name"]] != "gpu__time_duration.sum":
        continue
    name = r[ix["Kernel Name"]].replace("rcull::", "")
    tot[name] += float(r[ix["Metric Value"]].replace(",", ""))
    cnt[name] += 1
total = sum(tot.values())
tag = os.path.basename(path).replace(".csv", "")
print(f"# {tag}: kernels launched by `{cmd}` under ncu (gpu__time_duration.sum, ns; cold-cache, serialised)")
for name, t in tot.most_common():
    print(f"{cnt[name]:5d} x {t / cnt[name]:12.1f} avg  {100 * t / total:5.1f}%  {name[:70]}")
