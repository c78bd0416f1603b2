#!/usr/bin/env python
"""Short text summary of one `ncu --set full` capture (first kernel in the report): duration,
DRAM traffic, pipe utilisation, issue rate, stall reasons, top stalled SASS instructions.
usage: ncu_summary.py <report.ncu-rep> [workload-key-for-r01_traffic.json]
Writes nothing unless a workload key is given (then profiles/r01_traffic.json is updated with
dram read+write bytes per launch, the figure bench.py reports as roofline.traffic)."""
import csv, io, json, os, subprocess, sys, collections

rep = sys.argv[1]
key = sys.argv[2] if len(sys.argv) > 2 else None
NCU = os.environ.get("NCU", "ncu")


def page(*args):
    return subprocess.run([NCU, "-i", rep, *args], capture_output=True, text=True).stdout


rows = list(csv.reader(io.StringIO(page("--page", "raw", "--csv"))))
h, u, v = rows[0], rows[1], rows[2]
m = {n: (v[i], u[i]) for i, n in enumerate(h)}


def num(name):
    return float(m[name][0].replace(",", "")) if name in m and m[name][0] not in ("", "n/a") else float("nan")


UNIT = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "us": 1e-6, "ms": 1e-3, "ns": 1e-9, "s": 1, "second": 1}


def scaled(name):
    return num(name) * UNIT.get(m[name][1], 1)


dur = scaled("gpu__time_duration.sum")
rd, wr = scaled("dram__bytes_read.sum"), scaled("dram__bytes_write.sum")
print(f"kernel      : {m['Kernel Name'][0] if 'Kernel Name' in m else ''}")
print(f"duration    : {dur * 1e6:.1f} us (under ncu, cold caches)")
print(f"dram        : read {rd / 1e6:.1f} MB + write {wr / 1e6:.1f} MB = {(rd + wr) / 1e6:.1f} MB -> {(rd + wr) / dur / 1e9:.0f} GB/s")
print(f"grid/block  : {m.get('Grid Size', ('?',))[0]} x {m.get('Block Size', ('?',))[0]}, regs/thread {m.get('launch__registers_per_thread', ('?',))[0]}, "
      f"dyn smem {m.get('launch__shared_mem_per_block_dynamic', ('?', ''))[0]} {m.get('launch__shared_mem_per_block_dynamic', ('', ''))[1]}")
for name in ("smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
             "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
             "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.per_cycle_active",
             "smsp__inst_executed.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed"):
    if name in m:
        print(f"{name:75s} {m[name][0]} {m[name][1]}")
stalls = {n[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")]: num(n) for n in m
          if n.startswith("smsp__average_warps_issue_stalled_") and n.endswith("_per_issue_active.ratio")}
print("stalls per issue: " + ", ".join(f"{k}={x:.2f}" for k, x in sorted(stalls.items(), key=lambda kv: -kv[1])[:8]))

# SASS view: opcode mix and the most-sampled instructions
rows = list(csv.reader(io.StringIO(page("--page", "source", "--csv", "--print-source", "sass"))))
try:
    start = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[start]
    ix = {x: i for i, x in enumerate(hdr)}
    body = [r for r in rows[start + 1:] if r and len(r) >= len(hdr) - 5]
    ops = collections.Counter()
    top = []
    for r in body:
        t = r[ix["Source"]].strip()
        parts = t.split()
        if not parts:
            continue
        op = (parts[1] if parts[0].startswith("@") else parts[0]).split(".")[0]
        ops[op] += int(r[ix["Instructions Executed"]] or 0)
        top.append((int(r[ix["# Samples"]] or 0), t))
    tot = sum(ops.values())
    print(f"SASS: {len(body)} static instructions; executed mix (instrumented pass, spin loops over-counted): " +
          ", ".join(f"{k} {100 * c / tot:.1f}%" for k, c in ops.most_common(12)))
    ts = sum(s for s, _ in top)
    print("most-sampled instructions:")
    for s, t in sorted(top, reverse=True)[:8]:
        print(f"   {100 * s / max(ts, 1):5.1f}%  {t[:90]}")
    has = lambda o: any(k.startswith(o) for k in ops)
    print("TMA / packed-FP32 mnemonics present: " + ", ".join(o for o in ("UBLKCP", "UTMASTG", "SYNCS", "FMUL2", "FADD2", "REDUX") if has(o)))
except StopIteration:
    print("(no SASS page in this report)")

if key:
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "r01_traffic.json")
    d = json.load(open(path)) if os.path.exists(path) else {}
    d[key] = int(rd + wr)
    json.dump(d, open(path, "w"), indent=1, sort_keys=True)
