#!/bin/bash
# GPU-box helper: parity tests + the three bench workloads (one summary line each).
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
for w in ${WORKLOADS:-16m_5view 16m_1view 1m_1view}; do
  timeout 400 python bench.py --steps 20 --warmup 5 --workload $w > gpurun_out/bench_$w.log 2>&1
  python - "$w" <<'PY'
import json, sys
w = sys.argv[1]
txt = open(f"gpurun_out/bench_{w}.log").read()
try:
    d = json.loads(txt.strip().splitlines()[-1])
    print(w, "ms/step %.4f" % d["ms_per_step"], "frac %.3f" % d["roofline"]["frac"], "value %.4g" % d["value"], "e2e %.3g" % d["e2e"]["value"], "cpu %.3g" % d["cpu_baseline"]["value"])
except Exception:
    print(txt[-2000:])
PY
done
