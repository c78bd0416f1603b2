#!/bin/bash
# A/B helper: bench the variant libraries under renderer_b200/variants against the default build.
mkdir -p gpurun_out
for lib in "" $(ls renderer_b200/variants/*.so 2>/dev/null); do
  for w in ${WORKLOADS:-16m_5view 16m_1view}; do
    RCULL_LIB=${lib:+$PWD/$lib} timeout 400 python bench.py --steps 30 --warmup 5 --workload $w > gpurun_out/ab.log 2>&1
    python - "$w" "${lib:-default}" <<'PY'
import json, sys
txt = open("gpurun_out/ab.log").read()
try:
    d = json.loads(txt.strip().splitlines()[-1])
    print(sys.argv[2], sys.argv[1], "ms/step %.4f" % d["ms_per_step"], "frac %.3f" % d["roofline"]["frac"])
except Exception:
    print(txt[-1500:])
PY
  done
done
