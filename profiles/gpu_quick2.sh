#!/bin/bash
# quick check of a frame-kernel change: the parity tests that exercise it, then the resident timings
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_zy_golden_stage2.py -m gpu -x -q -k "not 16m or rank0" 2>&1 | tail -5
timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu > gpurun_out/bench_quick.log 2>&1; tail -1 gpurun_out/bench_quick.log | python -c "
import sys, json
l = json.loads(sys.stdin.read())
print('16m_5view', round(l['ms_per_step'], 4), 'frac', round(l['roofline']['frac'], 3))
for a in l.get('also', []):
    print(a['workload'], round(a['ms_per_step'], 4), a.get('roofline_frac') or (a.get('roofline') or {}).get('frac'))
"
