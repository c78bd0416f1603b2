#!/bin/bash
# GPU-box helper: the per-mesh tests, the Sponza per-mesh timing, then the whole GPU suite and the default bench line.
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "per_mesh" 2>&1 | tail -4
timeout 200 python profiles/per_mesh_bench.py > gpurun_out/per_mesh_1view.log 2>&1; tail -1 gpurun_out/per_mesh_1view.log | cut -c1-700
RCULL_NO_PDL=1 timeout 200 python profiles/per_mesh_bench.py > gpurun_out/per_mesh_1view_nopdl.log 2>&1; tail -1 gpurun_out/per_mesh_1view_nopdl.log | cut -c1-700
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 | tee gpurun_out/pytest_gpu.txt
timeout 600 python bench.py > gpurun_out/bench_default.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/bench_default.log | python -c "
import sys, json
d = json.loads(sys.stdin.read())
print(round(d['ms_per_step'], 4), round(d['roofline']['frac'], 4), [(a.get('workload'), round(a['ms_per_step'], 4)) for a in d['also']])"
