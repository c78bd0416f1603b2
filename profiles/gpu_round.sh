#!/bin/bash
# GPU-box helper: parity tests, bench workloads, ncu launch list + full capture of the frame kernel.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.max.mem --format=csv > gpurun_out/smi.txt; nproc >> gpurun_out/smi.txt
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 | tee gpurun_out/pytest_gpu.txt
for w in 16m_5view 16m_1view 1m_1view; do
  timeout 400 python bench.py --steps 20 --warmup 5 --workload $w > gpurun_out/bench_$w.log 2>&1
  tail -1 gpurun_out/bench_$w.log | cut -c1-1500
done
timeout 400 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.log 2>&1; tail -1 gpurun_out/bench_ref.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_16m5view.csv python bench.py --steps 3 --warmup 3 > gpurun_out/ncu_launch.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:transform_cull_compact -s 3 -c 1 -o gpurun_out/frame_16m5view -f python bench.py --steps 3 --warmup 3 > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/ncu_full.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:transform_cull_compact -s 3 -c 1 -o gpurun_out/frame_16m1view -f python bench.py --steps 3 --warmup 3 --workload 16m_1view > gpurun_out/ncu_full1.log 2>&1
tail -3 gpurun_out/ncu_full1.log
