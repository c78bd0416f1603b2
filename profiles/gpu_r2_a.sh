#!/bin/bash
# round 2, call A: smoke, full GPU parity suite, bench (packed-sum plane tests vs scalar sums), pipe micro-benchmark,
# one full ncu capture of the 5-view frame kernel
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.max.mem --format=csv > gpurun_out/smi.txt; nproc >> gpurun_out/smi.txt
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -2
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 | tee gpurun_out/pytest_gpu.txt
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_default.log 2>&1; tail -1 gpurun_out/bench_default.log | cut -c1-600
RCULL_NO_PACKED_SUMS=1 timeout 300 python bench.py --steps 20 --warmup 5 --no-also --no-cpu > gpurun_out/bench_scalar_sums.log 2>&1; tail -1 gpurun_out/bench_scalar_sums.log | cut -c1-300
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.log 2>&1; tail -1 gpurun_out/bench_ref.log | cut -c1-300
timeout 120 profiles/micro/pipes > gpurun_out/pipes.txt 2>&1; cat gpurun_out/pipes.txt
timeout 300 python bench.py --no-also --no-cpu --steps 3 --warmup 3 > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:transform_cull_compact -s 3 -c 1 -o gpurun_out/r02_frame_16m5view -f python bench.py --no-also --no-cpu --steps 3 --warmup 3 > gpurun_out/ncu_frame.log 2>&1
tail -2 gpurun_out/ncu_frame.log | cut -c1-200
