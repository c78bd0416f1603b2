#!/bin/bash
# GPU-box helper (round end): smoke, parity tests, bench lines, ncu launch list + full captures.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.max.mem --format=csv > gpurun_out/smi.txt; nproc >> gpurun_out/smi.txt
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -2
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 | tee gpurun_out/pytest_gpu.txt
for w in 16m_5view 16m_1view 1m_1view 1m_tri 1m_tri_random; do
  timeout 400 python bench.py --steps 30 --warmup 5 --workload $w > gpurun_out/bench_$w.log 2>&1
  tail -1 gpurun_out/bench_$w.log | cut -c1-300
done
timeout 200 python profiles/per_mesh_bench.py > gpurun_out/per_mesh_1view.log 2>&1; tail -1 gpurun_out/per_mesh_1view.log | cut -c1-300
timeout 200 python profiles/per_mesh_bench.py --views 5 > gpurun_out/per_mesh_5view.log 2>&1; tail -1 gpurun_out/per_mesh_5view.log | cut -c1-300
timeout 400 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.log 2>&1; tail -1 gpurun_out/bench_ref.log | cut -c1-300
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_16m5view.csv python bench.py --no-also --steps 3 --warmup 3 > gpurun_out/ncu_launch.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_1mtri.csv python bench.py --no-also --steps 3 --warmup 3 --workload 1m_tri > gpurun_out/ncu_launch_tri.log 2>&1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_per_mesh.csv python profiles/per_mesh_bench.py --steps 3 --warmup 3 > gpurun_out/ncu_launch_per_mesh.log 2>&1
for spec in "16m_5view:transform_cull_compact:frame_16m5view" "16m_1view:transform_cull_compact:frame_16m1view" "1m_tri:tri_cull:tri_1m"; do
  IFS=: read w k tag <<< "$spec"
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:$k -s 3 -c 1 -o gpurun_out/$tag -f python bench.py --no-also --steps 3 --warmup 3 --workload $w > gpurun_out/ncu_$tag.log 2>&1
  tail -1 gpurun_out/ncu_$tag.log | cut -c1-120
done
