#!/bin/bash
# GPU-box helper (round 2): launch lists of the bench commands + one full ncu capture per dominant kernel.
# Every profiled command first runs plainly and must exit 0.
mkdir -p gpurun_out
for w in 16m_5view 1m_tri; do
  timeout 300 python bench.py --no-cpu --steps 3 --warmup 3 --workload $w > gpurun_out/pre_$w.log 2>&1 || { echo "plain run of $w failed"; tail -5 gpurun_out/pre_$w.log; exit 1; }
done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r02_launches_16m5view.csv python bench.py --no-cpu --steps 3 --warmup 3 > gpurun_out/ncu_launch.log 2>&1; echo "launch list 5view rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r02_launches_1mtri.csv python bench.py --no-cpu --no-also --steps 3 --warmup 3 --workload 1m_tri > gpurun_out/ncu_launch_tri.log 2>&1; echo "launch list tri rc=$?"
timeout 200 python profiles/per_mesh_bench.py --steps 3 --warmup 3 > gpurun_out/pre_per_mesh.log 2>&1 && \
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/r02_launches_per_mesh.csv python profiles/per_mesh_bench.py --steps 3 --warmup 3 > gpurun_out/ncu_launch_per_mesh.log 2>&1; echo "launch list per-mesh rc=$?"
for spec in "16m_5view:transform_cull_compact:3:frame_16m5view" "16m_1view:transform_cull_compact:3:frame_16m1view" "1m_tri:tri_cull:3:tri_cull_1m"; do
  IFS=: read w k skip tag <<< "$spec"
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:$k -s $skip -c 1 -o gpurun_out/r02_$tag -f python bench.py --no-cpu --no-also --steps 3 --warmup 3 --workload $w > gpurun_out/ncu_$tag.log 2>&1
  echo "$tag rc=$?"; tail -1 gpurun_out/ncu_$tag.log | cut -c1-120
done
ls -la gpurun_out
