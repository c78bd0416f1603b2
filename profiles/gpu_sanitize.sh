#!/bin/bash
# one compute-sanitizer tool per gpurun call (B200_PROFILING.md): TOOL=racecheck|memcheck|initcheck|synccheck
# on BASELINE configs[0] (10 k entities, 5 views), the triangle stage on a small scene, the per-mesh mode and the
# two-rank record merge (ranks emulated in one process)
TOOL=${TOOL:-racecheck}
mkdir -p gpurun_out
SEL="config1_10k_five_views or test_per_mesh_instanced_mode_synthetic or record_merge_ranks_in_one_process or frozen_frame_on_the_other or packed_sum"
timeout 240 python -m pytest tests/test_gpu_parity.py tests/test_gpu_tri.py -m gpu -x -q -k "$SEL or tri_cull_small or small" > gpurun_out/sanitize_plain.log 2>&1 || { tail -5 gpurun_out/sanitize_plain.log; echo "plain run failed"; exit 1; }
tail -1 gpurun_out/sanitize_plain.log
timeout 1500 compute-sanitizer --tool $TOOL --print-limit 20 --log-file gpurun_out/r02_sanitizer_$TOOL.txt python -m pytest tests/test_gpu_parity.py tests/test_gpu_tri.py -m gpu -x -q -k "$SEL or tri_cull_small or small" 2>&1 | tail -3
echo "== $TOOL log:"; tail -8 gpurun_out/r02_sanitizer_$TOOL.txt
