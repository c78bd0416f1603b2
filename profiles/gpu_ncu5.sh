#!/bin/bash
# one full ncu capture of the frame kernel (16m_5view unless WORKLOAD is set) after a plain run exited 0
W=${WORKLOAD:-16m_5view}; TAG=${TAG:-frame_$W}
timeout 300 python bench.py --steps 5 --warmup 3 --workload $W > gpurun_out/pre_$TAG.log 2>&1 || { tail -5 gpurun_out/pre_$TAG.log; exit 1; }
timeout 900 ncu --set full --clock-control none --import-source on -k regex:transform_cull_compact -s 3 -c 1 -o gpurun_out/$TAG -f python bench.py --steps 3 --warmup 3 --workload $W > gpurun_out/ncu_$TAG.log 2>&1
tail -2 gpurun_out/ncu_$TAG.log | cut -c1-300
