#!/bin/bash
# N = $NG (default 2): record merge through bench.py under torchrun -- IPC peer stores (default), split-chip expand,
# multicast, NCCL fallback; then the N = 1 line on the same box for the efficiency
NG=${NG:-2}
mkdir -p gpurun_out
run() { tag=$1; shift; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $NG --steps 20 --warmup 5 "$@" > gpurun_out/n${NG}_$tag.log 2>&1; echo "== $tag rc=$?"; tail -1 gpurun_out/n${NG}_$tag.log | python -c "
import sys, json
try:
    l = json.loads(sys.stdin.read())
    m = l['merge']
    print(' ms/step', round(l['ms_per_step'], 4), 'value', round(l['value'] / 1e9, 2), 'G/s frame', round(l['roofline']['ms_per_launch'], 4), 'expand', round(m.get('expand_ms_per_step', m.get('ms_per_step', 0)), 4), 'e2e ms', round(l['e2e']['ms_per_step'], 2), 'nvlink in GB/s', round(m['nvlink_inbound_gbs_per_gpu'], 1), 'verified', m['verified_against_oracle_per_rank'], m['verified_against_nccl_allgatherv'], m['transport'])
except Exception as e:
    print(' no line:', e)
"; }
run auto
[ -n "$QUICK" ] || run ipc_pipeline --pipeline --no-verify
[ -n "$QUICK" ] || run split --expand-sms ${SPLIT:-8} --no-verify
[ -n "$QUICK" ] || run ipc --transport ipc --no-verify
[ -n "$QUICK" ] || run nccl --merge nccl
timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu --no-also > gpurun_out/n${NG}_n1.log 2>&1; tail -1 gpurun_out/n${NG}_n1.log | python -c "
import sys, json
l = json.loads(sys.stdin.read()); print('== N=1 on this box: ms/step', round(l['ms_per_step'], 4))"
grep -h -i "error\|Traceback" gpurun_out/n${NG}_*.log | head -5
