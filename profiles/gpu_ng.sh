#!/bin/bash
# N = $NG GPUs of one box (default 2): the record merge through bench.py under torchrun, one summary line per
# variant -- copy-engine records (ce), stores from the frame kernel (push), then N = 1 on the same box.
# VARIANTS selects ("ce push" by default); every run is under its own short timeout.
NG=${NG:-2}
VARIANTS=${VARIANTS:-"ce push"}
STEPS=${STEPS:-30}
mkdir -p gpurun_out
run() { tag=$1; shift; timeout ${TMO:-240} python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 300)) bench.py --gpus $NG --steps $STEPS --warmup 5 "$@" > gpurun_out/n${NG}_$tag.log 2>&1; echo "== $tag rc=$?"; tail -1 gpurun_out/n${NG}_$tag.log | python -c "
import sys, json
try:
    l = json.loads(sys.stdin.read())
    m = l['merge']
    print(' ms/step', round(l['ms_per_step'], 4), 'value', round(l['value'] / 1e9, 2), 'G/s frame', round(l['roofline']['ms_per_launch'], 4), 'expand', round(m.get('expand_ms_per_step', m.get('ms_per_step', 0)), 4), 'e2e ms', round(l['e2e']['ms_per_step'], 2), 'send', m.get('send_ms_per_step'), 'verified', m['verified_against_oracle_per_rank'], m['verified_against_nccl_allgatherv'], m['transport'])
except Exception as e:
    print(' no line:', e)
"; }
if [ -n "$PYTEST" ]; then timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "record_merge" 2>&1 | tail -6; fi
for v in $VARIANTS; do
  case $v in
    ce) run ce --records ce ;;
    ce_streams) RCULL_CE_STREAMS=1 run ce_streams --records ce --no-verify ;;
    ce_early) run ce_early --records ce --send-early --no-verify ;;
    ce_bound) run ce_bound --records ce --send-bound --no-verify ;;
    ce_noverify) run ce_noverify --records ce --no-verify ;;
    push) run push --records push ;;
    default) run default ;;
    push_pipeline) run push_pipeline --records push --pipeline --no-verify ;;
    ipc_ce) run ipc_ce --records ce --transport ipc --no-verify ;;
    ipc) run ipc --transport ipc --no-verify ;;
    nccl) run nccl --merge nccl ;;
  esac
done
[ -n "$SKIP_N1" ] && exit 0
timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu --no-also > gpurun_out/n${NG}_n1.log 2>&1; tail -1 gpurun_out/n${NG}_n1.log | python -c "
import sys, json
l = json.loads(sys.stdin.read()); print('== N=1 on this box: ms/step', round(l['ms_per_step'], 4))"
grep -h -i "error\|Traceback" gpurun_out/n${NG}_*.log | head -5
