#!/bin/bash
# 8-GPU box helper: the default bench line plus exchange-CTA variants (one summary line each).
mkdir -p gpurun_out
run() {  # name, env assignments..., then bench args after --
  name=$1; shift
  envs=(); while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  env "${envs[@]}" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 200)) \
      bench.py --gpus 8 --steps 20 --warmup 5 "$@" > gpurun_out/n8_$name.log 2>&1
  python - "$name" <<'PY'
import json, sys
txt = open(f"gpurun_out/n8_{sys.argv[1]}.log").read()
try:
    d = json.loads([l for l in txt.splitlines() if l.startswith("{")][-1])
    m = d.get("merge") or {}
    print(sys.argv[1], "ms/step %.4f" % d["ms_per_step"], "frame %.4f" % d["roofline"]["ms_per_launch"], "value %.4g" % d["value"],
          {k: (round(v, 4) if isinstance(v, float) else v) for k, v in m.items() if k in ("exposed_ms_per_step", "ms_per_step", "transport", "verified_against_nccl_allgatherv")})
except Exception:
    print(sys.argv[1], "FAILED", txt[-1200:])
PY
}
run default RCULL_X=0 --
run ctas64 RCULL_EXCHANGE_CTAS=64 --
run ctas16 RCULL_EXCHANGE_CTAS=16 --
run serial RCULL_X=0 -- --no-overlap
