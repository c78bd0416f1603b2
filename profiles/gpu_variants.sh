#!/bin/bash
# A/B of frame-kernel shapes: every librcull_*.so under renderer_b200/variants (same ABI, RCULL_LIB selects it):
# a parity subset, then the resident 16 M timings
mkdir -p gpurun_out
for lib in renderer_b200/librcull.so renderer_b200/variants/librcull_*.so; do
  name=$(basename $lib .so)
  export RCULL_LIB=$PWD/$lib
  t=$(timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "config1 or ragged or every_view or all_visible or none_visible or extreme or packed_sum or wide_sums or sponza" 2>&1 | tail -1)
  for w in 16m_5view 16m_1view 1m_1view; do
    timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu --no-also --workload $w > gpurun_out/bench_${name}_$w.log 2>&1
  done
  python - "$name" "$t" <<'PY'
import json, sys
name, t = sys.argv[1], sys.argv[2]
out = []
for w in ("16m_5view", "16m_1view", "1m_1view"):
    try:
        l = json.loads(open(f"gpurun_out/bench_{name}_{w}.log").read().strip().splitlines()[-1])
        out.append(f"{w} {l['roofline']['ms_per_launch']:.4f} ms frac {l['roofline']['frac']:.3f}")
    except Exception as e:
        out.append(f"{w} FAILED")
print(f"{name:24s} | {' | '.join(out)} | tests: {t}")
PY
done
