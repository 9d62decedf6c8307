#!/bin/bash
mkdir -p gpurun_out
for c in nccl fused nccl fused; do
  timeout 200 python bench.py --steps 10 --warmup 3 --no-e2e --comm $c > gpurun_out/bench_ab_$c.json 2>/dev/null
  python - "$c" <<'PY'
import json, sys
d = json.loads(open(f'gpurun_out/bench_ab_{sys.argv[1]}.json').read().strip().splitlines()[-1])
print(sys.argv[1], d["ms_per_step"], d["tflops_tf32_aggregate"], d["clocks"], d["gpu_launches"])
PY
done
for c in nccl fused; do
  timeout 200 python bench.py --steps 20 --warmup 3 --no-e2e --comm $c --points-per-gpu 2000000 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('2M', d['config']['comm'], d['ms_per_step'], d['clocks'])"
done
