#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_kmeans_gpu.py tests/test_gram_gpu.py tests/test_zz_no_alloc.py -x -q 2>&1 | tail -3
for v in "$@"; do
  HB_KMEANS_VARIANT=$v timeout 200 python bench.py --steps 10 --warmup 3 --no-e2e > gpurun_out/bench_$v.json 2>/dev/null
  python - "$v" <<'PY'
import json, sys
d = json.loads(open(f'gpurun_out/bench_{sys.argv[1]}.json').read().strip().splitlines()[-1])
print(sys.argv[1], d["ms_per_step"], d["tflops_tf32_aggregate"], d["clocks"])
PY
done
timeout 200 python bench.py --workload gram --steps 5 --warmup 3 2>/dev/null | tail -1
