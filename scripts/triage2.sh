#!/bin/bash
for dbg in "$@"; do
  HB_KMEANS_DBG=$dbg timeout 120 python scripts/bench_kmeans_kernel.py 2000000 128 10000 3 pair > /tmp/o.json 2>/dev/null
  python - "$dbg" <<'PY'
import json, sys
d = json.loads(open('/tmp/o.json').read().strip().splitlines()[-1])
print(f"dbg={sys.argv[1]} ms={d['ms_median']:.3f} TF={d['tflops_tf32']:.1f}")
PY
done
