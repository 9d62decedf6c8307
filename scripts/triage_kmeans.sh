#!/bin/bash
# perf triage: per-SM throughput of the pair kernel vs number of active CTA pairs
mkdir -p gpurun_out
for mc in 74 37 18; do
  for dbg in 0 7; do
    HB_KMEANS_MAX_CLUSTERS=$mc HB_KMEANS_DBG=$dbg timeout 120 python scripts/bench_kmeans_kernel.py 1000000 128 10000 3 pair > /tmp/o.json 2>/dev/null
    python - "$mc" "$dbg" <<'PY'
import json, sys
d = json.loads(open('/tmp/o.json').read().strip().splitlines()[-1])
mc = int(sys.argv[1])
print(f"clusters={mc} dbg={sys.argv[2]} ms={d['ms_median']:.3f} TF={d['tflops_tf32']:.1f} TF_scaled_to_74={d['tflops_tf32']*74/mc:.1f}")
PY
  done
done
