#!/bin/bash
# N-GPU validation ($1 = N): NVLink collectives + fused kernels, engines, then the four benches at N
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 400 $TR --master-port 29533 tests/mp_symm_worker.py auto > gpurun_out/symm$N.log 2>&1; echo "symm$N rc=$?"; grep -E "backend=|allreduce 5|fused|SYMM_OK|Error|error" gpurun_out/symm$N.log | head
timeout 400 $TR --master-port 29534 tests/mp_engines_worker.py > gpurun_out/engines$N.log 2>&1; echo "engines$N rc=$?"; grep -E "ENGINES_OK|Error" gpurun_out/engines$N.log | head -3
for w in kmeans sgd lda gram; do
  steps=5; [ $w = kmeans ] && steps=10
  timeout 400 $TR --master-port 29551 bench.py --gpus $N --steps $steps --warmup 3 --workload $w 2>gpurun_out/scale_${w}_$N.err | grep "^{" > gpurun_out/scale_${w}_$N.json
  python - $w $N <<'PY'
import json, sys
w, n = sys.argv[1], sys.argv[2]
try:
    d = json.loads(open(f"gpurun_out/scale_{w}_{n}.json").read().strip().splitlines()[-1])
    print(w, f"N={n}", d["metric"], d["value"], "ms/step", round(d["ms_per_step"], 3), d.get("config", {}).get("comm"), d.get("clocks"))
except Exception as e:
    print(w, "FAILED", repr(e)); print(open(f"gpurun_out/scale_{w}_{n}.err").read()[-1500:])
PY
done
