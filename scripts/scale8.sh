#!/bin/bash
# 8-GPU validation: NVLink collectives at world=8, K-means scaling 1/2/4/8, secondary workloads at 8
mkdir -p gpurun_out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 tests/mp_symm_worker.py auto > gpurun_out/symm8.log 2>&1; echo "symm8 rc=$?"; grep -E "backend=|allreduce 5|SYMM_OK|Error" gpurun_out/symm8.log | head
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29534 tests/mp_engines_worker.py > gpurun_out/engines8.log 2>&1; echo "engines8 rc=$?"; grep -E "ENGINES_OK|Error" gpurun_out/engines8.log | head -3
for n in 1 2 4 8; do
  if [ $n -eq 1 ]; then timeout 300 python bench.py --gpus 1 --steps 10 --warmup 3 > gpurun_out/scale_kmeans_$n.json 2>gpurun_out/scale_kmeans_$n.err
  else timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2954$n bench.py --gpus $n --steps 10 --warmup 3 2>gpurun_out/scale_kmeans_$n.err | grep "^{" > gpurun_out/scale_kmeans_$n.json; fi
  python - $n <<'PY'
import json, sys
n = sys.argv[1]
try:
    d = json.loads(open(f"gpurun_out/scale_kmeans_{n}.json").read().strip().splitlines()[-1])
    print("kmeans N=%s ms/step %.2f e2e %.2f comm %s launches %s clocks %s" % (n, d["ms_per_step"], d["e2e"]["value"] * 1e3, d["config"]["comm"], d["gpu_launches"], d["clocks"]))
except Exception as e:
    print("kmeans N=%s FAILED %r" % (n, e)); print(open(f"gpurun_out/scale_kmeans_{n}.err").read()[-1500:])
PY
done
for w in sgd lda gram; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29551 bench.py --gpus 8 --steps 3 --warmup 1 --workload $w 2>gpurun_out/scale_${w}_8.err | grep "^{" > gpurun_out/scale_${w}_8.json
  python - $w <<'PY'
import json, sys
w = sys.argv[1]
try:
    d = json.loads(open(f"gpurun_out/scale_{w}_8.json").read().strip().splitlines()[-1])
    print(w, "N=8", d["metric"], d["value"], "ms/step", d["ms_per_step"])
except Exception as e:
    print(w, "FAILED", repr(e)); print(open(f"gpurun_out/scale_{w}_8.err").read()[-1500:])
PY
done
