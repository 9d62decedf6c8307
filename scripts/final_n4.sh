#!/bin/bash
N=${1:-4}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 400 $TR --master-port 29551 bench.py --gpus $N --steps 10 --warmup 3 2>gpurun_out/scale_kmeans_$N.err | grep "^{" > gpurun_out/scale_kmeans_$N.json
python - $N <<'PY'
import json, sys
n = sys.argv[1]
d = json.loads(open(f"gpurun_out/scale_kmeans_{n}.json").read().strip().splitlines()[-1])
print("kmeans", f"N={n}", d["value"], "e2e", d["e2e"]["value"], d["config"]["comm"], d["clocks"])
PY
# README-sized K-means (1000 points over the ranks, 10 centroids, dim 100): launch-latency regime
timeout 300 $TR --master-port 29552 bench.py --gpus $N --steps 200 --warmup 20 --points-per-gpu $((1000 / N)) --dim 100 --k 10 --no-e2e 2>/dev/null | grep "^{" | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('readme-config', d['n_gpus'], 's/iter', d['value'])"
