#!/bin/bash
N=2
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 400 python -m pytest tests/test_symm_gpu.py -x -q 2>&1 | tail -2
for w in sgd lda gram; do
  timeout 300 $TR --master-port 29551 bench.py --gpus $N --steps 5 --warmup 3 --workload $w 2>gpurun_out/scale_${w}_$N.err | grep "^{" > gpurun_out/scale_${w}_$N.json
  python -c "
import json
d=json.loads(open('gpurun_out/scale_${w}_$N.json').read().strip().splitlines()[-1]); print('$w', d['value'], d['ms_per_step'])"
done
