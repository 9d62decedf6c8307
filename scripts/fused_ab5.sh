#!/bin/bash
timeout 300 python -m pytest tests/test_kmeans_gpu.py tests/test_zz_no_alloc.py -x -q 2>&1 | tail -2
run() { timeout 200 python bench.py --steps 20 --warmup 3 --no-e2e --comm $1 --points-per-gpu $2 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$2', d['config']['comm'], d['ms_per_step'], d['clocks']['sm_mhz'])"; }
run nccl 2000000; run fused 2000000; run nccl 200000; run fused 200000; run nccl 20000; run fused 20000
run nccl 12500000; run fused 12500000
