#!/bin/bash
run() { timeout 200 python bench.py --steps 20 --warmup 3 --no-e2e --comm $1 --points-per-gpu $2 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$2', d['config']['comm'], d['ms_per_step'], d['clocks']['sm_mhz'])"; }
HB_KMEANS_DBG=4 run nccl 2000000; HB_FUSED_DBG=4 run fused 2000000; run nccl 2000000; run fused 2000000
run nccl 200000; run fused 200000; run nccl 20000; run fused 20000
