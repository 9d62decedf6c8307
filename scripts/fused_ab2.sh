#!/bin/bash
run() { timeout 200 python bench.py --steps 20 --warmup 3 --no-e2e --comm $1 --points-per-gpu 2000000 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('2M', d['config']['comm'], d['ms_per_step'], d['clocks']['sm_mhz'])"; }
run nccl; run fused; HB_FUSED_NO_COOP=1 run fused; run nccl; HB_FUSED_NO_COOP=1 run fused; run fused
