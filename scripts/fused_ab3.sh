#!/bin/bash
HB_FUSED_DBG=16 timeout 200 python bench.py --steps 1 --warmup 3 --no-e2e --comm fused --points-per-gpu 2000000 2>&1 | grep "^prof" | sort | uniq | awk 'NR%3==1' | head -14
HB_KMEANS_DBG=16 timeout 200 python bench.py --steps 1 --warmup 3 --no-e2e --comm nccl --points-per-gpu 2000000 2>&1 | grep "^prof" | sort | uniq | awk 'NR%3==1' | head -14
