#!/bin/bash
# correctness first, then timing of the pair kernel under a few debug masks, then the profile printf
timeout 600 python -m pytest tests/test_kmeans_gpu.py tests/test_zz_no_alloc.py -x -q 2>&1 | tail -3
bash scripts/triage2.sh "$@"
HB_KMEANS_DBG=16 timeout 120 python scripts/bench_kmeans_kernel.py 2000000 128 10000 1 pair 2>&1 | grep "^prof" | sort | uniq | head -24
