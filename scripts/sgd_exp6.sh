#!/bin/bash
timeout 600 python -m pytest tests/test_mf_gpu.py -x -q 2>&1 | tail -2
timeout 300 python bench.py --workload sgd --steps 3 --warmup 2 2>/dev/null | tail -1 | cut -c1-250
echo "N=8 one-slice panel"; python scripts/bench_sgd_kernel.py 1560000 60000 2212 128 37,74,148 1 2>&1 | tail -3 | cut -c1-130
