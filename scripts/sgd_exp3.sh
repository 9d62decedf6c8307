#!/bin/bash
timeout 300 python -m pytest tests/test_mf_gpu.py tests/test_lda_gpu.py -x -q 2>&1 | tail -3
echo "full slice (N=1)"; python scripts/bench_sgd_kernel.py 50000000 480000 8850 128 148,296 2>&1 | tail -2 | cut -c1-150
echo "N=8 two-slice panel"; python scripts/bench_sgd_kernel.py 780000 60000 1106 128 37,74,148 2>&1 | tail -3 | cut -c1-150
echo "N=8 one-slice panel"; python scripts/bench_sgd_kernel.py 1560000 60000 2212 128 37,74,148 2>&1 | tail -3 | cut -c1-150
timeout 300 python bench.py --workload sgd --steps 3 --warmup 2 2>/dev/null | tail -1 | cut -c1-250
timeout 300 python bench.py --workload lda --steps 3 --warmup 2 2>/dev/null | tail -1 | cut -c1-250
