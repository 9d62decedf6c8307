#!/bin/bash
timeout 300 python -m pytest tests/test_mf_gpu.py -x -q 2>&1 | tail -3
echo "full slice (N=1)"; python scripts/bench_sgd_kernel.py 50000000 480000 8850 128 296 1,1.5,2,3 2>&1 | tail -4 | cut -c1-130
echo "N=8 two-slice panel"; python scripts/bench_sgd_kernel.py 780000 60000 1106 128 37,74,148 1,2 2>&1 | tail -6 | cut -c1-130
echo "N=8 one-slice panel"; python scripts/bench_sgd_kernel.py 1560000 60000 2212 128 37,74,148 1,2 2>&1 | tail -6 | cut -c1-130
