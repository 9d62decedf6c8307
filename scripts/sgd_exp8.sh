#!/bin/bash
timeout 600 python -m pytest tests/test_mf_gpu.py -x -q 2>&1 | tail -2
echo "N=8 one-slice panel"; python scripts/bench_sgd_kernel.py 1560000 60000 2212 128 37,56,74,110,148 1 2>&1 | tail -5 | cut -c1-130
echo "N=8 two-slice panel"; python scripts/bench_sgd_kernel.py 780000 60000 1106 128 37,56,74,110,148 1 2>&1 | tail -5 | cut -c1-130
echo "N=2 two-slice panel"; python scripts/bench_sgd_kernel.py 12500000 240000 4425 128 74,110,148 1 2>&1 | tail -3 | cut -c1-130
