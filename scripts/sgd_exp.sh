#!/bin/bash
echo "full slice (N=1)"; python scripts/bench_sgd_kernel.py 50000000 480000 8850 128 148,296 2>&1 | tail -2
echo "quarter rows"; python scripts/bench_sgd_kernel.py 12500000 120000 8850 128 148,296 2>&1 | tail -2
echo "half rows"; python scripts/bench_sgd_kernel.py 25000000 240000 8850 128 148,296 2>&1 | tail -2
echo "N=8 two-slice panel"; python scripts/bench_sgd_kernel.py 780000 60000 1106 128 37,74,148,296 2>&1 | tail -4
echo "N=8 one-slice panel"; python scripts/bench_sgd_kernel.py 1560000 60000 2212 128 37,74,148,296 2>&1 | tail -4
