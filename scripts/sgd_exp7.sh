#!/bin/bash
echo "full"; python scripts/bench_sgd_kernel.py 50000000 480000 8850 128 148 1 2>&1 | tail -1 | cut -c1-130
echo "half rows"; python scripts/bench_sgd_kernel.py 25000000 240000 8850 128 148 1 2>&1 | tail -1 | cut -c1-130
echo "third rows"; python scripts/bench_sgd_kernel.py 16666666 160000 8850 128 148 1 2>&1 | tail -1 | cut -c1-130
echo "quarter rows"; python scripts/bench_sgd_kernel.py 12500000 120000 8850 128 148 1 2>&1 | tail -1 | cut -c1-130
