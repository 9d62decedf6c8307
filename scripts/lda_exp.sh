#!/bin/bash
timeout 300 python -m pytest tests/test_lda_gpu.py tests/test_mf_gpu.py -x -q 2>&1 | tail -2
echo "N=8-like panels (125K docs, 1.5M tokens, vocab 12500)"
python scripts/bench_lda_kernel.py 125000 12500 1000 12 16,37,74,148 1,2 2>&1 | grep "^{" | cut -c1-120
echo "N=1 full (1M docs x 100, vocab 100K)"
python scripts/bench_lda_kernel.py 1000000 100000 1000 100 74,148 2 2>&1 | grep "^{" | cut -c1-120
