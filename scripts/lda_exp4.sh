#!/bin/bash
timeout 600 python -m pytest tests/test_lda_gpu.py -x -q 2>&1 | tail -2
for dk in int32 int16; do
HB_LDA_DK=$dk timeout 300 python bench.py --workload lda --steps 3 --warmup 2 2>/dev/null | tail -1 | cut -c1-220
done
echo "N=8-like panels"; python scripts/bench_lda_kernel.py 125000 12500 1000 12 37,74,148 1 2>&1 | grep "^{" | cut -c1-120
