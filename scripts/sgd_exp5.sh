#!/bin/bash
timeout 600 python -m pytest tests/test_mf_gpu.py tests/test_lda_gpu.py -x -q 2>&1 | tail -2
timeout 300 python bench.py --workload sgd --steps 3 --warmup 2 2>/dev/null | tail -1 | cut -c1-250
timeout 300 python bench.py --workload lda --steps 3 --warmup 2 2>/dev/null | tail -1 | cut -c1-250
