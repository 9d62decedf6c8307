#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_kmeans_gpu.py tests/test_zz_no_alloc.py -x -q 2>&1 | tail -5
for v in auto; do
  timeout 200 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_fused.json 2>gpurun_out/bench_fused.err
  tail -1 gpurun_out/bench_fused.json | cut -c1-1500
done
