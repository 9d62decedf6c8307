#!/bin/bash
# full bench line, then one ncu capture of the flagship kernel (only after the plain run exited 0)
mkdir -p gpurun_out
timeout 400 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err
tail -1 gpurun_out/bench_n1.json
timeout 200 python scripts/bench_kmeans_kernel.py 2000000 128 10000 3 pair > gpurun_out/plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:assign_tc_pair -c 1 -f -o gpurun_out/kmeans_pair3 \
  python scripts/bench_kmeans_kernel.py 2000000 128 10000 1 pair > gpurun_out/ncu.log 2>&1
tail -2 gpurun_out/plain.log | cut -c1-300
