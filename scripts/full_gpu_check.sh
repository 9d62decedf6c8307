#!/bin/bash
# what the driver runs at round end on one GPU: the gpu test tier, smoke(), the headline bench
mkdir -p gpurun_out
timeout 900 python -m pytest tests/ -x -q -m gpu > gpurun_out/pytest_gpu_all.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu_all.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('SMOKE_OK')" 2>&1 | tail -2
timeout 300 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_n1.json 2>gpurun_out/bench_n1.err; tail -1 gpurun_out/bench_n1.json | cut -c1-1200
