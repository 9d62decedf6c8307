#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/ -x -q -m gpu > gpurun_out/pytest_gpu_all.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu_all.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('SMOKE_OK')" 2>&1 | tail -2
bash scripts/ncu_one.sh lda_pass lda_pass_kernel bench.py --workload lda --steps 1 --warmup 2
