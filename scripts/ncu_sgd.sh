#!/bin/bash
bash scripts/ncu_one.sh sgd_pass sgd_pass_kernel bench.py --workload sgd --steps 2 --warmup 3
