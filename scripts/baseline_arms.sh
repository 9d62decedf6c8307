#!/bin/bash
mkdir -p gpurun_out
timeout 600 python bench.py --impl nccl_cublas --steps 3 --warmup 3 > gpurun_out/base_kmeans_1.json 2>gpurun_out/base_kmeans_1.err; tail -1 gpurun_out/base_kmeans_1.json | cut -c1-400; tail -3 gpurun_out/base_kmeans_1.err
timeout 600 python bench.py --impl nccl_cublas --workload gram --steps 5 --warmup 3 > gpurun_out/base_gram_1.json 2>gpurun_out/base_gram_1.err; tail -1 gpurun_out/base_gram_1.json | cut -c1-400; tail -3 gpurun_out/base_gram_1.err
bash scripts/ncu_one.sh gram_tc gram_tc_kernel bench.py --workload gram --steps 2 --warmup 3
