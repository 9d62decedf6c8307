#!/bin/bash
# usage: ncu_one.sh <name> <kernel-regex> <python args...>   — plain run first, ncu only if it exits 0
name=$1; regex=$2; shift 2
mkdir -p gpurun_out
timeout 400 python "$@" > gpurun_out/plain_$name.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:$regex -c 1 -f -o gpurun_out/$name \
  python "$@" > gpurun_out/ncu_$name.log 2>&1
echo "rc=$?"; tail -1 gpurun_out/plain_$name.log | cut -c1-600
