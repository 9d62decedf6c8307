#!/bin/bash
for v in 512_pf1 512_pf3 1024_pf1 1024_pf3 768_pf2 512_pf1 1024_pf3; do
  echo -n "$v: "; HB_CUDA_LIB=$PWD/ab/lib_$v.so timeout 300 python bench.py --workload sgd --steps 5 --warmup 2 2>/dev/null | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['value']/1e9)"
done
