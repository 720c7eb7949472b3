#!/bin/bash
cd /root/repo; mkdir -p gpurun_out
V=$1
for n in 16000 64000 256000; do for rep in 1 2; do for lib in default $V; do
  if [ $lib = default ]; then unset RPC_B200_LIB; else export RPC_B200_LIB=/root/repo/$lib; fi
  timeout 300 python bench.py --workload kitti --points $n --batch 2 --steps 100 --warmup 10 --no-cpu 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1]); print($n, '$lib', 'us/step', round(d['ms_per_step'] * 1e3, 2))"
done; done; done 2>&1 | tee gpurun_out/ab_small.log
