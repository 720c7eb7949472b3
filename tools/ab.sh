#!/bin/bash
# A/B of two builds of the library on the same box: tools/ab.sh <variant.so> [workloads...]
cd /root/repo; mkdir -p gpurun_out
V=$1; shift
WL=${@:-kitti nuscenes}
for wl in $WL; do
  for rep in 1 2; do
    for lib in default $V; do
      if [ $lib = default ]; then unset RPC_B200_LIB; else export RPC_B200_LIB=/root/repo/$lib; fi
      timeout 300 python bench.py --workload $wl --steps 200 --warmup 20 --no-cpu 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1]); r = d['roofline']
print('$wl', '$lib', 'us/step', round(d['ms_per_step'] * 1e3, 2), r['stage_us'], r['fwd_phase_us'])"
    done
  done
done 2>&1 | tee gpurun_out/ab.log
