#!/bin/bash
# A/B of several builds on one box: tools/ab_multi.sh "<lib1> <lib2> ..." [workloads...]   ("default" = in-tree build)
cd /root/repo; mkdir -p gpurun_out
LIBS=$1; shift
WL=${@:-kitti nuscenes}
for wl in $WL; do
  for rep in 1 2; do
    for lib in $LIBS; do
      if [ $lib = default ]; then unset RPC_B200_LIB; else export RPC_B200_LIB=/root/repo/$lib; fi
      timeout 300 python bench.py --workload $wl --steps 200 --warmup 20 --no-cpu 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1]); r = d['roofline']
print('$wl', '$lib', 'us/step', round(d['ms_per_step'] * 1e3, 2), r['stage_us'], r['fwd_phase_us'])"
    done
  done
done 2>&1 | tee gpurun_out/ab.log
