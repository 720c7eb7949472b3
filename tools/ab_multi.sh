#!/bin/bash
# A/B of several builds of the library on the same box: tools/ab_multi.sh "<variant.so> ..." [workloads...]
# (each variant is also run through the hard-voxelization parity tests first)
cd /root/repo; mkdir -p gpurun_out
VS=$1; shift
WL=${@:-kitti nuscenes}
for lib in $VS; do
  RPC_B200_LIB=/root/repo/$lib timeout 300 python -m pytest tests/test_gpu_hard.py tests/test_gpu_host.py -x -q 2>&1 | tail -n 1 | sed "s|^|$lib parity: |"
done 2>&1 | tee gpurun_out/ab.log
for wl in $WL; do
  for rep in 1 2; do
    for lib in default $VS; do
      if [ $lib = default ]; then unset RPC_B200_LIB; else export RPC_B200_LIB=/root/repo/$lib; fi
      timeout 300 python bench.py --workload $wl --steps 200 --warmup 20 --no-cpu 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1]); r = d['roofline']
print('$wl', '$lib', 'us/step', round(d['ms_per_step'] * 1e3, 2), r['stage_us'], r['fwd_phase_us'])"
    done
  done
done 2>&1 | tee -a gpurun_out/ab.log
