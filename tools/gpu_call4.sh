#!/bin/bash
mkdir -p gpurun_out
for wl in kitti nuscenes nuscenes-dynamic; do
  for fl in write write+read; do
    timeout 300 python bench.py --workload $wl --flush $fl --steps 50 --no-cpu 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('$wl','$fl',d['ms_per_step'], d['roofline']['stage_us'], d['roofline']['fwd_phase_us'], 'e2e', d['e2e']['ms_per_step'])
    else: print(l.rstrip())
"
  done
done
CMD="python bench.py --workload nuscenes-dynamic --steps 3 --warmup 3 --no-cpu --no-graph"
$CMD > gpurun_out/plain_dyn.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_dyn_scatter -s 4 -c 1 -o gpurun_out/prof_dyn $CMD > gpurun_out/ncu_dyn.log 2>&1
echo "ncu rc=$?"
