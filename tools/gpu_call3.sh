#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_dynamic.py -m gpu -x -q 2>&1 | tail -n 15
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
