#!/bin/bash
for nullf in 0.005 0; do
for wl in kitti nuscenes nuscenes-dynamic; do
    RPC_SYNTH_NULL=$nullf timeout 300 python bench.py --workload $wl --steps 50 --no-cpu 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('$wl','null=$nullf',d['ms_per_step'], d['roofline']['stage_us'], d['roofline']['fwd_phase_us'], d['config']['voxels'], d['config']['kept_points'])
    else: print(l.rstrip())
"
done
done
