#!/bin/bash
# BASELINE.json configs[4]: throughput sweep 16k-2M points per sweep (batch 2 per GPU), KITTI and nuScenes
# parameter sets, full step (perturb + hard voxelize + VFE fwd/bwd), with the 1-core CPU port beside it.
mkdir -p gpurun_out
: > gpurun_out/sweep.jsonl
for wl in kitti nuscenes; do
  for n in 16000 32000 64000 128000 256000 512000 1000000 2000000; do
    timeout 600 python bench.py --workload $wl --points $n --batch 2 --steps 50 --warmup 5 2>>gpurun_out/sweep.err | tail -n 1 >> gpurun_out/sweep.jsonl
  done
done
python - <<'PY'
import json
print(f"{'workload':9s} {'N/sweep':>8s} {'B':>2s} {'voxels':>7s} {'kept':>8s} {'us/step':>8s} {'Mpts/s':>9s} {'step %HBM':>9s} {'cpu1 Mpts/s':>11s}")
for l in open('gpurun_out/sweep.jsonl'):
    if not l.startswith('{'): continue
    d = json.loads(l); c = d['config']
    print(f"{'kitti' if 'KITTI' in c['workload'] else 'nuscenes':9s} {c['points_per_sweep']:8d} {c['batch_per_gpu']:2d} {c['voxels']:7d} {c['kept_points']:8d} "
          f"{d['ms_per_step']*1e3:8.1f} {d['value']:9.1f} {100*d['roofline']['step']['frac']:8.2f}% {d.get('cpu_baseline',{}).get('value',0):11.3f}")
PY
