#!/bin/bash
mkdir -p gpurun_out
CMD="python bench.py --workload nuscenes --steps 3 --warmup 3 --no-cpu --no-graph"
$CMD > gpurun_out/plain_nus.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_hard_fwd -s 4 -c 1 -o gpurun_out/prof_hard_nus $CMD > gpurun_out/ncu_nus.log 2>&1
echo "ncu rc=$?"
