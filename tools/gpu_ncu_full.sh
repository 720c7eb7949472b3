#!/bin/bash
# one ncu --set full capture of one kernel of one workload: tools/gpu_ncu_full.sh <workload> <kernel regex> <out name>
mkdir -p gpurun_out
CMD="python bench.py --workload $1 --steps 3 --warmup 3 --no-cpu --no-graph --no-also"
$CMD > gpurun_out/plain_$3.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:$2 -s 4 -c 1 -o gpurun_out/prof_$3 $CMD > gpurun_out/ncu_$3.log 2>&1
echo "ncu rc=$?"
