#!/bin/bash
# compute-sanitizer over the voxelization kernels: memcheck, racecheck, initcheck and
# synccheck on the small parity cases, plus an opcode histogram per kernel of the shipped library.  One gpurun call.
mkdir -p gpurun_out
OUT=gpurun_out/r02_sanitizer.txt
: > $OUT
SEL_HARD='single and kitti and lidar or edge or ragged and kitti'
SEL_DYN='single and kitti and lidar and True'
for tool in memcheck racecheck initcheck synccheck; do
  for spec in "tests/test_gpu_hard.py::$SEL_HARD" "tests/test_gpu_dynamic.py::$SEL_DYN" "tests/test_zz_no_alloc.py::twice"; do
    f=${spec%%::*}; k=${spec##*::}
    echo "=== compute-sanitizer --tool $tool  pytest $f -k \"$k\"" >> $OUT
    timeout 900 compute-sanitizer --tool $tool --error-exitcode 9 --print-limit 20 \
      python -m pytest $f -x -q -k "$k" > gpurun_out/san_$tool.log 2>&1
    echo "exit code $?" >> $OUT
    grep -E "passed|failed|ERROR SUMMARY|RACECHECK SUMMARY|=========     at|Race reported|hazard|Uninitialized|Invalid|Barrier error" gpurun_out/san_$tool.log | sort | uniq -c | sort -rn | head -n 25 >> $OUT
  done
done
tail -n 60 $OUT
