#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -n 30 gpurun_out/pytest_gpu.log
timeout 300 python bench.py --workload nuscenes-dynamic > gpurun_out/bench_nus_dyn.json 2> gpurun_out/bench_nus_dyn.err; echo "bench rc=$?"
cat gpurun_out/bench_nus_dyn.json; tail -n 5 gpurun_out/bench_nus_dyn.err
