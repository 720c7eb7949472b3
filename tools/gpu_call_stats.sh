#!/bin/bash
cd /root/repo
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_perturber_stats.py tests/test_gpu_host.py -x -q > gpurun_out/stats_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/stats_tests.log
tail -15 gpurun_out/stats_tests.log
PYTHONPATH=/root/repo timeout 300 python tools/bench_perturber_stats.py > gpurun_out/stats_bench.log 2>&1
cat gpurun_out/stats_bench.log
