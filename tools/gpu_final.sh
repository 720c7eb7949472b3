#!/bin/bash
# round-end evidence run: GPU parity tests, smoke, all bench workloads, reference arm, ncu launch list
# (the ncu --set full captures are separate calls: tools/gpu_ncu_full.sh <workload> <kernel regex> <name>)
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/smoke.log
for wl in kitti nuscenes nuscenes-dynamic; do
  python bench.py --workload $wl > gpurun_out/bench_$wl.json 2> gpurun_out/bench_$wl.err; echo "bench $wl rc=$?"
done
python bench.py --impl reference --steps 5 --warmup 3 > gpurun_out/bench_ref_kitti.json 2> gpurun_out/bench_ref.err; echo "bench ref rc=$?"
CMD="python bench.py --steps 3 --warmup 3 --no-cpu --no-graph"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "ncu list rc=$?"
tail -n 3 gpurun_out/pytest_gpu.log gpurun_out/smoke.log; cat gpurun_out/bench_*.json | cut -c1-1500
