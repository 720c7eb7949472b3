#!/bin/bash
# development run of the hard-voxelization kernel: parity tests, per-CTA stamps, bench lines, A/B against a saved build
# usage: tools/gpu_try.sh [tests|alltests|stamps|stampsx|bench|small|ab ...]
mkdir -p gpurun_out
for what in "$@"; do
case $what in
tests)
  timeout 900 python -m pytest tests/test_gpu_hard.py tests/test_gpu_host.py tests/test_gpu_pipeline.py -x -q > gpurun_out/pytest_hard.log 2>&1
  echo "pytest hard rc=$?"; tail -n 15 gpurun_out/pytest_hard.log ;;
alltests)
  timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1
  echo "pytest all rc=$?"; tail -n 15 gpurun_out/pytest_gpu.log ;;
stamps)
  for wl in kitti nuscenes; do
    echo "== stamps $wl"; timeout 200 python tools/dbg_stamps.py $wl 2>&1 | tail -n 10
  done | tee gpurun_out/stamps.log
  echo "== stamps kitti 2x16k"; BATCH=2 POINTS=16000 NCTA=16 timeout 200 python tools/dbg_stamps.py kitti 2>&1 | tail -n 10 | tee -a gpurun_out/stamps.log ;;
stampsx)
  for v in "RPC_TABLE_PREFETCH=2" "RPC_TABLE_PREFETCH=1" "RPC_SYNTH_NULL=0" "NOFLUSH=1"; do
    for wl in kitti nuscenes; do
      echo "== stamps $wl $v"; env $v timeout 200 python tools/dbg_stamps.py $wl 2>&1 | tail -n 10
    done
  done | tee gpurun_out/stampsx.log ;;
bench)
  for wl in kitti nuscenes; do
    timeout 600 python bench.py --workload $wl --no-cpu --no-also > gpurun_out/bench_$wl.json 2> gpurun_out/bench_$wl.err; echo "bench $wl rc=$?"
    python - <<PY
import json
try:
    d = json.loads(open('gpurun_out/bench_$wl.json').read().strip().splitlines()[-1]); r = d['roofline']
    print('$wl', 'us/step', round(d['ms_per_step'] * 1e3, 2), 'e2e ms', d['e2e'].get('ms_per_step'), r['stage_us'], r['fwd_phase_us'], 'step frac', r['step']['frac'])
except Exception as e:
    print('bench $wl: no line', e); print(open('gpurun_out/bench_$wl.err').read()[-1500:])
PY
  done ;;
small)
  for n in 16000 64000; do
    timeout 300 python bench.py --workload kitti --batch 2 --points $n --no-cpu 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1]); r = d['roofline']
print('kitti 2 x $n', 'us/step', round(d['ms_per_step'] * 1e3, 2), r['stage_us'], r['fwd_phase_us'])"
  done ;;
ab|abr)
  fl=write; [ $what = abr ] && fl=write+read
  for wl in kitti nuscenes; do for rep in 1 2; do for lib in default build_ab/librpc_r01.so; do
    if [ $lib = default ]; then unset RPC_B200_LIB; else export RPC_B200_LIB=/root/repo/$lib; fi
    timeout 300 python bench.py --workload $wl --steps 200 --warmup 20 --no-cpu --no-also --flush $fl 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1]); r = d['roofline']
print('$wl', '$fl', '$lib', 'us/step', round(d['ms_per_step'] * 1e3, 2), r['stage_us'])"
  done; done; done 2>&1 | tee gpurun_out/ab.log ;;
esac
done
