#!/bin/bash
# peer all-reduce parity + timing (both variants), then the bench line with each: tools/gpu_ar.sh N
N=$1
mkdir -p gpurun_out
for mode in ll oneshot; do
  AR_MODE=$mode timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29621 tools/test_peer_allreduce.py 2>&1 | grep -v "^W\|^\*\*\*\|^$" | tail -3
done | tee gpurun_out/ar_$N.log
for ar in ll p2p nccl; do
  timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29622 bench.py --gpus $N --steps 200 --warmup 10 --no-cpu --allreduce $ar 2>gpurun_out/ar_bench_$ar.err | tail -n 1 | python -c "
import sys, json
d = json.loads(sys.stdin.read()); print('$ar', d['n_gpus'], 'us/step', round(d['ms_per_step']*1e3, 2), 'M pts/s', d['value'], d['config']['collective'][:60])"
done | tee -a gpurun_out/ar_$N.log
