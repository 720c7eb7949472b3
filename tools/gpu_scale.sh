#!/bin/bash
# multi-GPU bench line at N ranks, as the driver launches it: tools/gpu_scale.sh N [extra bench args]
N=$1; shift
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 \
  bench.py --gpus $N --steps 200 --warmup 10 --no-cpu "$@" > gpurun_out/scale_$N.json 2> gpurun_out/scale_$N.err
echo "rc=$?"; tail -n 1 gpurun_out/scale_$N.json | cut -c1-400
