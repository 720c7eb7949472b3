#!/bin/bash
cd /root/repo; mkdir -p gpurun_out
python - <<'PY'
import os
from robustpointclouds_b200.hostmem import gpu_local_cpus
print("allowed", len(os.sched_getaffinity(0)), "cpu_count", os.cpu_count())
for i in range(8):
    c = gpu_local_cpus(i)
    print("gpu", i, "local cpus", (c[0], c[-1], len(c)) if c else None)
PY
nvidia-smi topo -m 2>/dev/null | head -12 | cut -c1-200
for f in "" "--no-numa"; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29631 bench.py --gpus 8 --steps 50 --warmup 5 --no-cpu $f 2>/dev/null | tail -n 1 | python -c "
import sys, json
d = json.loads(sys.stdin.read()); print('numa' if '$f' == '' else 'no-numa', 'us/step', round(d['ms_per_step']*1e3,2), 'e2e ms', d['e2e']['ms_per_step'], d['e2e']['value'], d['e2e']['how'][-50:])"
done
