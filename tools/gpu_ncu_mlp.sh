#!/bin/bash
# one ncu --set full capture each of the perturber-MLP kernels at nuScenes size (K = 835 k): tools/gpu_ncu_mlp.sh
mkdir -p gpurun_out
export PYTHONPATH=.
cat > /tmp/mlp_kernels.py <<'PY'
import torch
from robustpointclouds_b200 import ops
dev = "cuda:0"
K = 835414
g = torch.randn(K, 128, device=dev); a = torch.randn(K, 64, device=dev)
y = torch.randn(K, 128, device=dev); go = torch.randn(K, 128, device=dev)
ga, be = torch.rand(128, device=dev) + 0.5, torch.randn(128, device=dev)
for _ in range(3):
    ops.linear_wgrad(g, a)
    z, mean, invstd = ops.bn_relu_fwd(y, ga, be, 1e-3, 0.1)
    ops.bn_relu_bwd(go, y, mean, invstd, ga, be)
torch.cuda.synchronize()
PY
python /tmp/mlp_kernels.py > gpurun_out/plain_mlp.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'k_linear_wgrad|k_bn_' -s 5 -c 5 -o gpurun_out/prof_mlp python /tmp/mlp_kernels.py > gpurun_out/ncu_mlp.log 2>&1
echo "ncu rc=$?"
ncu -i gpurun_out/prof_mlp.ncu-rep --page raw --csv > gpurun_out/prof_mlp_raw.csv 2>/dev/null; wc -c gpurun_out/prof_mlp_raw.csv
