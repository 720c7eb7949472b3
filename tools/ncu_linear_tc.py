"""a few launches of rpc_linear_fwd (64 -> 128 with BatchNorm prologue + moments, 128 -> 64) and rpc_linear_wgrad (tensor-core
path) for an ncu capture"""
import torch
from robustpointclouds_b200 import ops
dev = torch.device("cuda:0"); K = 835414
for NI, NO in [(64, 128), (128, 64)]:
    x = torch.randn(K, NI, device=dev); w = torch.randn(NO, NI, device=dev) * 0.1; b = torch.randn(NO, device=dev)
    y = torch.empty(K, NO, device=dev); mom = torch.empty(2 * NO, dtype=torch.float64, device=dev)
    bn = tuple(torch.rand(NI, device=dev) + 0.5 for _ in range(4))
    for _ in range(3):
        ops.linear_fwd(x, w, b, out=y, bn=bn, moments=mom)
    g = torch.randn(K, NO, device=dev)
    for _ in range(3):
        ops.linear_wgrad(g, x, bn=bn)
torch.cuda.synchronize()
print("ok")
