"""Times training-mode BatchNorm1d + ReLU, forward and backward, through rpc_bn_relu_fwd/bwd and through ATen
(F.batch_norm + relu and autograd); CUDA events, median of 30 after 5 warm-ups.
    python tools/bench_bn_relu.py"""
import statistics

import torch
import torch.nn.functional as F

from robustpointclouds_b200 import ops

dev = "cuda:0"


def t(fn, n=30):
    for _ in range(5):
        fn()
    ts = []
    for _ in range(n):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); e1.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    return statistics.median(ts)


for K, H in [(131687, 64), (131687, 128), (835414, 64), (835414, 128)]:
    y = torch.randn(K, H, device=dev)
    ga, be = torch.rand(H, device=dev) + 0.5, torch.randn(H, device=dev)
    rm, rv = torch.zeros(H, device=dev), torch.ones(H, device=dev)
    go = torch.randn(K, H, device=dev)
    z, mean, invstd = ops.bn_relu_fwd(y, ga, be, 1e-3, 0.1, rm, rv)
    yr = y.clone().requires_grad_(True)
    gar, ber = ga.clone().requires_grad_(True), be.clone().requires_grad_(True)

    def aten_fwd():
        return torch.relu(F.batch_norm(yr, rm, rv, gar, ber, True, 0.1, 1e-3))

    out = aten_fwd()

    def aten_bwd():
        return torch.autograd.grad(out, (yr, gar, ber), go, retain_graph=True)

    tf, tb = t(lambda: ops.bn_relu_fwd(y, ga, be, 1e-3, 0.1, rm, rv)), t(lambda: ops.bn_relu_bwd(go, y, mean, invstd, ga, be))
    af, ab = t(aten_fwd), t(aten_bwd)
    b = K * H * 4
    print(f"K={K} H={H}: fused fwd {tf:.1f} us ({3 * b / tf / 1e3:.0f} GB/s) bwd {tb:.1f} us ({5 * b / tb / 1e3:.0f} GB/s); "
          f"ATen fwd {af:.1f} us bwd {ab:.1f} us")
