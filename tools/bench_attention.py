"""Times the perturber's attention gate (raw * sigmoid(Linear(relu(Linear(xn)))), forward + backward) through the fused
kernels and through the ATen modules the reference runs; CUDA events, median of 50 after 10 warm-ups.
    python tools/bench_attention.py [K ...]"""
import statistics
import sys

import torch
from torch import nn

from robustpointclouds_b200 import ops

dev = torch.device("cuda:0")


def timed(fn, n=50, warm=10):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(n):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); b.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    return statistics.median(ts)


for K in [int(a) for a in sys.argv[1:]] or [100_000, 480_000, 1_200_000]:
    for C in (4, 5):
        mid = C // 2
        att = nn.Sequential(nn.Linear(C, mid), nn.ReLU(inplace=True), nn.Linear(mid, 1), nn.Sigmoid()).to(dev)
        xn = torch.randn(K, C, device=dev)
        t0 = torch.tanh(torch.randn(K, C, device=dev))
        go = torch.randn(K, C, device=dev)
        w = [att[0].weight, att[0].bias, att[2].weight, att[2].bias]

        def aten():
            t = t0.detach().requires_grad_(True)
            for q in w:
                q.grad = None
            (t * att(xn)).backward(go)
            return t.grad

        def fused():
            with torch.no_grad():
                out = ops.attention_gate_fwd(xn, t0, *w)
                return out, ops.attention_gate_bwd(go, xn, t0, *w)

        def fused_fwd():
            with torch.no_grad():
                return ops.attention_gate_fwd(xn, t0, *w)

        def fused_bwd():
            with torch.no_grad():
                return ops.attention_gate_bwd(go, xn, t0, *w)

        ta, tf, t1, t2 = timed(aten), timed(fused), timed(fused_fwd), timed(fused_bwd)
        b_fwd, b_bwd = K * C * 4 * 3, K * C * 4 * 4             # xn, t in, out written; g, xn, t in, grad_t written
        print(f"K={K} C={C}: ATen fwd+bwd {ta:.1f} us; fused fwd+bwd {tf:.1f} us (fwd {t1:.1f} us = "
              f"{b_fwd / t1 / 1e3:.0f} GB/s, bwd {t2:.1f} us = {b_bwd / t2 / 1e3:.0f} GB/s algorithmic); "
              f"speed-up {ta / tf:.1f}x")
