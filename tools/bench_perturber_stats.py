"""Times the perturber's statistics (std-normalise + four losses + their gradient) through the fused kernels and
through the ATen formulas of the mirror (what the reference runs), CUDA events, median of 50 after 10 warm-ups.
    python tools/bench_perturber_stats.py [K ...]"""
import sys
import statistics

import torch

from robustpointclouds_b200 import ops
from robustpointclouds_b200.voxel_perturber import VoxelPerturber, _PerturbLosses

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
        x = torch.randn(K, C, device=dev) * 10
        p0 = (torch.randn(K, C, device=dev) * 0.1)
        coef = torch.tensor([1.0, 2.0, 3.0, 4.0], device=dev)

        def aten():
            scale = torch.std(x, dim=0, keepdim=True) + 1e-6
            xn = torch.clamp(x / scale, -10.0, 10.0)
            p = p0.detach().requires_grad_(True)
            ld = VoxelPerturber.losses(p, fused=False)
            (ld["l2_norm"] + 2 * ld["intensity_loss"] + 3 * ld["bias_loss"] + 4 * ld["imbalance_loss"]).backward()
            return xn, p.grad

        def fused():
            xn = ops.normalize_fwd(x)
            p = p0.detach().requires_grad_(True)
            (_PerturbLosses.apply(p) * coef).sum().backward()
            return xn, p.grad

        def fused_raw():
            xn = ops.normalize_fwd(x)
            losses, consts = ops.perturb_losses_fwd(p0)
            return xn, ops.perturb_losses_bwd(p0, consts, coef)

        ta, tf, tr = timed(aten), timed(fused), timed(fused_raw)
        algo = K * C * 4 * (2 + 1 + 2)                            # x read twice + xn written; p read twice + grad written
        algo += K * C * 4
        print(f"K={K} C={C}: ATen {ta:.1f} us, fused+autograd {tf:.1f} us, fused C-ABI only {tr:.1f} us "
              f"({algo / tr / 1e3:.0f} GB/s algorithmic), speed-up {ta / tr:.1f}x")
