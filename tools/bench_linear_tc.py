"""Per-kernel timing of the tensor-core Linear kernels against cuBLAS (torch.addmm / mm) at the perturber's layer shapes.
    PYTHONPATH=. python tools/bench_linear_tc.py [K ...]"""
import statistics, sys
import torch
from robustpointclouds_b200 import ops
dev = torch.device("cuda:0")
def ev(fn, n=20, warm=3):
    for _ in range(warm): fn()
    ts = []
    for _ in range(n):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); b.synchronize(); ts.append(a.elapsed_time(b) * 1e3)
    return statistics.median(ts)
for K in [int(a) for a in sys.argv[1:]] or [131687, 835414]:
    print(f"K = {K}")
    for NI, NO in [(4, 64), (64, 128), (128, 64), (64, 4)]:
        x = torch.randn(K, NI, device=dev); w = torch.randn(NO, NI, device=dev) * 0.1; b = torch.randn(NO, device=dev)
        y = torch.empty(K, NO, device=dev); wt = w.t().contiguous()
        mom = torch.empty(2 * NO, dtype=torch.float64, device=dev)
        bn = tuple(torch.rand(NI, device=dev) + 0.5 for _ in range(4))
        byt = 4 * K * (NI + NO)
        t0 = ev(lambda: ops.linear_fwd(x, w, b, out=y))
        t1 = ev(lambda: ops.linear_fwd(x, w, b, out=y, bn=bn, moments=mom if NO > 8 else None))
        t2 = ev(lambda: torch.addmm(b, x, wt, out=y))
        g = torch.randn(K, NO, device=dev); gx = torch.empty(K, NI, device=dev)
        t3 = ev(lambda: ops.linear_fwd(g, w, transposed=True, out=gx)) if NI > 8 or NO >= 16 else float("nan")
        t4 = ev(lambda: torch.mm(g, w, out=gx))
        print(f"  {NI:3d} -> {NO:3d}: fwd {t0:7.1f} us ({byt / t0 / 1e3:6.0f} GB/s)  fwd+bn+moments {t1:7.1f} us  cuBLAS addmm {t2:7.1f} us | "
              f"dgrad {t3:7.1f} us  cuBLAS mm {t4:7.1f} us")
    for H in (64, 128):
        yv = torch.randn(K, H, device=dev); g = torch.randn(K, H, device=dev); z = torch.empty_like(yv)
        st = tuple(torch.rand(H, device=dev) + 0.5 for _ in range(4))
        a = torch.randn(K, 64 if H == 128 else 128, device=dev)
        print(f"  H = {H}: bn_relu_apply {ev(lambda: ops.bn_relu_apply(yv, *st, out=z)):7.1f} us  bn_relu_bwd {ev(lambda: ops.bn_relu_bwd(g, yv, *st)):7.1f} us  "
              f"wgrad {ev(lambda: ops.linear_wgrad(g, a)):7.1f} us")
