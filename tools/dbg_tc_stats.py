"""cycle counters of k_linear_tc's roles in CTA 0 (RPC_TC_DBG=4): where a tile's time goes"""
import ctypes, os, torch
os.environ["RPC_TC_DBG"] = "4"
from robustpointclouds_b200 import ops, _lib
dev = torch.device("cuda:0"); K = 835414
for NI, NO in [(64, 128), (128, 64)]:
    x = torch.randn(K, NI, device=dev); w = torch.randn(NO, NI, device=dev) * 0.1; b = torch.randn(NO, device=dev)
    y = torch.empty(K, NO, device=dev)
    mom = torch.empty(2 * NO, dtype=torch.float64, device=dev)
    bn = tuple(torch.rand(NI, device=dev) + 0.5 for _ in range(4))
    for tag, kw in (("plain", {}), ("moments", dict(moments=mom)), ("bn prologue", dict(bn=bn)), ("tanh", dict(tanh=True))):
        for _ in range(3):
            ops.linear_fwd(x, w, b, out=y, **kw)
        torch.cuda.synchronize()
        a, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); ops.linear_fwd(x, w, b, out=y, **kw); e.record(); torch.cuda.synchronize()
        out = (ctypes.c_ulonglong * 16)()
        _lib.lib().rpc_debug_tc_stats(out)
        s = list(out)
        jobs, tiles = max(s[2], 1), max(s[9], 1)
        print(f"{NI} -> {NO} [{tag}]: {a.elapsed_time(e) * 1e3:.1f} us; cycles per stage: loader wait-free {s[0] / jobs:.0f} work {s[1] / jobs:.0f} | "
              f"MMA wait-acc {s[3] / jobs:.0f} wait-full {s[4] / jobs:.0f} issue {s[5] / jobs:.0f} | per tile: epilogue wait-full {s[6] / tiles:.0f} ld {s[7] / tiles:.0f} slab+store {s[8] / tiles:.0f}")
