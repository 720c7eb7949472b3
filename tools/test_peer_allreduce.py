"""torchrun --nproc-per-node N tools/test_peer_allreduce.py : PeerAllReduce against NCCL, then timing of both"""
import os, sys, time
import torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from robustpointclouds_b200 import PeerAllReduce
rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local); dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
n = int(os.environ.get("AR_N", 34641))
mode = os.environ.get("AR_MODE", "ll")
par = PeerAllReduce(n, dev, mode=mode)
ok = True
for it in range(20):
    x = torch.randn(n, device=dev, generator=torch.Generator(dev).manual_seed(100 * it + rank))
    ref = x.clone(); dist.all_reduce(ref)
    got = torch.empty_like(x); par.all_reduce(x, got)
    torch.cuda.synchronize()
    ok &= bool(torch.allclose(got, ref, rtol=1e-5, atol=1e-5))
    g = [torch.empty_like(got) for _ in range(dist.get_world_size())]; dist.all_gather(g, got)
    ok &= all(torch.equal(g[0], t) for t in g)                     # bit-identical on every rank
par.check()
x = torch.randn(n, device=dev)
def timeit(fn, k=200):
    for _ in range(20): fn()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(k): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / k * 1e3
t_nccl = timeit(lambda: dist.all_reduce(x))
y = torch.empty_like(x)
t_peer = timeit(lambda: par.all_reduce(x, y))
g = torch.cuda.CUDAGraph()
s2 = torch.cuda.Stream()
with torch.cuda.stream(s2):
    par.all_reduce(x, y)
torch.cuda.synchronize(); dist.barrier()
with torch.cuda.graph(g):
    par.all_reduce(x, y)
t_graph = timeit(g.replay)
ref = x.clone(); dist.all_reduce(ref); torch.cuda.synchronize()
ok &= bool(torch.allclose(y, ref, rtol=1e-5, atol=1e-5))
par.check()
if rank == 0:
    print(f"world {dist.get_world_size()} n {n}: parity {'OK' if ok else 'FAIL'}  nccl {t_nccl:.1f} us  peer {mode} {t_peer:.1f} us (eager), {t_graph:.1f} us (graph replay)")
dist.barrier(); dist.destroy_process_group()
