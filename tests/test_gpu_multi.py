"""The path's only exchange on real GPUs: rpc_allreduce_ll / rpc_allreduce_oneshot over NVLink peer memory against an fp64
sum of rank-dependent patterns, eagerly and across CUDA-graph replays (call numbers advance inside the graph).  What it
replaces in the reference is DDP's gradient all-reduce (/root/reference/custom_hook.py:24-28).  Needs two GPUs: skipped on
the one-GPU test box; the world-2 host logic runs on CPU with gloo in tests/test_host.py."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r'''
import os, sys
import torch, torch.distributed as dist
sys.path.insert(0, os.environ["RPC_ROOT"])
from robustpointclouds_b200 import PeerAllReduce
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank); dev = torch.device("cuda", rank)
dist.init_process_group("nccl", device_id=dev)
n = 34641
idx = torch.arange(n, device=dev, dtype=torch.float64)
def pattern(r, call):                                   # exact in fp32: small integers and multiples of 2^-10
    return ((r + 1) * (call + 1) + (idx % 1024) / 1024.0)
ok = True
for mode in ("ll", "oneshot"):
    par = PeerAllReduce(n, dev, mode=mode)
    for call in range(6):
        x = pattern(rank, call).float()
        got = torch.empty_like(x)
        par.all_reduce(x, got)
        ref = sum(pattern(r, call) for r in range(world))
        ok &= bool(torch.equal(got.double(), ref))          # the sum is exact in fp32: bit-for-bit
    x, y = torch.empty(n, device=dev), torch.empty(n, device=dev)
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        par.all_reduce(x, y)                                 # warm-up on a side stream before the capture
    torch.cuda.synchronize(); dist.barrier()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        par.all_reduce(x, y)
    for call in range(5):
        x.copy_(pattern(rank, 10 + call).float())
        g.replay()
        torch.cuda.synchronize()
        ok &= bool(torch.equal(y.double(), sum(pattern(r, 10 + call) for r in range(world))))
    par.check()
flag = torch.tensor([1 if ok else 0], device=dev)
dist.all_reduce(flag, op=dist.ReduceOp.MIN)
if rank == 0:
    print("PEER_ALLREDUCE", "OK" if int(flag) == 1 else "MISMATCH")
dist.barrier(); dist.destroy_process_group()
'''


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs with a peer mapping")
def test_peer_allreduce_two_ranks(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    env = dict(os.environ, RPC_ROOT=ROOT)
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                          "--master-addr", "127.0.0.1", "--master-port", "29731", str(script)],
                         capture_output=True, text=True, timeout=600, env=env)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    assert "PEER_ALLREDUCE OK" in out.stdout, out.stdout[-2000:]
