"""Per-rank PCIe rates with R ranks copying at the same time: each rank moves what StepPipeline moves
per KITTI step (23 MB in, 22.7 MB out as single packed copies, both directions at once) between pinned host memory and
its own GPU.  Launch under torchrun with R = 1, 2, 4, 8; rank 0 prints min / median / max over the ranks.
    python -m torch.distributed.run --nproc-per-node R tools/pcie_by_ranks.py"""
import os
import statistics
import torch
import torch.distributed as dist

rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
local = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)


def run(h2d, d2h, reps=40):
    hs = torch.empty(max(h2d, 1), dtype=torch.uint8).pin_memory(); ds = torch.empty(max(h2d, 1), dtype=torch.uint8, device=dev)
    do = torch.empty(max(d2h, 1), dtype=torch.uint8, device=dev); ho = torch.empty(max(d2h, 1), dtype=torch.uint8).pin_memory()
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()

    def once():
        if h2d:
            with torch.cuda.stream(s1): ds.copy_(hs, non_blocking=True)
        if d2h:
            with torch.cuda.stream(s2): ho.copy_(do, non_blocking=True)
    for _ in range(5): once()
    torch.cuda.synchronize()
    if world > 1: dist.barrier()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); s1.wait_event(a); s2.wait_event(a)
    for _ in range(reps): once()
    e1, e2 = torch.cuda.Event(), torch.cuda.Event(); e1.record(s1); e2.record(s2)
    torch.cuda.current_stream().wait_event(e1); torch.cuda.current_stream().wait_event(e2)
    b.record(); b.synchronize()
    return a.elapsed_time(b) / reps


for name, h2d, d2h in [("H2D alone 23.0 MB", 23040000, 0), ("D2H alone 22.7 MB", 0, 22656256), ("both directions", 23040000, 22656256),
                       ("e2e_real: 11.5 MB in, 3.5 MB out", 11520000, 3456256)]:
    ms = run(h2d, d2h)
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        g = [torch.empty_like(t) for _ in range(world)]; dist.all_gather(g, t); all_ms = [float(x) for x in g]
    else:
        all_ms = [ms]
    if rank == 0:
        med = statistics.median(all_ms)
        print(f"ranks {world}  {name:34s}: per rank {min(all_ms):.3f} / {med:.3f} / {max(all_ms):.3f} ms (min / median / max)  "
              f"H2D {h2d / med / 1e6:5.1f} GB/s  D2H {d2h / med / 1e6:5.1f} GB/s per rank,  node total {world * (h2d + d2h) / max(all_ms) / 1e6:6.1f} GB/s")
if world > 1:
    dist.barrier(); dist.destroy_process_group()
