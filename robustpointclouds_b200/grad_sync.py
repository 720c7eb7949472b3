"""Multi-GPU plumbing of the path.  The batch of sweeps shards across ranks with no data-path
collective; the only exchange is the all-reduce of the perturber's parameter gradients at step end, which
the reference gets implicitly from mmengine's DDP wrapper (/root/reference/custom_hook.py:24-28,
configs/adversarial/adversarial-centerpoint_voxel-nuscenes.py:110).  Here it is ONE collective over a flat
fp32 buffer (34 641 values for hidden [64,128,64], C=4), issued after the +-0.1 clamp hooks have run (they
fire in backward, before this is called -- same order as hook-then-DDP-reduce in the reference)."""
from __future__ import annotations

from typing import Iterable, List, Optional, Sequence

import torch
import torch.distributed as dist


def shard_batch(num_sweeps: int, rank: int, world_size: int) -> List[int]:
    """sweeps of the global batch owned by ``rank``: r, r + R, r + 2R, ..."""
    return list(range(rank, num_sweeps, world_size))


class FlatGradSync:
    """All-reduce (mean) of the gradients of ``params`` through one persistent flat buffer."""

    def __init__(self, params: Iterable[torch.nn.Parameter], group: Optional[dist.ProcessGroup] = None):
        self.params = [p for p in params if p.requires_grad]
        self.group = group
        n = sum(p.numel() for p in self.params)
        dev = self.params[0].device if self.params else torch.device("cpu")
        self.flat = torch.zeros(n, dtype=torch.float32, device=dev)
        self.offsets, o = [], 0
        for p in self.params:
            self.offsets.append(o)
            o += p.numel()

    def numel(self) -> int:
        return self.flat.numel()

    def pack(self) -> torch.Tensor:
        for p, o in zip(self.params, self.offsets):
            v = self.flat[o:o + p.numel()]
            if p.grad is None:
                v.zero_()
            else:
                v.copy_(p.grad.reshape(-1))
        return self.flat

    def unpack(self) -> None:
        for p, o in zip(self.params, self.offsets):
            g = self.flat[o:o + p.numel()].view_as(p)
            if p.grad is None:
                p.grad = g.clone()
            else:
                p.grad.copy_(g)

    def all_reduce(self, async_op: bool = False):
        """pack -> all_reduce(sum) -> / world -> unpack.  With ``async_op`` returns the work handle and the
        caller finishes with ``finish(handle)`` (overlap with the detector's backward)."""
        if not (dist.is_available() and dist.is_initialized()):
            return None
        self.pack()
        work = dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=self.group, async_op=async_op)
        if async_op:
            return work
        self.finish(None)
        return None

    def finish(self, work) -> None:
        if work is not None:
            work.wait()
        self.flat.div_(dist.get_world_size(self.group))
        self.unpack()
