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


class PeerAllReduce:
    """Sum all-reduce of a small fp32 buffer over NVLink peer memory in ONE launch (``rpc_allreduce_ll`` /
    ``rpc_allreduce_oneshot``, csrc/allreduce.cuh), sums taken in rank order (bit-identical on every rank).  The
    symmetric, peer-mapped allocation comes from ``torch.distributed._symmetric_memory`` (plumbing); the
    exchange itself is this library's kernel.  One NCCL-free launch instead of a ~16 us NCCL all-reduce."""

    def __init__(self, numel: int, device: torch.device, group: Optional[dist.ProcessGroup] = None,
                 mode: str = "ll"):
        """``mode="ll"``: push variant, value + call number in one 8-byte store (``rpc_allreduce_ll``, one one-way
        NVLink latency); ``mode="oneshot"``: staged buffers + flags, peers read (``rpc_allreduce_oneshot``)."""
        import ctypes

        import torch.distributed._symmetric_memory as symm_mem

        from . import _lib
        if mode not in ("ll", "oneshot"):
            raise ValueError(mode)
        self._lib, self._ct, self.mode = _lib, ctypes, mode
        self.group = group if group is not None else dist.group.WORLD
        self.n = int(numel)
        self.npad = (self.n + 3) // 4 * 4
        world = dist.get_world_size(self.group)
        if mode == "ll":
            words = _lib.lib().rpc_allreduce_ll_bytes(world, self.n) // 4
        else:
            words = 2 * self.npad + _lib.lib().rpc_allreduce_flag_bytes() // 4
        self.sym = symm_mem.empty(words, dtype=torch.float32, device=device)
        self.sym.zero_()
        self.hdl = symm_mem.rendezvous(self.sym, self.group)
        self.rank, self.world = self.hdl.rank, self.hdl.world_size
        torch.cuda.synchronize(device)
        dist.barrier(self.group)                                       # every rank's flags are zero from here on
        ptrs = [int(q) for q in self.hdl.buffer_ptrs]
        vp = ctypes.c_void_p
        self._recv = (vp * self.world)(*ptrs)
        self._stage = [(vp * self.world)(*[q + 4 * self.npad * b for q in ptrs]) for b in (0, 1)]
        self._flags = (vp * self.world)(*[q + 8 * self.npad for q in ptrs])
        self.state = torch.zeros(4, dtype=torch.int32, device=device)   # [0] completed calls, [1] ticket, [2] status
        self.dev = device

    def all_reduce(self, src: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """out = sum over ranks of src (out defaults to src, in place); one stream-ordered launch with fixed
        parameters (CUDA-graph capturable), no host sync"""
        out = src if out is None else out
        if src.numel() != self.n or src.dtype != torch.float32 or not src.is_contiguous():
            raise ValueError(f"expected a contiguous fp32 tensor of {self.n} values")
        stream = torch.cuda.current_stream(self.dev).cuda_stream
        if self.mode == "ll":
            st = self._lib.lib().rpc_allreduce_ll(self._recv, self.rank, self.world, self.n, src.data_ptr(),
                                                 out.data_ptr(), self.state.data_ptr(), self.state.data_ptr() + 8, stream)
            self._lib.check(st, "rpc_allreduce_ll")
            return out
        st = self._lib.lib().rpc_allreduce_oneshot(self._stage[0], self._stage[1], self._flags, self.rank, self.world,
                                                  self.n, src.data_ptr(), out.data_ptr(), self.state.data_ptr(),
                                                  self.state.data_ptr() + 8, stream)
        self._lib.check(st, "rpc_allreduce_oneshot")
        return out

    def check(self) -> None:
        """host sync: raise if a peer timed out in any call so far"""
        if int(self.state[2].item()) != 0:
            raise RuntimeError("peer all-reduce: a peer did not arrive within 2 s")


class FlatGradSync:
    """All-reduce (mean) of the gradients of ``params`` through one persistent flat buffer."""

    def __init__(self, params: Iterable[torch.nn.Parameter], group: Optional[dist.ProcessGroup] = None,
                 peer: bool = False):
        """``peer=True`` (CUDA, one node, NVLink): the sum goes through ``PeerAllReduce`` instead of NCCL."""
        self.params = [p for p in params if p.requires_grad]
        self.group = group
        self.peer_requested = peer
        self.peer: Optional[PeerAllReduce] = None
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
        if self.peer_requested and self.flat.is_cuda and not async_op:
            if self.peer is None:
                self.peer = PeerAllReduce(self.flat.numel(), self.flat.device, self.group)
            self.peer.all_reduce(self.flat)
            self.finish(None)
            return None
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
