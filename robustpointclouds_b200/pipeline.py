"""Host-buffer front end of the fused adversarial step: pinned host tensors in, pinned host tensors out.

The reference runs this path on the host (mmcv's CPU voxelizer inside the data preprocessor), so a
drop-in caller that still lives on the host hands over host buffers.  At the named sizes the step itself takes
tens of microseconds while the PCIe copies take hundreds (23 MB in, 23 MB out for a KITTI batch), so the only
thing that matters end to end is to keep both PCIe directions busy at once: ``StepPipeline`` runs the
host->device copies, the kernels and the device->host copies on three CUDA streams over ``depth`` sets of
device buffers, ordered by events only (no host synchronisation until a result is asked for).

    pipe = StepPipeline(cfg, bounds, n_per_sample, device)
    t = pipe.submit(h_points, h_raw, grad_feats)      # asynchronous; returns a ticket
    res = pipe.result(t)                              # dict of pinned host tensors, valid until the slot is reused
"""
from __future__ import annotations

import dataclasses
from typing import Dict, List, Optional, Sequence

import torch

from . import ops
from ._lib import PerturbCfg, VoxelCfg


class StepPipeline:
    _HOST_FIELDS = ("voxels", "coors", "voxel_coors", "num_points", "voxel_num", "feats")
    LEAN_FIELDS = ("coors", "voxel_coors", "num_points", "voxel_num", "feats")      # everything but the padded voxels

    def __init__(self, cfg: VoxelCfg, bounds: PerturbCfg, n_per_sample: Sequence[int], device: torch.device,
                 depth: int = 2, dynamic: bool = False, use_graph: bool = True, pre_step=None,
                 host_fields: Optional[Sequence[str]] = None, raw_on_device: bool = False, grad_to_host: bool = True):
        """``pre_step``: optional callable run on the compute stream in front of every step (bench.py's L2 flush).
        ``host_fields``: which outputs travel back to the host (default: all of ``_HOST_FIELDS``; ``grad_raw`` always
        does).  The zero-padded ``voxels`` tensor is 35 % (KITTI) to 70 % (nuScenes) of the device->host bytes and is
        only an intermediate once the VFE mean is fused -- what the reference hands on after HardSimpleVFE is
        ``feats`` + ``coors`` (adversarial_voxelnet.py:135-140) -- so a caller that continues with the middle encoder
        leaves it out: ``host_fields=StepPipeline.LEAN_FIELDS``.  Outputs left out stay device tensors in ``outs``.
        ``raw_on_device=True`` / ``grad_to_host=False``: what the reference's training step moves and nothing else --
        the perturbation is produced on the device by the perturber (``d_raw[k]``, written by the caller on ``s_cmp`` or
        set once with ``set_raw``) and its gradient feeds the perturber's backward on the device (``graw[k]``), so only
        ``points`` cross PCIe on the way in and the voxel outputs on the way out (``submit(h_points, None, grad)``)."""
        self.raw_on_device, self.grad_to_host = raw_on_device, grad_to_host
        self.host_fields = tuple(self._HOST_FIELDS if host_fields is None else host_fields)
        unknown = set(self.host_fields) - set(self._HOST_FIELDS)
        if unknown:
            raise ValueError(f"unknown host_fields {sorted(unknown)}; choose from {self._HOST_FIELDS}")
        self.cfg, self.bounds, self.n = cfg, bounds, [int(x) for x in n_per_sample]
        self.dev, self.depth, self.dynamic = device, depth, dynamic
        self.pre_step = pre_step
        N, C, F = sum(self.n), cfg.C, cfg.F
        self.s_in, self.s_cmp, self.s_out = (torch.cuda.Stream(device) for _ in range(3))
        f32 = dict(dtype=torch.float32, device=device)
        # points and raw of a slot are the two halves of ONE device buffer, every result of a slot lives in ONE
        # packed device buffer with the same layout as its pinned host twin: one PCIe copy per direction and step
        # (measured on this pool: 0.50 ms for 23 MB each way as single copies, 0.56 ms split 2 + 6 ways)
        self.d_in = [torch.empty((2, N, C), **f32) for _ in range(depth)]
        self.d_pts = [b[0] for b in self.d_in]
        self.d_raw = [b[1] for b in self.d_in]
        self.outs: List = []
        self.graphs: List[Optional[torch.cuda.CUDAGraph]] = [None] * depth
        # every slot owns its workspace (a captured graph bakes the address in) and a persistent copy of the gradient the caller
        # hands in, so that ONE capture per slot serves every step
        ws_bytes = (ops.dynamic_workspace_bytes if dynamic else ops.hard_workspace_bytes)(cfg, self.n)
        self.ws = [torch.empty((ws_bytes + 256,), dtype=torch.uint8, device=device) for _ in range(depth)]
        self.d_gf: List[Optional[torch.Tensor]] = [None] * depth
        self.d_pack: List[torch.Tensor] = []
        self.graw: List[torch.Tensor] = []
        self._layout: Dict[str, tuple] = {}
        with torch.cuda.stream(self.s_cmp):
            for k in range(depth):
                self.d_in[k].zero_()
                o = self._fwd(k, None)                                   # shapes of the outputs at capacity
                named = {f.name: getattr(o, f.name) for f in dataclasses.fields(o)
                         if isinstance(getattr(o, f.name), torch.Tensor) and f.name in self.host_fields}
                if grad_to_host:
                    named["grad_raw"] = torch.empty((N, C), **f32)
                if not self._layout:
                    off = 0
                    for name, t in named.items():
                        self._layout[name] = (off, t.numel() * t.element_size(), t.dtype, tuple(t.shape))
                        off = (off + t.numel() * t.element_size() + 255) // 256 * 256
                    self._pack_bytes = off
                pack = torch.empty((self._pack_bytes,), dtype=torch.uint8, device=device)
                views = self._views(pack)
                self.d_pack.append(pack)
                self.graw.append(views.pop("grad_raw") if grad_to_host else torch.empty((N, C), **f32))
                self.outs.append(dataclasses.replace(o, **views))
        torch.cuda.synchronize(device)
        self.h_pack = [torch.empty((self._pack_bytes,), dtype=torch.uint8).pin_memory() for _ in range(depth)]
        self.h_out: List[Dict[str, torch.Tensor]] = [self._views(h) for h in self.h_pack]
        ev = lambda: [torch.cuda.Event() for _ in range(depth)]          # noqa: E731
        self.ev_in, self.ev_cmp, self.ev_out = ev(), ev(), ev()
        self.use_graph = use_graph
        self.count = 0

    # ---------------------------------------------------------------------------------------------
    def _fwd(self, k: int, out):
        if self.dynamic:
            return ops.dynamic_scatter_mean(self.d_pts[k], self.n, self.cfg, raw=self.d_raw[k], bounds=self.bounds, out=out,
                                            workspace=self.ws[k])
        if out is None:
            return ops.hard_voxelize(self.d_pts[k], self.n, self.cfg, raw=self.d_raw[k], bounds=self.bounds,
                                     want_feats=True, want_maps=True, want_slots=False, workspace=self.ws[k])
        return ops.hard_voxelize(self.d_pts[k], self.n, self.cfg, raw=self.d_raw[k], bounds=self.bounds, out=out,
                                 workspace=self.ws[k])

    def _views(self, pack: torch.Tensor) -> Dict[str, torch.Tensor]:
        return {name: pack[off:off + nbytes].view(dtype).view(shape)
                for name, (off, nbytes, dtype, shape) in self._layout.items()}

    def alloc_host_inputs(self):
        """(h_points, h_raw): the two halves of one pinned [2, N, C] buffer -- ``submit`` moves such a pair with a
        single copy"""
        h = torch.empty(tuple(self.d_in[0].shape), dtype=torch.float32).pin_memory()
        return h[0], h[1]

    def _path(self, k: int):
        o = self._fwd(k, self.outs[k])
        ops.step_bwd(self.d_gf[k], o.num_points, o.point2voxel, self.cfg, raw=self.d_raw[k], bounds=self.bounds,
                     out=self.graw[k])

    def _run(self, k: int, grad_feats: torch.Tensor):
        # the gradient is copied into the slot's own buffer (stream-ordered after whatever produced it: submit makes
        # s_cmp wait for the caller's stream), so the captured graph always reads the same address
        if self.d_gf[k] is None or self.d_gf[k].shape != grad_feats.shape:
            self.d_gf[k] = torch.empty(tuple(grad_feats.shape), dtype=torch.float32, device=self.dev)
            self.graphs[k] = None
        self.d_gf[k].copy_(grad_feats, non_blocking=True)
        if not self.use_graph:
            return self._path(k)
        if self.graphs[k] is None:
            self._path(k)                                                # warm-up outside capture
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=self.s_cmp):
                self._path(k)
            self.graphs[k] = g
        self.graphs[k].replay()

    # ---------------------------------------------------------------------------------------------
    def set_raw(self, raw: torch.Tensor) -> None:
        """``raw_on_device``: the perturbation of every slot := ``raw`` [sum n, C] (device), stream-ordered on ``s_cmp``"""
        with torch.cuda.stream(self.s_cmp):
            self.s_cmp.wait_stream(torch.cuda.current_stream(self.dev))
            for d in self.d_raw:
                d.copy_(raw, non_blocking=True)

    @property
    def h2d_bytes(self) -> int:
        return (1 if self.raw_on_device else 2) * self.d_pts[0].numel() * 4

    @property
    def d2h_bytes(self) -> int:
        return int(self._pack_bytes)                                 # what the one device->host copy moves

    def submit(self, h_points: torch.Tensor, h_raw: Optional[torch.Tensor], grad_feats: torch.Tensor) -> int:
        """Enqueue one step: ``h_points`` / ``h_raw`` pinned host [sum n, C] (``h_raw=None`` with ``raw_on_device``);
        ``grad_feats`` device [cap, F] (the gradient the detector sends back).  Returns a ticket for ``result``."""
        if (h_raw is None) != self.raw_on_device:
            raise ValueError("h_raw must be None exactly when the pipeline was built with raw_on_device=True")
        k = self.count % self.depth
        with torch.cuda.stream(self.s_in):
            self.s_in.wait_event(self.ev_cmp[k])                         # kernels that read this slot's inputs are done
            nb = h_points.numel() * h_points.element_size()
            if h_raw is None:
                self.d_pts[k].copy_(h_points, non_blocking=True)
            elif (h_raw.data_ptr() == h_points.data_ptr() + nb and h_points.is_contiguous() and h_raw.is_contiguous()
                    and h_points.dtype == torch.float32 and h_raw.dtype == torch.float32):
                both = torch.as_strided(h_points, tuple(self.d_in[k].shape), self.d_in[k].stride())
                self.d_in[k].copy_(both, non_blocking=True)              # the halves of one pinned buffer: one copy
            else:
                self.d_pts[k].copy_(h_points, non_blocking=True)
                self.d_raw[k].copy_(h_raw, non_blocking=True)
            self.ev_in[k].record(self.s_in)
        caller = torch.cuda.current_stream(self.dev)                     # whatever produced grad_feats runs here
        with torch.cuda.stream(self.s_cmp):
            self.s_cmp.wait_stream(caller)
            self.s_cmp.wait_event(self.ev_in[k])
            self.s_cmp.wait_event(self.ev_out[k])                        # this slot's previous results have left
            if self.pre_step is not None:
                self.pre_step()
            self._run(k, grad_feats)
            self.ev_cmp[k].record(self.s_cmp)
        with torch.cuda.stream(self.s_out):
            self.s_out.wait_event(self.ev_cmp[k])
            self.h_pack[k].copy_(self.d_pack[k], non_blocking=True)
            self.ev_out[k].record(self.s_out)
        self.count += 1
        return self.count - 1

    def result(self, ticket: int) -> Dict[str, torch.Tensor]:
        """Pinned host results of a submitted step (waits for its device->host copies).  The tensors are reused
        ``depth`` submissions later."""
        if ticket < self.count - self.depth or ticket >= self.count:
            raise ValueError("ticket is no longer (or not yet) resident")
        k = ticket % self.depth
        self.ev_out[k].synchronize()
        return self.h_out[k]

    def drain(self) -> None:
        for e in self.ev_out:
            e.synchronize()
