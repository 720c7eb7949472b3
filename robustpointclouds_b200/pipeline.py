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

from typing import Dict, List, Optional, Sequence

import torch

from . import ops
from ._lib import PerturbCfg, VoxelCfg


class StepPipeline:
    def __init__(self, cfg: VoxelCfg, bounds: PerturbCfg, n_per_sample: Sequence[int], device: torch.device,
                 depth: int = 2, dynamic: bool = False, use_graph: bool = True, pre_step=None):
        """``pre_step``: optional callable run on the compute stream in front of every step (bench.py's L2 flush)."""
        self.cfg, self.bounds, self.n = cfg, bounds, [int(x) for x in n_per_sample]
        self.dev, self.depth, self.dynamic = device, depth, dynamic
        self.pre_step = pre_step
        N, C, F = sum(self.n), cfg.C, cfg.F
        self.s_in, self.s_cmp, self.s_out = (torch.cuda.Stream(device) for _ in range(3))
        f32 = dict(dtype=torch.float32, device=device)
        self.d_pts = [torch.empty((N, C), **f32) for _ in range(depth)]
        self.d_raw = [torch.empty((N, C), **f32) for _ in range(depth)]
        self.graw = [torch.empty((N, C), **f32) for _ in range(depth)]
        self.outs: List = []
        self.graphs: List[Optional[torch.cuda.CUDAGraph]] = [None] * depth
        self._gf: List[Optional[torch.Tensor]] = [None] * depth
        with torch.cuda.stream(self.s_cmp):
            for k in range(depth):
                self.d_pts[k].zero_(); self.d_raw[k].zero_()
                self.outs.append(self._fwd(k, None))
        torch.cuda.synchronize(device)
        self.h_out: List[Dict[str, torch.Tensor]] = [
            {name: torch.empty_like(t, device="cpu").pin_memory() for name, t in self._device_results(k).items()}
            for k in range(depth)]
        ev = lambda: [torch.cuda.Event() for _ in range(depth)]          # noqa: E731
        self.ev_in, self.ev_cmp, self.ev_out = ev(), ev(), ev()
        self.use_graph = use_graph
        self.count = 0

    # ---------------------------------------------------------------------------------------------
    def _fwd(self, k: int, out):
        if self.dynamic:
            return ops.dynamic_scatter_mean(self.d_pts[k], self.n, self.cfg, raw=self.d_raw[k], bounds=self.bounds, out=out)
        if out is None:
            return ops.hard_voxelize(self.d_pts[k], self.n, self.cfg, raw=self.d_raw[k], bounds=self.bounds,
                                     want_feats=True, want_maps=True, want_slots=False)
        return ops.hard_voxelize(self.d_pts[k], self.n, self.cfg, raw=self.d_raw[k], bounds=self.bounds, out=out)

    def _device_results(self, k: int) -> Dict[str, torch.Tensor]:
        o = self.outs[k]
        if self.dynamic:
            return dict(coors=o.coors, voxel_coors=o.voxel_coors, num_points=o.num_points, voxel_num=o.voxel_num,
                        feats=o.feats, grad_raw=self.graw[k])
        return dict(voxels=o.voxels, coors=o.coors, num_points=o.num_points, voxel_num=o.voxel_num, feats=o.feats,
                    grad_raw=self.graw[k])

    def _path(self, k: int, grad_feats: torch.Tensor):
        o = self._fwd(k, self.outs[k])
        ops.step_bwd(grad_feats, o.num_points, o.point2voxel, self.cfg, raw=self.d_raw[k], bounds=self.bounds,
                     out=self.graw[k])

    def _run(self, k: int, grad_feats: torch.Tensor):
        if not self.use_graph:
            return self._path(k, grad_feats)
        if self.graphs[k] is None or self._gf[k] is not grad_feats:      # (re)capture for this gradient buffer
            self._path(k, grad_feats)                                    # warm-up outside capture
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=self.s_cmp):
                self._path(k, grad_feats)
            self.graphs[k], self._gf[k] = g, grad_feats
        self.graphs[k].replay()

    # ---------------------------------------------------------------------------------------------
    @property
    def h2d_bytes(self) -> int:
        return 2 * self.d_pts[0].numel() * 4

    @property
    def d2h_bytes(self) -> int:
        return sum(t.numel() * t.element_size() for t in self.h_out[0].values())

    def submit(self, h_points: torch.Tensor, h_raw: torch.Tensor, grad_feats: torch.Tensor) -> int:
        """Enqueue one step: ``h_points`` / ``h_raw`` pinned host [sum n, C]; ``grad_feats`` device [cap, F] (the
        gradient the detector sends back).  Returns a ticket for ``result``."""
        k = self.count % self.depth
        with torch.cuda.stream(self.s_in):
            self.s_in.wait_event(self.ev_cmp[k])                         # kernels that read this slot's inputs are done
            self.d_pts[k].copy_(h_points, non_blocking=True)
            self.d_raw[k].copy_(h_raw, non_blocking=True)
            self.ev_in[k].record(self.s_in)
        with torch.cuda.stream(self.s_cmp):
            self.s_cmp.wait_event(self.ev_in[k])
            self.s_cmp.wait_event(self.ev_out[k])                        # this slot's previous results have left
            if self.pre_step is not None:
                self.pre_step()
            self._run(k, grad_feats)
            self.ev_cmp[k].record(self.s_cmp)
        with torch.cuda.stream(self.s_out):
            self.s_out.wait_event(self.ev_cmp[k])
            for name, t in self._device_results(k).items():
                self.h_out[k][name].copy_(t, non_blocking=True)
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
