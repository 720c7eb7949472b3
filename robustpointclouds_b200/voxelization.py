"""mmcv / mmdet3d-shaped front ends over the kernels:

* ``Voxelization(voxel_size, point_cloud_range, max_num_points, max_voxels=20000, deterministic=True)``
  -- mmcv ``ops.Voxelization``; ``forward(points[N,C])`` returns ``(voxels, coors, num_points)`` (hard) or
  ``coors`` (dynamic, when ``max_num_points == -1`` or ``max_voxels == -1``).
* ``voxelize(points_list, voxel_layer, voxel_type)`` -- mmdet3d ``Det3DDataPreprocessor.voxelize``: the whole
  batch in ONE launch, batch index in ``coors[:, 0]``, ``voxel_centers``; consumed by the reference at
  /root/reference/models/detectors/adversarial_voxelnet.py:66-71 and adversarial_centerpoint.py:62,191.
* ``HardSimpleVFE(num_features)`` -- mmdet3d voxel encoder, ``features[:, :, :F].sum(1) / num_points`` with
  autograd (called at adversarial_voxelnet.py:135-137, adversarial_centerpoint.py:100-102).

CUDA tensors only: these wrap sm_100a kernels and have no CPU path.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence, Tuple, Union

import torch
from torch import nn

from . import ops
from ._lib import VoxelCfg


def _pair(v) -> Tuple[int, int]:
    if isinstance(v, (tuple, list)):
        return int(v[0]), int(v[1])
    return int(v), int(v)


class Voxelization(nn.Module):
    def __init__(self, voxel_size: Sequence[float], point_cloud_range: Sequence[float], max_num_points: int,
                 max_voxels: Union[int, Sequence[int]] = 20000, deterministic: bool = True):
        super().__init__()
        self.voxel_size = list(voxel_size)
        self.point_cloud_range = list(point_cloud_range)
        self.max_num_points = int(max_num_points)
        self.max_voxels = _pair(max_voxels)
        # the B200 path is always deterministic (first-occurrence voxel order, point-order slots);
        # the flag is accepted for signature parity
        self.deterministic = deterministic
        pcr = torch.tensor(self.point_cloud_range, dtype=torch.float32)
        vs = torch.tensor(self.voxel_size, dtype=torch.float32)
        grid = torch.round((pcr[3:] - pcr[:3]) / vs).long()
        self.grid_size = grid                                   # (gx, gy, gz) like mmcv
        self.pcd_shape = [*grid.tolist()[:2], 1][::-1]

    @property
    def dynamic(self) -> bool:
        return self.max_num_points == -1 or self.max_voxels[0] == -1

    def cfg(self, C: int, F: Optional[int] = None) -> VoxelCfg:
        V = self.max_voxels[0] if self.training else self.max_voxels[1]
        if self.dynamic:
            return VoxelCfg.make(self.voxel_size, self.point_cloud_range, 1, 1, C, F)
        return VoxelCfg.make(self.voxel_size, self.point_cloud_range, self.max_num_points, V, C, F)

    def forward(self, points: torch.Tensor):
        C = points.shape[1]
        if self.dynamic:
            return ops.dynamic_voxelize(points, [points.shape[0]], self.cfg(C), ncol=3)
        out = ops.hard_voxelize(points, [points.shape[0]], self.cfg(C))
        m = out.total()                                         # host sync, as mmcv's voxel_num.item()
        return out.voxels[:m], out.coors[:m, 1:].contiguous(), out.num_points[:m]

    def __repr__(self) -> str:
        return (f"{self.__class__.__name__}(voxel_size={self.voxel_size}, point_cloud_range={self.point_cloud_range}, "
                f"max_num_points={self.max_num_points}, max_voxels={self.max_voxels}, deterministic={self.deterministic})")


@torch.no_grad()
def voxelize(points: List[torch.Tensor], voxel_layer: Voxelization, voxel_type: str = "hard",
             sync: bool = True) -> Dict[str, torch.Tensor]:
    """``Det3DDataPreprocessor.voxelize`` for a list of B point clouds: one batched launch instead of B.
    Returns ``voxels``, ``coors`` (batch, z, y, x), and for hard voxelization ``num_points`` and
    ``voxel_centers``.  ``sync=False`` keeps the outputs at capacity (no host sync) and adds the device
    tensor ``voxel_num`` ([B+1], last = total); rows beyond the total are undefined."""
    n = [int(p.shape[0]) for p in points]
    cat = torch.cat(points, dim=0) if len(points) > 1 else points[0]
    C = cat.shape[1]
    cfg = voxel_layer.cfg(C)
    if voxel_type == "hard":
        out = ops.hard_voxelize(cat, n, cfg)
        m = out.total() if sync else out.capacity
        coors = out.coors[:m]
        centers = ops.voxel_centers(coors, cfg, None if sync else out.voxel_num[-1:])
        d = dict(voxels=out.voxels[:m], coors=coors, num_points=out.num_points[:m], voxel_centers=centers)
        if not sync:
            d["voxel_num"] = out.voxel_num
        return d
    if voxel_type == "dynamic":
        coors = ops.dynamic_voxelize(cat, n, cfg, ncol=4)
        return dict(voxels=cat, coors=coors)
    raise ValueError(f"Invalid voxelization type {voxel_type}")


class _VfeMean(torch.autograd.Function):
    @staticmethod
    def forward(ctx, features: torch.Tensor, num_points: torch.Tensor, num_features: int):
        ctx.save_for_backward(num_points)
        ctx.shape = features.shape
        return ops.vfe_mean_fwd(features, num_points, num_features)

    @staticmethod
    def backward(ctx, grad_out):
        (num_points,) = ctx.saved_tensors
        _, P, C = ctx.shape
        return ops.vfe_mean_bwd(grad_out.contiguous(), num_points, P, C), None, None


class HardSimpleVFE(nn.Module):
    def __init__(self, num_features: int = 4):
        super().__init__()
        self.num_features = num_features
        self.fp16_enabled = False

    def forward(self, features: torch.Tensor, num_points: torch.Tensor, coors: torch.Tensor = None, *args, **kwargs):
        return _VfeMean.apply(features, num_points, self.num_features)


class _ScatterMean(torch.autograd.Function):
    @staticmethod
    def forward(ctx, points, coors, cfg: VoxelCfg, B: int):
        dv = ops.dynamic_scatter_mean(points, [points.shape[0]] + [0] * (B - 1), cfg, coors_in=coors)
        m = dv.total()                                          # host sync, as mmcv's unique() has
        ctx.save_for_backward(dv.num_points, dv.point2voxel)
        ctx.cfg = cfg
        ctx.mark_non_differentiable(dv.voxel_coors)
        return dv.feats[:m], dv.voxel_coors[:m]

    @staticmethod
    def backward(ctx, grad_feats, _):
        num_points, p2v = ctx.saved_tensors
        return ops.step_bwd(grad_feats.contiguous(), num_points, p2v, ctx.cfg), None, None, None


class DynamicScatter(nn.Module):
    """mmcv ``ops.DynamicScatter(voxel_size, point_cloud_range, average_points)``: ``forward(points[N,C],
    coors[N,4] (batch, z, y, x)) -> (voxel_feats[M,C], voxel_coors[M,4])``, voxels per sample in sorted
    (z, y, x) order, rows with a negative coordinate dropped.  Only the mean reduction
    (``average_points=True``) is on the path (DynamicSimpleVFE); max raises."""

    def __init__(self, voxel_size: Sequence[float], point_cloud_range: Sequence[float], average_points: bool):
        super().__init__()
        if not average_points:
            raise NotImplementedError("DynamicScatter(max) is not on the adversarial input path")
        self.voxel_size = list(voxel_size)
        self.point_cloud_range = list(point_cloud_range)
        self.average_points = average_points

    def forward(self, points: torch.Tensor, coors: torch.Tensor):
        C = points.shape[1]
        cfg = VoxelCfg.make(self.voxel_size, self.point_cloud_range, 1, 1, C, C)
        if coors.shape[-1] == 3:                                # single sample, (z, y, x)
            coors4 = torch.nn.functional.pad(coors, (1, 0), value=0)
            f, c = _ScatterMean.apply(points, coors4, cfg, 1)
            return f, c[:, 1:]
        B = int(coors[-1, 0].item()) + 1 if coors.shape[0] else 1   # mmcv reads the batch size the same way
        return _ScatterMean.apply(points, coors, cfg, max(B, 1))
