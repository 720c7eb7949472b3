"""GPU versions of the two pre-path pipeline transforms the reference's configs run on the host in front of the
voxelizer (/root/reference/configs/_base_/kitti-3d-car.py:62-64: ``PointsRangeFilter``,
``PointShuffle``), with mmdet3d's call convention (``transform(input_dict)`` -> ``input_dict``, key ``points``)."""
from __future__ import annotations

from typing import Dict, Optional, Sequence

import torch

from . import ops


class PointsRangeFilter:
    """``points[in_range_3d(point_cloud_range)]``: the open box ``lo < p < hi``, order kept
    (``rpc_points_range_filter``)."""

    def __init__(self, point_cloud_range: Sequence[float]):
        self.pcd_range = [float(v) for v in point_cloud_range]

    def filter(self, points: torch.Tensor) -> torch.Tensor:
        out, _, count = ops.points_range_filter(points, self.pcd_range)
        return out[: int(count)]                                     # the one host sync boolean indexing has too

    def __call__(self, input_dict: Dict) -> Dict:
        input_dict["points"] = self.filter(input_dict["points"])
        return input_dict


class PointShuffle:
    """``points[randperm(N)]``: the permutation comes from torch's generator on the device, the row gather is
    ``rpc_gather_rows``."""

    def __init__(self, generator: Optional[torch.Generator] = None):
        self.generator = generator

    def shuffle(self, points: torch.Tensor) -> torch.Tensor:
        N, C = points.shape
        perm = torch.randperm(N, device=points.device, generator=self.generator).to(torch.int32)
        return ops.gather_rows(points, perm, N, C)

    def __call__(self, input_dict: Dict) -> Dict:
        input_dict["points"] = self.shuffle(input_dict["points"])
        return input_dict
