"""B200-native (sm_100a) implementation of the per-step adversarial input path of
temex12/RobustPointClouds: VoxelPerturber's bounded offset, hard / dynamic voxelization and the
HardSimpleVFE mean forward / backward, behind the reference's own Python API."""
from . import _lib, ops  # noqa: F401
from ._lib import PerturbCfg, RpcError, VoxelCfg  # noqa: F401

__all__ = ["ops", "VoxelCfg", "PerturbCfg", "RpcError"]
