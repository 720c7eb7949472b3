"""B200-native (sm_100a) implementation of the per-step adversarial input path of
temex12/RobustPointClouds: VoxelPerturber's bounded offset, hard / dynamic voxelization and the
HardSimpleVFE mean forward / backward, behind the reference's own Python API."""
from . import _lib, ops  # noqa: F401
from ._lib import PerturbCfg, RpcError, VoxelCfg  # noqa: F401
from .adversarial import adversarial_step, dynamic_step, extract_voxel_features, perturb_voxels  # noqa: F401
from .builder import ADVERSARIES, build_adversary  # noqa: F401
from .grad_sync import FlatGradSync, PeerAllReduce, shard_batch  # noqa: F401
from .pipeline import StepPipeline  # noqa: F401
from .strong import EnhancedPerturbation  # noqa: F401
from .transforms import PointShuffle, PointsRangeFilter  # noqa: F401
from .voxel_perturber import VoxelPerturber, perturber_bounds  # noqa: F401
from .voxelization import DynamicScatter, HardSimpleVFE, Voxelization, voxelize  # noqa: F401

__all__ = ["ops", "VoxelCfg", "PerturbCfg", "RpcError", "ADVERSARIES", "build_adversary", "VoxelPerturber",
           "perturber_bounds", "Voxelization", "voxelize", "HardSimpleVFE", "perturb_voxels",
           "extract_voxel_features", "adversarial_step", "dynamic_step", "DynamicScatter", "FlatGradSync", "PeerAllReduce", "shard_batch", "StepPipeline", "EnhancedPerturbation", "PointsRangeFilter", "PointShuffle"]
