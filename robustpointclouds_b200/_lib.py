"""ctypes binding of ``librpc_b200.so`` (the C-ABI declared in ``include/rpc_b200.h``).

The library is the product: there is no CPU or PyTorch fallback.  If it has not been built
(``python -c "import __graft_entry__ as g; g.build()"`` or ``make -C robustpointclouds_b200/csrc``)
importing this module raises, and every op that needs a kernel raises when handed a non-CUDA tensor.
"""
from __future__ import annotations

import ctypes
import os
from typing import Optional, Sequence

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("RPC_B200_LIB") or os.path.join(_HERE, "librpc_b200.so")   # override: kernel experiments

RPC_MAX_BATCH = 32
RPC_MAX_CHANNELS = 8

# every symbol include/rpc_b200.h declares (tests check the header against this list and the .so)
SYMBOLS = (
    "rpc_abi_version", "rpc_strerror", "rpc_grid_size", "rpc_workspace_bytes", "rpc_voxel_capacity",
    "rpc_voxelize_hard", "rpc_step_bwd", "rpc_voxelize_dynamic", "rpc_vfe_mean_fwd", "rpc_vfe_mean_bwd",
    "rpc_profile_enable", "rpc_profile_read", "rpc_workspace_stamps_offset",
    "rpc_valid_rows_workspace_bytes", "rpc_valid_rows", "rpc_scatter_rows", "rpc_gather_rows",
    "rpc_perturb_apply_fwd", "rpc_perturb_apply_bwd",
    "rpc_dynamic_workspace_bytes", "rpc_dynamic_scatter_mean_fwd", "rpc_dynamic_stamps_offset", "rpc_allreduce_flag_bytes", "rpc_allreduce_oneshot", "rpc_allreduce_ll_bytes", "rpc_allreduce_ll",
    "rpc_voxel_centers", "rpc_points_range_filter", "rpc_points_range_filter_static", "rpc_momentum_apply_fwd", "rpc_momentum_apply_bwd", "rpc_moments_bytes", "rpc_normalize_fwd", "rpc_perturb_losses_fwd", "rpc_perturb_losses_bwd",
    "rpc_attention_scratch_bytes", "rpc_attention_gate_fwd", "rpc_attention_gate_bwd", "rpc_linear_wgrad",
    "rpc_bn_scratch_bytes", "rpc_bn_relu_fwd", "rpc_bn_relu_bwd",
    "rpc_linear_fwd", "rpc_bn_finalize", "rpc_bn_relu_apply", "rpc_trunk_fwd", "rpc_trunk_fwd_eval", "rpc_trunk_bwd", "rpc_linear_wgrad_bn", "rpc_linear_dgrad_bnred", "rpc_bn_relu_bwd_sums",
)


class VoxelCfg(ctypes.Structure):
    """``rpc_voxel_cfg``: mirrors mmcv ``Voxelization(voxel_size, point_cloud_range, max_num_points,
    max_voxels)`` plus the point width C and HardSimpleVFE's ``num_features`` F."""
    _fields_ = [("voxel_size", ctypes.c_float * 3), ("pcr", ctypes.c_float * 6),
                ("max_points", ctypes.c_int32), ("max_voxels", ctypes.c_int32),
                ("C", ctypes.c_int32), ("F", ctypes.c_int32)]

    @classmethod
    def make(cls, voxel_size: Sequence[float], pcr: Sequence[float], max_points: int, max_voxels: int,
             C: int, F: Optional[int] = None) -> "VoxelCfg":
        c = cls()
        c.voxel_size[:] = [float(v) for v in voxel_size]
        c.pcr[:] = [float(v) for v in pcr]
        c.max_points, c.max_voxels, c.C, c.F = int(max_points), int(max_voxels), int(C), int(F or C)
        return c


class PerturbCfg(ctypes.Structure):
    """``rpc_perturb_cfg``: per-channel scale ``eb`` and clamp ``fb`` of VoxelPerturber."""
    _fields_ = [("eb", ctypes.c_float * RPC_MAX_CHANNELS), ("fb", ctypes.c_float * RPC_MAX_CHANNELS)]

    @classmethod
    def make(cls, eb: Sequence[float], fb: Sequence[float]) -> "PerturbCfg":
        c = cls()
        for i, (a, b) in enumerate(zip(eb, fb)):
            c.eb[i], c.fb[i] = float(a), float(b)
        return c


RPC_TRUNK_MAX_HIDDEN = 6


class TrunkDesc(ctypes.Structure):
    """``rpc_trunk_desc``: the layers of VoxelPerturber's MLP trunk (voxel_perturber.py:82-103) as device pointers."""
    _fields_ = [("C", ctypes.c_int32), ("nl", ctypes.c_int32), ("width", ctypes.c_int32 * RPC_TRUNK_MAX_HIDDEN),
                ("w", ctypes.c_void_p * (RPC_TRUNK_MAX_HIDDEN + 1)), ("b", ctypes.c_void_p * (RPC_TRUNK_MAX_HIDDEN + 1)),
                ("gamma", ctypes.c_void_p * RPC_TRUNK_MAX_HIDDEN), ("beta", ctypes.c_void_p * RPC_TRUNK_MAX_HIDDEN),
                ("running_mean", ctypes.c_void_p * RPC_TRUNK_MAX_HIDDEN), ("running_var", ctypes.c_void_p * RPC_TRUNK_MAX_HIDDEN),
                ("eps", ctypes.c_float * RPC_TRUNK_MAX_HIDDEN), ("momentum", ctypes.c_float * RPC_TRUNK_MAX_HIDDEN)]


class RpcError(RuntimeError):
    def __init__(self, status: int, where: str):
        self.status = status
        super().__init__(f"{where}: {strerror(status)} (status {status})")


_lib: Optional[ctypes.CDLL] = None


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build the sm_100a kernels first "
            "(make -C robustpointclouds_b200/csrc, or __graft_entry__.build()). "
            "There is no CPU fallback.")
    L = ctypes.CDLL(LIB_PATH)
    vp, i32, i64, sz = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_size_t
    cfgp, pcp = ctypes.POINTER(VoxelCfg), ctypes.POINTER(PerturbCfg)
    i32p = ctypes.POINTER(ctypes.c_int32)
    L.rpc_abi_version.restype = i32
    L.rpc_strerror.restype = ctypes.c_char_p
    L.rpc_strerror.argtypes = [i32]
    L.rpc_grid_size.argtypes = [cfgp, i32p]
    L.rpc_workspace_bytes.restype = sz
    L.rpc_workspace_bytes.argtypes = [cfgp, i32, i32p]
    L.rpc_voxel_capacity.restype = i64
    L.rpc_voxel_capacity.argtypes = [cfgp, i32, i32p]
    L.rpc_workspace_stamps_offset.restype = i64
    L.rpc_workspace_stamps_offset.argtypes = [cfgp, i32, i32p]
    L.rpc_voxelize_hard.argtypes = [cfgp, vp, i32p, i32, vp, pcp, vp, vp, vp, vp, vp, vp, vp, vp, sz, vp]
    L.rpc_step_bwd.argtypes = [cfgp, vp, vp, vp, vp, pcp, i64, vp, vp]
    L.rpc_voxelize_dynamic.argtypes = [cfgp, vp, i32p, i32, vp, pcp, vp, i32, vp]
    L.rpc_dynamic_workspace_bytes.restype = sz
    L.rpc_dynamic_workspace_bytes.argtypes = [cfgp, i32, i32p]
    L.rpc_dynamic_stamps_offset.restype = i64
    L.rpc_dynamic_stamps_offset.argtypes = [cfgp, i32, i32p]
    L.rpc_dynamic_scatter_mean_fwd.argtypes = [cfgp, vp, i32p, i32, vp, pcp, vp, vp, vp, vp, vp, vp, vp, vp, sz, vp]
    L.rpc_vfe_mean_fwd.argtypes = [vp, vp, i64, i32, i32, i32, vp, vp, vp]
    L.rpc_vfe_mean_bwd.argtypes = [vp, vp, i64, i32, i32, i32, vp, vp, vp]
    L.rpc_valid_rows_workspace_bytes.restype = sz
    L.rpc_valid_rows_workspace_bytes.argtypes = [i64, i32]
    L.rpc_valid_rows.argtypes = [vp, vp, i64, i32, i32, i32, vp, vp, vp, vp, vp, vp, sz, vp]
    L.rpc_scatter_rows.argtypes = [vp, vp, vp, i64, i32, i32, vp, vp]
    L.rpc_gather_rows.argtypes = [vp, vp, i64, vp, i32, vp, vp]
    L.rpc_perturb_apply_fwd.argtypes = [vp, vp, pcp, i64, vp, i32, vp, vp, vp]
    L.rpc_perturb_apply_bwd.argtypes = [vp, vp, pcp, i64, vp, i32, vp, vp]
    L.rpc_allreduce_flag_bytes.restype = sz
    L.rpc_allreduce_ll_bytes.restype = sz
    L.rpc_allreduce_ll_bytes.argtypes = [i32, i64]
    L.rpc_allreduce_ll.argtypes = [ctypes.POINTER(vp), i32, i32, i64, vp, vp, vp, vp, vp]
    L.rpc_allreduce_oneshot.argtypes = [ctypes.POINTER(vp), ctypes.POINTER(vp), ctypes.POINTER(vp), i32, i32, i64, vp, vp, vp, vp, vp]
    L.rpc_voxel_centers.argtypes = [ctypes.POINTER(VoxelCfg), vp, i32, i64, vp, vp, vp]
    L.rpc_momentum_apply_fwd.argtypes = [vp, vp, vp, i32, ctypes.c_float, ctypes.c_float, i64, vp, vp, vp, vp, vp]
    L.rpc_momentum_apply_bwd.argtypes = [vp, vp, vp, vp, i32, ctypes.c_float, i64, vp, vp, vp]
    L.rpc_points_range_filter.argtypes = [vp, i64, i32, ctypes.POINTER(ctypes.c_float), vp, vp, vp, vp, sz, vp]
    L.rpc_points_range_filter_static.argtypes = [vp, i64, i32, ctypes.POINTER(ctypes.c_float), vp, vp, vp, vp]
    L.rpc_moments_bytes.restype = sz
    L.rpc_normalize_fwd.argtypes = [vp, i64, vp, i32, vp, vp, vp]
    L.rpc_perturb_losses_fwd.argtypes = [vp, i64, vp, i32, vp, vp, vp, vp]
    L.rpc_perturb_losses_bwd.argtypes = [vp, vp, vp, i64, vp, i32, vp, vp]
    L.rpc_attention_scratch_bytes.restype = ctypes.c_size_t
    L.rpc_attention_scratch_bytes.argtypes = []
    L.rpc_attention_gate_fwd.argtypes = [vp, vp, vp, vp, vp, vp, i64, i32, vp, vp, vp]
    L.rpc_attention_gate_bwd.argtypes = [vp] * 7 + [i64, i32] + [vp] * 8
    L.rpc_linear_wgrad.argtypes = [vp, vp, i64, i32, i32, vp, vp, vp]
    L.rpc_linear_dgrad_bnred.argtypes = [vp, i64, i32, i32, vp, vp, vp, vp, vp, vp, vp, vp, vp]
    L.rpc_bn_relu_bwd_sums.argtypes = [vp, vp, vp, vp, vp, vp, i64, i32, vp, vp, vp, vp, vp]
    L.rpc_linear_wgrad_bn.argtypes = [vp, vp, vp, vp, vp, vp, i64, i32, i32, vp, vp, vp]
    L.rpc_bn_scratch_bytes.restype = ctypes.c_size_t
    L.rpc_bn_scratch_bytes.argtypes = []
    L.rpc_bn_relu_fwd.argtypes = [vp, i64, i32, vp, vp, ctypes.c_float, ctypes.c_float, vp, vp, vp, vp, vp, vp, vp]
    L.rpc_bn_relu_bwd.argtypes = [vp, vp, vp, vp, vp, vp, i64, i32, vp, vp, vp, vp, vp]
    L.rpc_linear_fwd.argtypes = [vp, i64, vp, i32, i32, vp, i32, vp, vp, vp, vp, vp, i32, vp, vp, vp]
    L.rpc_bn_finalize.argtypes = [vp, i64, vp, i32, ctypes.c_float, ctypes.c_float, vp, vp, vp, vp, vp]
    L.rpc_bn_relu_apply.argtypes = [vp, i64, vp, i32, vp, vp, vp, vp, vp, vp]
    tdp, vpp = ctypes.POINTER(TrunkDesc), ctypes.POINTER(vp)
    L.rpc_trunk_fwd.argtypes = [tdp, vp, i64, vpp, vp, vp, vp, vp]
    L.rpc_trunk_fwd_eval.argtypes = [tdp, vp, i64, vp, vp, vp, vp]
    L.rpc_trunk_bwd.argtypes = [tdp, vp, i64, vpp, vp, vp, vp, vp, vp, vpp, vpp, vpp, vpp, vp, vp]
    L.rpc_profile_enable.restype = None
    L.rpc_profile_enable.argtypes = [i32]
    L.rpc_profile_read.argtypes = [ctypes.POINTER(ctypes.c_float), ctypes.POINTER(ctypes.c_char_p), i32]
    for name in SYMBOLS:
        getattr(L, name)                      # AttributeError if the .so is stale
    _lib = L
    return L


def strerror(status: int) -> str:
    return lib().rpc_strerror(int(status)).decode()


def check(status: int, where: str) -> None:
    if status != 0:
        raise RpcError(status, where)


def counts(n_per_sample: Sequence[int]):
    arr = (ctypes.c_int32 * len(n_per_sample))(*[int(n) for n in n_per_sample])
    return arr


def profile_enable(on: bool) -> None:
    lib().rpc_profile_enable(1 if on else 0)


def profile_read(max_stages: int = 16):
    """[(kernel name, ms)] of the most recent profiled call (waits for it to finish)"""
    ms = (ctypes.c_float * max_stages)()
    names = (ctypes.c_char_p * max_stages)()
    n = lib().rpc_profile_read(ms, names, max_stages)
    if n < 0:
        raise RpcError(n, "rpc_profile_read")
    return [(names[k].decode(), float(ms[k])) for k in range(n)]
