"""rpc_points_range_filter_static: the allocation-free variant of rpc_points_range_filter (scratch in a
``static __device__`` array of the library; include/rpc_b200.h).  No counterpart in the reference
-- so the check is against the existing entry point: identical outputs (bit-exact, like the rest
of the integer / index work), on two calls in a row (the array keeps the previous call's contents), and the documented
rejection beyond RPC_STATIC_FILTER_MAX_POINTS.  Collected last on purpose (pytest -x)."""
import ctypes

import numpy as np
import pytest
import torch

from robustpointclouds_b200 import _lib, ops, synth

DEV = torch.device("cuda:0")
MAX_POINTS = 4194304                       # RPC_STATIC_FILTER_MAX_POINTS
RPC_ERR_STATIC_LIMIT = -8


def _static_call(points: torch.Tensor, pcr, out, index, count) -> int:
    """straight through the C-ABI: caller-owned inputs and outputs, no workspace argument"""
    rng = (ctypes.c_float * 6)(*[float(v) for v in pcr])
    N, C = points.shape
    return _lib.lib().rpc_points_range_filter_static(
        points.data_ptr(), N, C, rng, out.data_ptr(), index.data_ptr(), count.data_ptr(),
        torch.cuda.current_stream(DEV).cuda_stream)


def test_static_entry_declared_and_rejects_without_gpu():
    """argument checks run before any CUDA call: usable on a CPU-only box"""
    L = _lib.lib()
    rng = (ctypes.c_float * 6)(0, -40, -3, 70.4, 40, 1)
    assert L.rpc_points_range_filter_static(None, 10, 2, rng, None, None, None, None) == -2          # needs x, y, z
    assert L.rpc_points_range_filter_static(None, MAX_POINTS + 1, 4, rng, None, None, None, None) == RPC_ERR_STATIC_LIMIT
    assert "static" in _lib.strerror(RPC_ERR_STATIC_LIMIT)
    assert L.rpc_points_range_filter_static(None, 10, 4, rng, None, None, None, None) == -1          # NULL pointers


@pytest.mark.gpu
def test_static_scratch_filter_matches_workspace_entry_twice():
    cases = [("nuscenes", 300001, 5), ("kitti", 120000, 4), ("kitti", 1025, 4), ("kitti", 1, 4)]
    for rep in range(2):                                            # two rounds: every size follows a different one
        for preset, N, C in cases:
            pcr = synth.PRESETS[preset]["point_cloud_range"]
            pts = synth.make_cloud(preset, N, "lidar", 11 + rep)
            pts[::97, 0] = pcr[0]                                   # on the lower face: dropped (open box)
            if N > 10:
                pts[7, 2] = np.nan
            t = torch.from_numpy(pts).to(DEV)
            ref_out, ref_index, ref_count = ops.points_range_filter(t, pcr)          # existing entry point
            k = int(ref_count)
            for call in range(2):                                   # two calls in a row on the shared array
                out = torch.full_like(t, -7.0)
                index = torch.full((N,), -7, dtype=torch.int32, device=DEV)
                count = torch.full((1,), -7, dtype=torch.int32, device=DEV)
                assert _static_call(t, pcr, out, index, count) == 0
                assert int(count) == k, (preset, N, rep, call)
                assert torch.equal(index[:k], ref_index[:k])
                assert np.array_equal(out[:k].cpu().numpy(), ref_out[:k].cpu().numpy(), equal_nan=True)
                assert bool((index[k:] == -7).all())                # rows beyond the count are not touched


@pytest.mark.gpu
def test_static_scratch_filter_limit_and_empty():
    pcr = synth.PRESETS["kitti"]["point_cloud_range"]
    t = torch.zeros((MAX_POINTS + 1, 3), device=DEV)
    out = torch.empty_like(t)
    index = torch.empty((MAX_POINTS + 1,), dtype=torch.int32, device=DEV)
    count = torch.full((1,), -7, dtype=torch.int32, device=DEV)
    assert _static_call(t, pcr, out, index, count) == RPC_ERR_STATIC_LIMIT
    assert int(count) == -7                                         # nothing was launched
    # exactly at the limit: every tile count of the array in use
    t = t[:MAX_POINTS]
    t[:, 0] = 1.0
    t[::3, 0] = -1.0                                                # outside
    assert _static_call(t, pcr, out, index, count) == 0
    ref = ops.points_range_filter(t, pcr)
    assert int(count) == int(ref[2]) and torch.equal(index[:int(count)], ref[1][:int(count)])
    assert _static_call(t[:0], pcr, out, index, count) == 0 and int(count) == 0
