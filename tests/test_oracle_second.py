"""A second, independently written restatement of mmcv's hard / dynamic voxelization semantics (consumed by the
reference at /root/reference/models/detectors/adversarial_voxelnet.py:66-71,135-137, requirements.txt:9) in vectorised
torch-CPU -- unique + first occurrence + stable sort, no serial walk and no hash table -- and a hypothesis property test
that it, the C oracle (both table modes) and the hand-derived vectors agree on random clouds with duplicate cells, cap
overflow, out-of-range points and NaNs.  This does NOT pin the oracle against mmcv (nothing here is mmcv's output:
"parity unpinned"); it removes the single-author risk of one restatement."""
import os

import numpy as np
import torch
from hypothesis import given, settings, strategies as st

from oracle import ref_path as orc


def torch_grid(voxel_size, pcr):
    vs, r = torch.tensor(voxel_size, dtype=torch.float32), torch.tensor(pcr, dtype=torch.float32)
    return torch.round((r[3:] - r[:3]) / vs).to(torch.int64)                # (gx, gy, gz), fp32 arithmetic like mmcv


def torch_cells(points, voxel_size, pcr):
    """per point (cz, cy, cx) int64, or -1 in all three when the point is dropped (outside the grid on any axis, or NaN)"""
    p = torch.as_tensor(points, dtype=torch.float32)[:, :3]
    vs, r = torch.tensor(voxel_size, dtype=torch.float32), torch.tensor(pcr, dtype=torch.float32)
    g = torch_grid(voxel_size, pcr)
    c = torch.floor((p - r[:3]) / vs)                                        # fp32 subtract, divide, floor
    ok = (~torch.isnan(c)).all(1) & (c >= 0).all(1) & (c < g.to(torch.float32)).all(1)
    c = torch.where(ok[:, None], c, torch.full_like(c, -1.0)).to(torch.int64)
    return c.flip(1), ok                                                     # (z, y, x)


def torch_hard_voxelize(points, voxel_size, pcr, P, V):
    """-> voxels [M, P, C], coors [M, 3] (z, y, x), num_points [M], point2voxel [N], point2slot [N]"""
    pts = torch.as_tensor(points, dtype=torch.float32)
    N, C = pts.shape
    zyx, ok = torch_cells(pts, voxel_size, pcr)
    g = torch_grid(voxel_size, pcr)
    p2v = torch.full((N,), -1, dtype=torch.int64)
    p2s = torch.full((N,), -1, dtype=torch.int64)
    idx = torch.nonzero(ok).flatten()
    if idx.numel() == 0:
        return (torch.zeros((0, P, C)), torch.zeros((0, 3), dtype=torch.int64), torch.zeros((0,), dtype=torch.int64), p2v, p2s)
    lin = (zyx[idx, 0] * g[1] + zyx[idx, 1]) * g[0] + zyx[idx, 2]
    uniq, inv = torch.unique(lin, return_inverse=True)
    first = torch.full((uniq.numel(),), N, dtype=torch.int64).scatter_reduce(0, inv, idx, reduce="amin")
    order = torch.argsort(first, stable=True)                                # cells in order of their first point
    rank = torch.empty_like(order)
    rank[order] = torch.arange(order.numel())
    vox_of = rank[inv]                                                       # voxel index of every in-range point
    # slot = number of earlier in-range points of the same cell: stable sort by cell keeps point order inside a cell
    by_cell = torch.argsort(vox_of, stable=True)
    sorted_v = vox_of[by_cell]
    start = torch.searchsorted(sorted_v, sorted_v, right=False)
    slot = torch.empty_like(by_cell)
    slot[by_cell] = torch.arange(by_cell.numel()) - start
    keep = (vox_of < V) & (slot < P)                                         # cells after the V-th are refused for good
    M = int(min(order.numel(), V))
    voxels = torch.zeros((M, P, C))
    voxels[vox_of[keep], slot[keep]] = pts[idx[keep]]
    coors = zyx[first[order[:M]]]
    num = torch.bincount(vox_of[keep], minlength=M)[:M]
    p2v[idx[keep]] = vox_of[keep]
    p2s[idx[keep]] = slot[keep]
    return voxels, coors, num, p2v, p2s


HERE = os.path.dirname(os.path.abspath(__file__))


def test_second_restatement_on_hand_vectors():
    g = dict(np.load(os.path.join(HERE, "golden", "voxel_micro.npz")))
    v, c, n, p2v, p2s = torch_hard_voxelize(g["pts"], g["vs"], g["pcr"], int(g["P"]), int(g["V"]))
    assert np.array_equal(c.numpy(), g["coors"]) and np.array_equal(n.numpy(), g["num"])
    assert np.array_equal(v.numpy(), g["vox"], equal_nan=True)
    assert np.array_equal(p2v.numpy(), g["p2v"]) and np.array_equal(p2s.numpy(), g["p2s"])
    zyx, _ = torch_cells(g["pts"], g["vs"], g["pcr"])
    assert np.array_equal(zyx.numpy(), g["dyn"])
    assert torch_grid(g["vs"], g["pcr"]).tolist() == g["grid"].tolist()


@st.composite
def clouds(draw):
    vs = draw(st.sampled_from([(0.5, 0.5, 1.0), (0.05, 0.05, 0.1), (0.25, 0.5, 0.75)]))
    pcr = draw(st.sampled_from([(0.0, -4.0, -3.0, 8.0, 4.0, 1.0), (-2.0, -2.0, -1.0, 2.0, 2.0, 1.0)]))
    n = draw(st.integers(0, 300))
    seed = draw(st.integers(0, 2 ** 31 - 1))
    spread = draw(st.sampled_from([0.05, 0.3, 1.0, 1.3]))                    # small: many duplicates; > 1: out of range too
    rng = np.random.default_rng(seed)
    lo, hi = np.array(pcr[:3]), np.array(pcr[3:])
    centre = lo + (hi - lo) * rng.random(3)
    pts = np.empty((n, 4), np.float32)
    pts[:, :3] = centre + (rng.random((n, 3)) - 0.5) * (hi - lo) * spread
    pts[:, 3] = rng.random(n)
    if n:
        k = draw(st.integers(0, min(n, 5)))
        pts[rng.integers(0, n, k), rng.integers(0, 3, k)] = np.nan           # NaN coordinates are dropped
        on = rng.integers(0, n, min(n, 4))
        pts[on, 0] = np.array(pcr)[rng.choice([0, 3], on.size)]             # exactly on a face: lo kept, hi dropped
        snap = rng.integers(0, n, min(n, 6))
        pts[snap, :3] = (lo + np.floor((pts[snap, :3] - lo) / np.array(vs)) * np.array(vs)).astype(np.float32)   # cell edges
    P = draw(st.integers(1, 6))
    V = draw(st.sampled_from([1, 3, 10, 1000]))
    return pts, vs, pcr, P, V


@settings(max_examples=150, deadline=None)
@given(clouds())
def test_three_restatements_agree(case):
    pts, vs, pcr, P, V = case
    v2, c2, n2, p2v2, p2s2 = torch_hard_voxelize(pts, vs, pcr, P, V)
    for table_mode in (0, 1):
        v, c, n, p2v, p2s = orc.hard_voxelize(pts, vs, pcr, P, V, table_mode, with_maps=True)
        assert np.array_equal(c, c2.numpy()), "coors"
        assert np.array_equal(n, n2.numpy()), "num_points"
        assert np.array_equal(v, v2.numpy(), equal_nan=True), "voxels"
        assert np.array_equal(p2v, p2v2.numpy()) and np.array_equal(p2s, p2s2.numpy()), "maps"
    zyx, _ = torch_cells(pts, vs, pcr)
    assert np.array_equal(orc.dynamic_voxelize(pts, vs, pcr), zyx.numpy())
    assert orc.grid_size(vs, pcr).tolist() == torch_grid(vs, pcr).tolist()
